"""`load_ops(tasks)`: resolve every requested op at start-up (reference ops_loader.py:8-19).

Unknown or disabled names raise ValueError here, before the agent starts leasing work.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List

from ops import get_op, list_ops  # noqa: F401  (list_ops re-exported like the reference)


def load_ops(tasks: List[str]) -> Dict[str, Callable[..., Any]]:
    return {name: get_op(name) for name in tasks}
