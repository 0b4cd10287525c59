"""Shim for the reference's unmodified agent loop.

`app.py` does `from tpu_ops import map_classify_tpu` and calls it with one positional argument
(app.py:118-123, 283-287). Putting this directory on PYTHONPATH next to the reference's `app.py`
makes that import resolve to the B200 op; nothing in `app.py` changes.
"""
from ops_loader import load_ops

_run = load_ops(["map_classify_tpu"])["map_classify_tpu"]


def map_classify_tpu(payload: dict) -> dict:
    return _run(payload)
