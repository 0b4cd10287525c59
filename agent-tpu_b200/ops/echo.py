"""Plumbing diagnostic: returns its payload (reference ops/echo.py:8-24, app.py:114-115)."""
from typing import Any, Dict

from . import register_op


@register_op("echo")
def echo_op(payload: Any) -> Dict[str, Any]:
    return {"ok": True, "echo": payload}
