"""Op registry with the reference's surface: `register_op`, `get_op`, `list_ops`, `OP_TO_MODULE`,
`OPS_REGISTRY`, `OPS_LOAD_ERRORS` (reference ops/__init__.py:8-111).

Behaviour kept from the reference:
  * an op name maps to a module under `ops/`, imported lazily on first `get_op`
    (ops/__init__.py:74-84); the module registers its handler with `@register_op(name)`
    (ops/__init__.py:35-39);
  * the `TASKS` environment variable is re-read on every call: unset, "*" or "all" enable every
    op, "none" disables all, otherwise a comma-separated allow-list (ops/__init__.py:42-57);
  * import failures are caught, printed with the `[ops]` prefix, remembered in OPS_LOAD_ERRORS and
    surface as ValueError from `get_op` (ops/__init__.py:81-84, 97-106).

Only the ops of the hot path exist here: the classifier, the tokenizer, the CSV
shard reader that feeds them, and the plumbing `echo`. The reference's phantom entries (fibonacci, prime_factor, ... — mapped but without
modules) are not reproduced.
"""
from __future__ import annotations

import importlib
import os
from typing import Any, Callable, Dict, List, Optional, Set, Tuple

Handler = Callable[..., Any]

OPS_REGISTRY: Dict[str, Handler] = {}
OPS_LOAD_ERRORS: List[Tuple[str, str]] = []

OP_TO_MODULE: Dict[str, str] = {
    "echo": "echo",
    "map_tokenize": "map_tokenize",
    "map_classify_tpu": "map_classify_tpu",
    # the reference lists the module as "csv_shard" but registers the handler as "read_csv_shard",
    # so neither name resolves there; both resolve here
    "csv_shard": "csv_shard",
    "read_csv_shard": "csv_shard",
}

_LOADED: Set[str] = set()


def register_op(name: str) -> Callable[[Handler], Handler]:
    """Decorator: make `fn` the handler of op `name`; returns `fn` unchanged."""

    def bind(fn: Handler) -> Handler:
        OPS_REGISTRY[name] = fn
        return fn

    return bind


def _enabled_set() -> Optional[Set[str]]:
    """None = everything enabled; otherwise the allow-list from $TASKS."""
    names = [piece.strip() for piece in os.getenv("TASKS", "").split(",")]
    names = [n for n in names if n]
    if not names:
        return None
    folded = {n.lower() for n in names}
    if folded & {"*", "all"}:
        return None
    if "none" in folded:
        return set()
    return set(names)


def list_ops() -> List[str]:
    allowed = _enabled_set()
    return sorted(n for n in OP_TO_MODULE if allowed is None or n in allowed)


def _load(module: str) -> None:
    if module in _LOADED:
        return
    try:
        importlib.import_module(f"{__name__}.{module}")
    except Exception as exc:  # the agent must keep starting: record and report
        text = f"{type(exc).__name__}: {exc}"
        OPS_LOAD_ERRORS.append((module, text))
        print(f"[ops] ERROR: failed to import ops.{module}: {text}", flush=True)
    else:
        _LOADED.add(module)


def get_op(name: str) -> Handler:
    allowed = _enabled_set()
    if allowed is not None and name not in allowed:
        raise ValueError(f"Op {name!r} is not enabled by TASKS. Enabled ops: {list_ops()}")
    module = OP_TO_MODULE.get(name)
    if not module:
        raise ValueError(f"Unknown op {name!r}. Allowed ops: {sorted(OP_TO_MODULE.keys())}")
    _load(module)
    handler = OPS_REGISTRY.get(name)
    if handler is not None:
        return handler
    registered = sorted(OPS_REGISTRY.keys())
    if OPS_LOAD_ERRORS:
        shown = "; ".join(f"{m} => {e}" for m, e in OPS_LOAD_ERRORS[:10])
        extra = f" (+{len(OPS_LOAD_ERRORS) - 10} more)" if len(OPS_LOAD_ERRORS) > 10 else ""
        raise ValueError(
            f"Unknown or failed op {name!r}. Registered ops: {registered}. "
            f"Also saw op import errors: {shown}{extra}"
        )
    raise ValueError(f"Unknown op {name!r}. Registered ops: {registered}")


__all__ = ["OPS_REGISTRY", "OPS_LOAD_ERRORS", "OP_TO_MODULE", "register_op", "list_ops", "get_op"]
