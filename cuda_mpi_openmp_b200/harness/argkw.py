"""Free-form ``--key value`` command-line pairs -> constructor kwargs.

Same rules as ref: arg_parsing.py:1-31 — a flag without a value is True, values
are coerced bool > int > float > str, positional leftovers are ignored.
"""
from __future__ import annotations

from typing import Any, Dict, List


def _coerce(text: str) -> Any:
    low = text.lower()
    if low in ("true", "false"):
        return low == "true"
    for cast in (int, float):
        try:
            return cast(text)
        except ValueError:
            pass
    return text


def unknown_args_to_kwargs(unknown: List[str]) -> Dict[str, Any]:
    kwargs: Dict[str, Any] = {}
    it = iter(range(len(unknown)))
    for i in it:
        tok = unknown[i]
        if not tok.startswith("--"):
            continue
        key = tok.lstrip("-")
        if i + 1 < len(unknown) and not unknown[i + 1].startswith("--"):
            kwargs[key] = _coerce(unknown[i + 1])
            next(it, None)
        else:
            kwargs[key] = True
    return kwargs
