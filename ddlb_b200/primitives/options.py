"""Implementation options: defaults, allowed values, scoped env mutation.

One shared module (the reference carries two identical copies:
``ddlb/primitives/TPColumnwise/utils.py`` and ``.../TPRowwise/utils.py``).
"""
from __future__ import annotations

import os
from typing import Any, Dict, Mapping, Optional


class EnvVarGuard:
    """Set environment variables now, restore the previous values on ``close()``.

    Also restores on garbage collection, like the reference
    (``TPColumnwise/utils.py:9-31``), but ``close()`` is explicit and idempotent.
    """

    def __init__(self, env_vars: Optional[Mapping[str, str]] = None):
        self.env_vars = dict(env_vars or {})
        self._saved: Dict[str, Optional[str]] = {}
        for key, value in self.env_vars.items():
            self._saved[key] = os.environ.get(key)
            os.environ[key] = str(value)
        self._closed = False

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        for key, old in self._saved.items():
            if old is None:
                os.environ.pop(key, None)
            else:
                os.environ[key] = old

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class OptionsManager:
    """Parse and validate per-implementation options.

    ``allowed_values[name]`` is either a list of permitted values or a
    ``(min, max)`` tuple for numeric ranges (reference
    ``TPColumnwise/utils.py:34-108``). Keys in ``BENCHMARK_OPTIONS`` belong to
    the harness and are ignored.
    """

    BENCHMARK_OPTIONS = {"implementation"}

    def __init__(self, default_options: Mapping[str, Any],
                 allowed_values: Optional[Mapping[str, Any]] = None):
        self.default_options = dict(default_options)
        self.allowed_values = dict(allowed_values or {})
        self.options = dict(default_options)

    def parse(self, kwargs: Mapping[str, Any]) -> "OptionsManager":
        given = {k: v for k, v in kwargs.items() if k not in self.BENCHMARK_OPTIONS}
        unknown = sorted(set(given) - set(self.default_options))
        if unknown:
            raise ValueError(
                f"Unknown options provided: {unknown}. "
                f"Valid options are: {sorted(self.default_options)}")
        self.options.update(given)
        for name, allowed in self.allowed_values.items():
            if name not in self.options:
                continue
            value = self.options[name]
            if isinstance(allowed, tuple) and len(allowed) == 2:
                lo, hi = allowed
                ok = (isinstance(value, (int, float)) and not isinstance(value, bool)
                      and lo <= value <= hi)
                if not ok:
                    raise ValueError(
                        f"Invalid value for {name}: {value!r}. "
                        f"Must be a number between {lo} and {hi}")
            elif value not in allowed:
                raise ValueError(
                    f"Invalid value for {name}: {value!r}. Must be one of {list(allowed)}")
        return self

    def get(self, key: str, default: Any = None) -> Any:
        return self.options.get(key, default)

    def __getitem__(self, key: str) -> Any:
        return self.options[key]

    def __contains__(self, key: str) -> bool:
        return key in self.options

    def as_dict(self) -> Dict[str, Any]:
        return dict(self.options)


def setup_ucc_env_vars(backend: str) -> Dict[str, str]:
    """Environment needed for a ``ucc/tl/<tl>`` backend string.

    Kept for config compatibility (reference ``TPColumnwise/utils.py:111-131``);
    UCC itself is not available in this image, so selecting it raises later.
    """
    env: Dict[str, str] = {}
    if backend.startswith("ucc/tl/"):
        tl = backend.rsplit("/", 1)[-1]
        env["UCC_CL_BASIC_TLS"] = tl
        if tl == "ucp":
            env["UCX_RNDV_THRESH"] = "0"
            env["UCX_TLS"] = "ib,cuda_copy"
    return env
