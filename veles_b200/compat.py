"""Python-version shims kept for API compatibility with user code written against the reference (veles/compat.py)."""
from_none = lambda exc: exc  # noqa: E731 - `raise from_none(e)` idiom


def is_interactive():
    import sys
    return hasattr(sys, "ps1") or "IPython" in sys.modules
