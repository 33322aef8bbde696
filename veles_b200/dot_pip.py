"""Per-user package directory for workflow plugins (reference: ``veles/dot_pip.py``: ``~/.veles/pip`` is
appended to ``sys.path`` so forge-fetched models can ship extra pure-Python dependencies)."""
from __future__ import annotations

import os
import sys


def dot_pip_dir():
    return os.path.join(os.path.expanduser("~"), ".veles_b200", "pip")


def install_dot_pip(create=False):
    """Put the per-user plugin directory on ``sys.path`` (last, so it never shadows the environment)."""
    path = dot_pip_dir()
    if create:
        os.makedirs(path, exist_ok=True)
    if os.path.isdir(path) and path not in sys.path:
        sys.path.append(path)
    return path
