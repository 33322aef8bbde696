"""Platform helpers (reference: ``veles/portable.py``): high-resolution timer alias and ``show_file``."""
from __future__ import annotations

import os
import subprocess
import sys
import time

#: monotonic high resolution clock used by unit timers
clock = time.perf_counter


def show_file(path):
    """Open ``path`` with the desktop's default viewer when there is one; no-op on headless hosts."""
    if sys.platform.startswith("linux"):
        if not os.environ.get("DISPLAY") and not os.environ.get("WAYLAND_DISPLAY"):
            return False
        opener = "xdg-open"
    elif sys.platform == "darwin":
        opener = "open"
    else:
        try:
            os.startfile(path)  # noqa: F821  (Windows only)
            return True
        except AttributeError:
            return False
    try:
        subprocess.Popen([opener, path], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return True
    except OSError:
        return False
