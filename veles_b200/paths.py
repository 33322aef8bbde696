"""Well-known paths (reference: veles/paths.py)."""
import os

__root__ = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
__home__ = os.path.join(os.environ.get("HOME", "/tmp"), ".veles_b200")
