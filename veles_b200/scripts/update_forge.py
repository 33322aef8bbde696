"""Upload every sample model directory (containing manifest.json) under a root to a forge server
(reference: veles/scripts/update_forge.py)."""
from __future__ import annotations

import os
import sys


def main(argv=None):
    from ..forge.forge_client import ForgeClient
    argv = argv if argv is not None else sys.argv[1:]
    if len(argv) < 2:
        print("usage: update_forge <models root> <server url> [token]")
        return 1
    client = ForgeClient(argv[1], argv[2] if len(argv) > 2 else None)
    for base, _, names in os.walk(argv[0]):
        if "manifest.json" in names:
            print(base, client.upload(base))
    return 0


if __name__ == "__main__":
    sys.exit(main())
