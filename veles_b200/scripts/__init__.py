"""Command line helper scripts (reference: veles/scripts/)."""
