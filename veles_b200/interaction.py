"""Shell unit: drops into an interactive console with the workflow in scope when its gate opens
(reference: ``veles/interaction.py:49-95`` embedded IPython on key press).  Non-interactive runs skip it."""
from __future__ import annotations

import code
import sys

from .distributable import TriviallyDistributable
from .units import Unit


class Shell(Unit, TriviallyDistributable):
    hide_from_registry = False
    BANNER = "veles_b200 interactive shell: `workflow`, `units`, `root` are in scope; Ctrl-D to continue"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(Shell, self).__init__(workflow, **kwargs)
        self.enabled = kwargs.get("enabled", True)

    def initialize(self, **kwargs):
        pass

    def run(self):
        if not self.enabled or not self.interactive or not sys.stdin.isatty():
            return
        from .config import root
        scope = {"workflow": self.workflow, "units": self.workflow.units, "root": root, "self": self}
        try:
            from IPython import embed
            embed(banner1=self.BANNER, user_ns=scope)
        except ImportError:
            code.interact(banner=self.BANNER, local=scope)
