"""Shell unit: a console with the workflow in scope, entered on a key press while the workflow runs
(reference: ``veles/interaction.py:49-95``: every time the unit runs it polls the terminal without blocking; an
``i`` + Enter typed since the last visit stops the training loop at this unit and opens an embedded IPython; leaving
the console resumes the run).  Runs that have no terminal, or that already live inside an interactive interpreter,
skip it.  :meth:`interact` opens the console unconditionally (used by the manhole and by tests with scripted input)."""
from __future__ import annotations

import code
import select
import sys

from .distributable import TriviallyDistributable
from .units import Unit


class Shell(Unit, TriviallyDistributable):
    """Console with the workflow in scope, opened by typing ``i`` + Enter while the workflow runs."""
    hide_from_registry = False
    KEY = "i"
    BANNER = ("veles_b200 interactive console: `workflow`, `units`, `root` are in scope; "
              "Ctrl-D resumes the run")

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(Shell, self).__init__(workflow, **kwargs)
        self.enabled = kwargs.get("enabled", True)
        self.entered = 0          # how many times the console was opened

    def initialize(self, **kwargs):
        pass

    def scope(self, extra_locals=None):
        from .config import root
        ns = {"workflow": self.workflow, "units": self.workflow.units, "root": root, "shell": self}
        ns.update(extra_locals or {})
        return ns

    def interact(self, extra_locals=None, stdin=None):
        """Open the console now.  ``stdin``: an iterable of source lines to run instead of reading the terminal."""
        ns = self.scope(extra_locals)
        self.entered += 1
        if stdin is not None:
            console = code.InteractiveConsole(ns)
            for line in stdin:
                console.push(line.rstrip("\n"))
            return ns
        try:
            from IPython import embed
            embed(banner1=self.BANNER, user_ns=ns)
        except ImportError:
            code.interact(banner=self.BANNER, local=ns)
        return ns

    @staticmethod
    def poll_key(stream=None):
        """First character of a line typed since the last poll ('' when nothing is waiting); never blocks."""
        stream = stream if stream is not None else sys.stdin
        try:
            ready, _, _ = select.select([stream], [], [], 0)
        except (ValueError, OSError, TypeError):      # closed / not selectable (a pipe object on some platforms)
            return ""
        if not ready:
            return ""
        line = stream.readline()
        return line[:1] if line else ""

    def run(self):
        if not self.enabled or self.interactive or not sys.stdin.isatty():
            return
        if self.poll_key() == self.KEY:
            self.interact()
