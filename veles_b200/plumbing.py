"""Control-flow plumbing units (reference: veles/plumbing.py:17-112)."""
import os

from .distributable import TriviallyDistributable
from .units import Unit, TrivialUnit


class Repeater(TrivialUnit):
    """Loop head: fires whenever ANY incoming link fires (``ignore_gate=True``)."""
    hide_from_registry = False
    VIEW_GROUP = "PLUMBING"

    def __init__(self, workflow, **kwargs):
        kwargs["ignore_gate"] = True
        kwargs.setdefault("view_group", "PLUMBING")
        super(Repeater, self).__init__(workflow, **kwargs)

    def link_from(self, *args):
        super(Repeater, self).link_from(*args)
        return self

    def run(self):
        # one beat per loop iteration for the rank supervisor (no-op unless VELES_HEARTBEAT_FILE is set)
        if os.environ.get("VELES_HEARTBEAT_FILE"):
            from .parallel.supervisor import heartbeat, maybe_die
            heartbeat()
            maybe_die()
        return super(Repeater, self).run()


class UttermostPoint(TrivialUnit):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "START_END")
        super(UttermostPoint, self).__init__(workflow, **kwargs)


class StartPoint(UttermostPoint):
    """Workflow entry: ``Workflow.run`` calls ``start_point.run_dependent()``."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("name", "Start")
        super(StartPoint, self).__init__(workflow, **kwargs)


class EndPoint(UttermostPoint):
    """Workflow exit: running it finishes the workflow."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("name", "End")
        super(EndPoint, self).__init__(workflow, **kwargs)

    def init_unpickled(self):
        super(EndPoint, self).init_unpickled()
        # EndPoint must be runnable even after the workflow asked units to stop

    def run(self):
        self.workflow.on_workflow_finished()

    def generate_data_for_master(self):
        return True

    def apply_data_from_slave(self, data, slave):
        # master side: the loop ends when the gate into EndPoint is open
        if not self.gate_block:
            self.workflow.on_workflow_finished()


class FireStarter(Unit, TriviallyDistributable):
    """Un-stops the given units each time it runs (lets a stopped sub-graph run again)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "PLUMBING")
        super(FireStarter, self).__init__(workflow, **kwargs)
        self._units = set(kwargs.get("units", ()))

    @property
    def units(self):
        return self._units

    def initialize(self, **kwargs):
        pass

    def run(self):
        for unit in self._units:
            unit.stopped = False
