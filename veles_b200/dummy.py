"""Test doubles (reference: veles/dummy.py:46-131): a launcher that is not a launcher, a
workflow pre-linked start->end, and a trivial unit."""
from .thread_pool import ThreadPool
from .units import TrivialUnit
from .workflow import Workflow


class DummyLauncher(object):
    def __init__(self, mode="standalone", **kwargs):
        self.mode = mode
        self.interactive = kwargs.get("interactive", False)
        self.testing = kwargs.get("testing", False)
        self.id = "dummy"
        self.log_id = "dummy"
        self.device = kwargs.get("device")
        self._thread_pool = None
        self.workflows = []
        self.finished = False

    @property
    def is_master(self):
        return self.mode == "master"

    @property
    def is_slave(self):
        return self.mode == "slave"

    @property
    def is_standalone(self):
        return self.mode == "standalone"

    @property
    def thread_pool(self):
        if self._thread_pool is None:
            self._thread_pool = ThreadPool(2, 2, name="dummy")
        return self._thread_pool

    def add_ref(self, workflow):
        self.workflows.append(workflow)

    def del_ref(self, workflow):
        if workflow in self.workflows:
            self.workflows.remove(workflow)

    def on_workflow_finished(self):
        self.finished = True

    def stop(self):
        if self._thread_pool is not None:
            self._thread_pool.shutdown()
            self._thread_pool = None


class DummyWorkflow(Workflow):
    hide_from_registry = True

    def __init__(self, **kwargs):
        launcher = kwargs.pop("launcher", None) or DummyLauncher(**{k: kwargs.pop(k) for k in ("mode", "device")
                                                                   if k in kwargs})
        super(DummyWorkflow, self).__init__(launcher, **kwargs)
        self.end_point.link_from(self.start_point)


class DummyUnit(TrivialUnit):
    hide_from_registry = True

    def __init__(self, **kwargs):
        wf = kwargs.pop("workflow", None) or DummyWorkflow()
        super(DummyUnit, self).__init__(wf, **{k: v for k, v in kwargs.items() if k in ("name", "view_group")})
        for k, v in kwargs.items():
            if k not in ("name", "view_group"):
                setattr(self, k, v)
