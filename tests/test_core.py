"""Graph-runtime behaviour specs (reference tests: veles/tests/test_units.py, test_workflow.py,
test_mutable.py, test_config.py, test_pickle.py)."""
import logging
import pickle
import threading

import numpy as np
import pytest

from veles_b200.config import Config, root, get
from veles_b200.distributable import Pickleable
from veles_b200.dummy import DummyLauncher, DummyUnit, DummyWorkflow
from veles_b200.mutable import Bool, LinkableAttribute
from veles_b200.plumbing import Repeater
from veles_b200.units import TrivialUnit, Unit, NotInitializedError, RunAfterStopError
from veles_b200.workflow import Workflow


# ---------------------------------------------------------------- Bool
def test_bool_algebra_is_live():
    a, b = Bool(False), Bool(True)
    c = a | b
    d = a & b
    e = ~a
    f = a ^ b
    assert c and not d and e and f
    a <<= True
    assert c and d and not e and not f
    b <<= False
    assert c and not d


def test_bool_callable_and_guard():
    state = {"v": False}
    a = Bool(lambda: state["v"])
    assert not a
    state["v"] = True
    assert a
    with pytest.raises(RuntimeError):
        (a | Bool(True)).__ilshift__(True)
    with pytest.raises(TypeError):
        Bool().__ilshift__("yes")


def test_bool_callbacks_and_pickle():
    fired = []
    a = Bool(False)
    c = ~a
    c.on_false = lambda b: fired.append("false")
    a <<= True
    assert fired == ["false"]
    a2, c2 = pickle.loads(pickle.dumps((a, c)))
    assert a2 and not c2
    a2 <<= False
    assert c2  # the expression stays live after unpickling


# ---------------------------------------------------------------- Config
def test_config_autovivify_update_protect():
    c = Config("test")
    c.a.b.c = 5
    assert c.a.b.c == 5
    c.update({"x": {"y": 1, "z": {"dict": True, "k": 2}}, "a": {"b": {"d": 7}}})
    assert c.x.y == 1 and c.x.z == {"k": 2} and c.a.b.c == 5 and c.a.b.d == 7
    assert get(c.nothing.here, "dflt") == "dflt" and get(c.x.y, 0) == 1
    c.protect("x")
    with pytest.raises(AttributeError):
        c.x = 3
    with pytest.raises(ValueError):
        root.update({"a": 1})
    c2 = pickle.loads(pickle.dumps(c))
    assert c2.a.b.d == 7 and c2.to_dict()["x"]["y"] == 1


# ---------------------------------------------------------------- Pickleable
class Holder(Pickleable):
    def __init__(self):
        self.count = 0
        super(Holder, self).__init__()
        self.kept = 1
        self.volatile_ = 2

    def init_unpickled(self):
        super(Holder, self).init_unpickled()
        self.count_ = getattr(self, "count_", 0) + 1
        self.lock_ = threading.Lock()


def test_trailing_underscore_attributes_are_not_pickled():
    h = Holder()
    assert h.count_ == 1
    h2 = pickle.loads(pickle.dumps(h))
    assert h2.kept == 1 and not hasattr(h2, "volatile_")
    assert h2.count_ == 1 and h2.lock_ is not h.lock_


# ---------------------------------------------------------------- Units / links
def test_link_attrs_one_and_two_way():
    wf = DummyWorkflow()
    a, b = TrivialUnit(wf, name="a"), TrivialUnit(wf, name="b")
    a.value = 1
    a.arr = [1, 2]
    b.link_attrs(a, "value", ("mine", "arr"))
    assert b.value == 1 and b.mine is a.arr
    a.value = 5
    assert b.value == 5
    with pytest.raises(RuntimeError):
        b.value = 9
    b.link_attrs(a, "value", two_way=True)
    b.value = 9
    assert a.value == 9
    c = TrivialUnit(wf, name="c")
    c.value = "own"          # other instances of the class keep plain attributes
    assert c.value == "own"
    with pytest.raises(TypeError):
        b.link_attrs(a, 3)


def test_demand_blocks_initialize():
    wf = DummyWorkflow()
    u = TrivialUnit(wf)
    u.demand("input")
    with pytest.raises(AttributeError):
        u.initialize()
    u.input = 3
    u.initialize()
    assert u.is_initialized


def test_gates_and_run_order():
    wf = DummyWorkflow()
    wf.end_point.unlink_all()
    log = []

    class Rec(TrivialUnit):
        def run(self):
            log.append(self.name)
    a, b, c = Rec(wf, name="a"), Rec(wf, name="b"), Rec(wf, name="c")
    a.link_from(wf.start_point)
    b.link_from(wf.start_point)
    c.link_from(a, b)          # c needs BOTH
    wf.end_point.link_from(c)
    wf.initialize()
    wf.run()
    assert sorted(log[:2]) == ["a", "b"] and log[2] == "c" and len(log) == 3
    log.clear()
    b.gate_skip = Bool(True)   # skipped units still propagate
    wf.run()
    assert "b" not in log and log[-1] == "c"
    log.clear()
    a.gate_block = Bool(True)  # blocked units stop propagation: c never gets both inputs
    t = threading.Thread(target=wf.run, daemon=True)
    t.start()
    t.join(1.0)
    assert "c" not in log
    wf.on_workflow_finished()
    t.join(2.0)
    wf.launcher.stop()


def test_run_guards():
    wf = DummyWorkflow()
    u = TrivialUnit(wf)
    u.link_from(wf.start_point)
    with pytest.raises(NotInitializedError):
        u._check_gate_and_run(wf.start_point)
    u.initialize()
    u.stop()
    with pytest.raises(RunAfterStopError):
        u.run()
    assert u.stopped


def test_repeater_loop_constant_stack_and_stats():
    wf = DummyWorkflow()
    wf.end_point.unlink_all()
    rep = Repeater(wf)
    rep.link_from(wf.start_point)

    class Counter(TrivialUnit):
        def __init__(self, w, **kw):
            super(Counter, self).__init__(w, **kw)
            self.n = 0
            self.done = Bool(False)

        def run(self):
            self.n += 1
            if self.n == 3000:
                self.done <<= True
    c = Counter(wf, name="counter")
    c.link_from(rep)
    rep.link_from(c)
    rep.gate_block = c.done
    wf.end_point.link_from(c)
    wf.end_point.gate_block = ~c.done
    wf.initialize()
    wf.run()
    assert c.n == 3000 and wf.launcher.finished
    stats = wf.print_stats(file=open("/dev/null", "w"))
    assert stats[0][0].name in ("counter", "Repeater")
    wf.launcher.stop()


def test_unknown_kwargs_warn_with_hint(caplog):
    class Knob(TrivialUnit):
        def __init__(self, workflow, **kwargs):
            super(Knob, self).__init__(workflow, **kwargs)
            self.learning_rate = kwargs.get("learning_rate", 0.1)
    with caplog.at_level(logging.WARNING):
        Knob(DummyWorkflow(), learnign_rate=0.5)
    assert any("misprint" in r.message and "learning_rate" in r.message for r in caplog.records)


# ---------------------------------------------------------------- Workflow
def test_workflow_container_graph_pickle():
    wf = DummyWorkflow()
    u = TrivialUnit(wf, name="mid")
    u.insert_after(wf.start_point)
    assert wf["mid"] is u and len(wf) == 3 and u in wf
    assert [x.name for x in wf.units_in_dependency_order] == ["Start", "mid", "End"]
    dot = wf.generate_graph(write_on_disk=False)
    assert dot.startswith("digraph Workflow {") and dot.count("->") == 2 and "mid\\nTrivialUnit" in dot
    wf2 = pickle.loads(pickle.dumps(wf))
    assert wf2.restored_from_snapshot and wf2["mid"].restored_from_snapshot
    assert [x.name for x in wf2.units_in_dependency_order] == ["Start", "mid", "End"]
    assert wf.checksum == wf2.checksum and wf.checksum.endswith("_3")
    other = TrivialUnit(wf, name="other")
    wf.change_unit("mid", other)
    assert [x.name for x in wf.units_in_dependency_order] == ["Start", "other", "End"]


def test_nested_workflow_runs_as_unit():
    outer = DummyWorkflow()
    outer.end_point.unlink_all()
    inner = Workflow(outer, name="inner")
    ran = []

    class R(TrivialUnit):
        def run(self):
            ran.append(1)
    r = R(inner)
    r.link_from(inner.start_point)
    inner.end_point.link_from(r)
    inner.link_from(outer.start_point)
    outer.end_point.link_from(inner)
    outer.initialize()
    outer.run()
    assert ran == [1] and outer.launcher.finished
    outer.launcher.stop()


def test_results_and_package_export(tmp_path):
    import json
    import zipfile
    from veles_b200.backends import NumpyDevice
    from veles_b200.models import build_workflow
    root.common.disable.snapshotting = True
    wf = build_workflow("mnist_mlp", minibatch_size=50, n_train=200, n_valid=100, max_epochs=1)
    wf.initialize(device=NumpyDevice())
    wf.run()
    res = wf.write_results(str(tmp_path / "r.json"))
    assert "EvaluationFitness" in res and "Total epochs" in res
    pkg = str(tmp_path / "mnist.zip")
    wf.package_export(pkg)
    with zipfile.ZipFile(pkg) as z:
        names = z.namelist()
        contents = json.loads(z.read("contents.json"))
    assert contents["units"][0]["class"]["uuid"] == "b3a2bd5c-3c01-46ef-978a-fef22e008f31"  # All2AllTanh
    assert contents["units"][1]["class"]["name"] == "All2AllSoftmax"
    assert contents["units"][0]["data"]["weights"] == "@0001_100x784" or contents["units"][0]["data"]["weights"].startswith("@")
    assert contents["units"][0]["links"] == [1] and contents["units"][1]["links"] == []
    assert any(n.endswith("100x784.npy") for n in names)
    wf.package_export(str(tmp_path / "mnist.tgz"), "tgz", precision=16)
    wf.launcher.stop()


def test_reference_import_paths_alias_to_this_package():
    """``import veles.x`` resolves to ``veles_b200.x`` (same module objects) so workflows and snapshots written
    against the reference's module names load."""
    import pickle
    from veles_b200.compat import install_veles_alias
    if not install_veles_alias():
        import pytest
        pytest.skip("a real 'veles' package is installed")
    import veles
    import veles_b200
    assert veles.__name__ == "veles_b200"
    from veles.znicz.all2all import All2AllTanh
    from veles_b200.znicz import all2all
    assert All2AllTanh is all2all.All2AllTanh
    import veles.loader.fullbatch as fb
    assert fb.__name__ == "veles_b200.loader.fullbatch"
    cls = pickle.loads(b"cveles.mutable\nBool\n.")   # a pickle GLOBAL opcode naming the reference path
    from veles_b200.mutable import Bool
    assert cls is Bool
