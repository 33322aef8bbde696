"""MeanDispNormalizer / InputJoiner / Avatar / normalizers / PRNG / loader behaviour (CPU)."""
import pickle

import numpy as np
import pytest

from veles_b200 import prng
from veles_b200.avatar import Avatar
from veles_b200.backends import NumpyDevice
from veles_b200.dummy import DummyWorkflow
from veles_b200.input_joiner import InputJoiner
from veles_b200.mean_disp_normalizer import MeanDispNormalizer
from veles_b200.memory import Array
from veles_b200.normalization import NormalizerRegistry, factory


def test_mean_disp_normalizer_numpy():
    rng = np.random.RandomState(0)
    inp = rng.randint(0, 256, size=(15, 3, 7, 5)).astype(np.uint8)
    mean = rng.rand(3, 7, 5).astype(np.float32) * 255
    rdisp = (rng.rand(3, 7, 5) + 0.1).astype(np.float32)
    u = MeanDispNormalizer(DummyWorkflow())
    u.input, u.mean, u.rdisp = Array(inp), Array(mean), Array(rdisp)
    u.initialize(device=NumpyDevice())
    u.run()
    assert np.abs(u.output.mem - (inp.astype(np.float32) - mean) * rdisp).max() < 1e-5


def test_input_joiner_numpy():
    a, b, c = np.random.rand(9, 4), np.random.rand(9, 2, 3), np.random.rand(9, 1)
    u = InputJoiner(DummyWorkflow(), inputs=[Array(a), Array(b), Array(c)])
    u.initialize(device=NumpyDevice())
    u.run()
    assert u.output.shape == (9, 11)
    assert np.allclose(u.output.mem, np.concatenate([a, b.reshape(9, 6), c], axis=1))
    assert u.offsets == [0, 4, 10]


def test_avatar_clones_arrays_and_scalars():
    from veles_b200.units import TrivialUnit
    wf = DummyWorkflow()
    real = TrivialUnit(wf)
    real.data = Array(np.arange(6, dtype=np.float32).reshape(2, 3))
    real.size = 2
    av = Avatar(wf)
    av.reals[real] = ["data", "size"]
    av.clone()
    av.initialize(device=NumpyDevice())
    real.data.mem[...] = 7
    real.size = 1
    av.run()
    assert (av.data.mem == 7).all() and av.size == 1 and av.data is not real.data


@pytest.mark.parametrize("name", sorted(n for n in NormalizerRegistry.normalizers))
def test_normalizers_roundtrip(name):
    rng = np.random.RandomState(1)
    data = rng.uniform(-3, 5, size=(40, 6, 2)).astype(np.float32) + 6.0  # positive: exp's inverse needs it
    orig = data.copy()
    kw = {"mean_source": orig.mean(axis=0)} if name == "external_mean" else {}
    norm = factory(name, **kw)
    norm.analyze(data)
    extra = norm.normalize(data) or {}
    if name == "linear":
        assert np.allclose(data.reshape(40, -1).min(axis=1), -1, atol=1e-5)
        assert np.allclose(data.reshape(40, -1).max(axis=1), 1, atol=1e-5)
    if name == "mean_disp":
        assert abs(data.mean()) < 1e-5
    if name == "exp":
        assert np.allclose(data.reshape(40, -1).sum(axis=1), 1, atol=1e-5)
    n2 = pickle.loads(pickle.dumps(norm))
    back = n2.denormalize(data.copy(), **extra)
    assert np.allclose(back, orig, atol=2e-3, rtol=2e-3)


def test_prng_streams_are_reproducible_and_independent():
    a, b = prng.get(101), prng.get(102)
    a.seed(5)
    b.seed(5)
    x = a.rand(4)
    b.rand(10)              # another user in between must not disturb generator a
    a.seed(5)
    assert np.array_equal(a.rand(4), x)
    arr = np.zeros((3, 5), dtype=np.float32)
    a.fill(arr, -2, 2)
    assert arr.min() >= -2 and arr.max() < 2 and arr.std() > 0
    a.fill_normal_real(arr, 0, 1.0)
    assert np.abs(arr).max() <= 5.0
    st = np.random.RandomState(3).randint(1, 2 ** 62, size=(2, 16)).astype(np.uint64)
    st2 = st.copy()
    out = prng.random_generator.xorshift1024star_fill(st2, 2)
    assert out.shape == (64,) and len(set(out.tolist())) == 64 and not np.array_equal(st, st2)


def test_loader_epoch_flags_and_pickle():
    from veles_b200.loader.base import TRAIN, VALID
    from veles_b200.loader.synthetic import SyntheticFullBatchLoader
    l = SyntheticFullBatchLoader(DummyWorkflow(), minibatch_size=64, n_train=200, n_valid=100, sample_shape=(8,))
    l.initialize(device=NumpyDevice())
    seq = []
    for _ in range(12):
        l.run()
        seq.append((l.minibatch_class, l.minibatch_size, bool(l.last_minibatch), bool(l.epoch_ended), l.epoch_number))
    assert seq[:6] == [(VALID, 64, False, False, 0), (VALID, 36, True, True, 0), (TRAIN, 64, False, False, 0),
                       (TRAIN, 64, False, False, 0), (TRAIN, 64, False, False, 0), (TRAIN, 8, True, False, 0)]
    assert seq[6][4] == 1 and seq[7][3]
    assert (l.minibatch_labels.mem[8:] == -1).all() and (l.minibatch_data.mem[8:] == 0).all()
    first_train = l.shuffled_indices.mem[100:].copy()
    l2 = pickle.loads(pickle.dumps(l))
    assert l2.global_offset == l.global_offset and np.array_equal(l2.shuffled_indices.mem[100:], first_train)
    assert sorted(first_train.tolist()) == list(range(100, 300))  # only the train part is shuffled


def test_loader_validation_ratio_is_stratified():
    from veles_b200.loader.synthetic import SyntheticFullBatchLoader

    class L(SyntheticFullBatchLoader):
        def load_data(self):
            super(L, self).load_data()
            self.resize_validation(0.25)
    l = L(DummyWorkflow(), minibatch_size=50, n_train=400, n_valid=0, sample_shape=(4,), n_classes=4)
    l.initialize(device=NumpyDevice())
    nv = l.class_lengths[1]
    assert abs(nv - 100) <= 2 and l.class_lengths[2] == 400 - nv   # per-class rounding
    valid_labels = np.array(l.original_labels[:nv])
    all_labels = np.array(l.original_labels)
    for c in range(4):
        assert abs((valid_labels == c).sum() - 0.25 * (all_labels == c).sum()) <= 1


def test_device_benchmark_and_workflow_computing_power():
    """DeviceBenchmark / AcceleratedWorkflow.computing_power on the numpy backend (reference:
    accelerated_units.py:706-866: power = 1000 / time of the GEMM batch, cached between measurements)."""
    from veles_b200.accelerated_units import DeviceBenchmark, AcceleratedWorkflow
    from veles_b200.backends import NumpyDevice
    from veles_b200.dummy import DummyLauncher
    launcher = DummyLauncher()
    wf = AcceleratedWorkflow(launcher)
    wf.initialize(device=NumpyDevice())
    bench = DeviceBenchmark(wf, size=64, repeats=3)
    bench.initialize(device=NumpyDevice())
    bench.run()
    assert bench.computing_power > 0
    p1 = wf.computing_power
    assert p1 > 0 and wf.computing_power == p1          # cached until the measurement interval passes
    launcher.stop()


def test_shell_unit_is_inert_without_a_terminal():
    """Shell drops into a console only for interactive runs on a tty (reference: interaction.py:49-95)."""
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.interaction import Shell
    wf = DummyWorkflow()
    sh = Shell(wf)
    sh.initialize()
    assert sh.run() is None                              # pytest has no tty: returns immediately
    sh.enabled = False
    assert sh.run() is None
    assert "workflow" in Shell.BANNER


def test_shell_key_poll_and_scripted_console():
    """An ``i`` typed while the workflow runs opens the console (reference: interaction.py:84-95); the console sees the
    workflow and can change it."""
    import os
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.interaction import Shell
    wf = DummyWorkflow()
    sh = Shell(wf)
    r, w = os.pipe()
    with os.fdopen(r, "r") as rf:
        assert Shell.poll_key(rf) == ""                  # nothing typed: returns at once
        os.write(w, b"i\n")
        assert Shell.poll_key(rf) == "i"
        os.write(w, b"x\n")
        assert Shell.poll_key(rf) == "x"
        os.close(w)
    ns = sh.interact({"extra": 5}, stdin=["workflow.visited_by_shell = extra + len(units)\n", "y = root.common.engine\n"])
    assert wf.visited_by_shell == 5 + len(wf.units) and sh.entered == 1 and "y" in ns
