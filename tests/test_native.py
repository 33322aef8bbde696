"""Native (C++) inference runtime: CPU checks of loader/planner/optimizer through the test binary, and a GPU run of
exported packages compared with the Python forward pass (reference: libVeles/tests/*.cc with mnist.zip/.tar.gz)."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build():
    sys.path.insert(0, os.path.join(ROOT, "native"))
    import importlib
    build = importlib.import_module("build")
    return build.build()


def _trained(model, device, **kw):
    from veles_b200 import prng
    from veles_b200.config import root
    from veles_b200.models import build_workflow
    root.common.disable.snapshotting = True
    prng.get().seed(9)
    wf = build_workflow(model, max_epochs=1, **kw)
    wf.initialize(device=device)
    wf.run()
    return wf


def test_native_cpu_checks_and_package_loading(tmp_path):
    from veles_b200.backends import NumpyDevice
    lib, test_bin = _build()
    wf = _trained("mnist_mlp", NumpyDevice(), minibatch_size=50, n_train=200, n_valid=50)
    z, t = str(tmp_path / "mnist.zip"), str(tmp_path / "mnist.tar.gz")
    wf.package_export(z, "zip")
    wf.package_export(t, "tgz", precision=16)
    r = subprocess.run([test_bin, z, t], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=120)
    out = r.stdout.decode()
    assert r.returncode == 0 and "0 failed" in out, out
    assert "All2All" not in out or "2 units" in out
    wf.launcher.stop()
    # a recurrent model: the factory knows the LSTM unit and plans its shapes (784 = 28 steps x 28 pixels -> 10)
    wf = _trained("mnist_lstm", NumpyDevice(), minibatch_size=50, n_train=100, n_valid=50)
    z2 = str(tmp_path / "mnist_lstm.zip")
    wf.package_export(z2, "zip")
    r = subprocess.run([test_bin, z2], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=120)
    out = r.stdout.decode()
    assert r.returncode == 0 and "0 failed" in out and "2 units" in out, out
    wf.launcher.stop()


REF_FIXTURES = "/root/reference/libVeles/tests/workflow_files"


@pytest.mark.skipif(not os.path.isdir(REF_FIXTURES), reason="the reference checkout is not on this machine")
def test_native_loads_the_reference_fixtures():
    """Format compatibility with the reference's own exported packages (libVeles/tests/workflow_files/mnist.zip and
    mnist.tar.gz, read in place): the loader finds the MNIST workflow's two units and their four .npy members, the
    planner sizes the arena, and the C API opens them."""
    lib, test_bin = _build()
    z, t = os.path.join(REF_FIXTURES, "mnist.zip"), os.path.join(REF_FIXTURES, "mnist.tar.gz")
    r = subprocess.run([test_bin, z, t], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=120)
    out = r.stdout.decode()
    assert r.returncode == 0 and "0 failed" in out, out
    assert out.count('workflow "MnistWorkflow", 2 units') == 2, out
    assert "0001_100x784.npy" in out and "0003_10x100.npy" in out
    L = _native(lib)
    for path in (z, t):
        h = L.veles_native_load(path.encode())
        assert h, L.veles_native_last_error()
        h = ctypes.c_void_p(h)
        assert L.veles_native_unit_count(h) == 2
        names = [L.veles_native_unit_name(h, i).decode() for i in range(2)]
        assert "All2AllTanh" in names[0] and "All2AllSoftmax" in names[1], names
        L.veles_native_free(h)


@pytest.mark.skipif(not os.path.isdir(REF_FIXTURES), reason="the reference checkout is not on this machine")
def test_native_logger_through_the_c_api():
    """The runtime's logger (reference: libVeles/src/logger.cc): an embedding application installs a sink, sets the
    thresholds and sees the loader's records; its own records go through the same path."""
    lib, _ = _build()
    L = _native(lib)
    records = []
    SINK = ctypes.CFUNCTYPE(None, ctypes.c_int, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p)
    sink = SINK(lambda level, domain, message, user: records.append((level, domain.decode(), message.decode())))
    L.veles_native_set_log_sink(sink, None)
    try:
        prev = L.veles_native_set_log_level(3)                      # info
        assert 0 <= prev <= 4 and L.veles_native_set_log_level(9) == -1
        h = L.veles_native_load(os.path.join(REF_FIXTURES, "mnist.zip").encode())
        assert h
        L.veles_native_free(ctypes.c_void_p(h))
        assert any(d == "WorkflowLoader" and lv == 3 and "MnistWorkflow" in m and "2 units" in m for lv, d, m in records)
        assert not any(lv == 4 for lv, _, _ in records)             # debug records stay below the threshold
        n = len(records)
        assert L.veles_native_set_domain_log_level(b"WorkflowLoader", 4) == 0
        h = L.veles_native_load(os.path.join(REF_FIXTURES, "mnist.tar.gz").encode())
        L.veles_native_free(ctypes.c_void_p(h))
        assert sum(1 for lv, d, _ in records[n:] if lv == 4 and d == "WorkflowLoader") >= 3   # opening + one per unit
        L.veles_native_log(1, b"App", b"from the host application")
        assert records[-1] == (1, "App", "from the host application")
        L.veles_native_set_log_level(prev)
    finally:
        L.veles_native_set_log_sink(None, None)


def test_header_declares_every_exported_kernel_entry_point():
    """include/veles_b200.h is the C ABI of libveles_b200.so: every exported ``vb_*`` symbol is declared there and
    every declared one is exported."""
    import re
    import shutil
    lib = os.path.join(ROOT, "veles_b200", "lib", "libveles_b200.so")
    if not os.path.exists(lib) or shutil.which("nm") is None:
        pytest.skip("kernel library or nm not available")
    out = subprocess.run(["nm", "-D", "--defined-only", lib], stdout=subprocess.PIPE, check=True).stdout.decode()
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith("vb_")}
    header = open(os.path.join(ROOT, "include", "veles_b200.h")).read()
    declared = set(re.findall(r"\b(vb_[a-z0-9_]+)\s*\(", header))
    assert exported - declared == set(), sorted(exported - declared)
    assert declared - exported == set(), sorted(declared - exported)
    assert len(exported) >= 70


def _native(lib_path):
    L = ctypes.CDLL(lib_path)
    L.veles_native_load.restype = ctypes.c_void_p
    L.veles_native_last_error.restype = ctypes.c_char_p
    L.veles_native_output_size.restype = ctypes.c_longlong
    L.veles_native_unit_name.restype = ctypes.c_char_p
    for fn in (L.veles_native_initialize, L.veles_native_run, L.veles_native_output_size, L.veles_native_free,
               L.veles_native_unit_count, L.veles_native_unit_name, L.veles_native_uses_graph, L.veles_native_plan):
        fn.argtypes = None
    return L


@pytest.mark.gpu
@pytest.mark.parametrize("model,kw,shape", [
    ("mnist_mlp", dict(minibatch_size=50, n_train=200, n_valid=50), (784,)),
    ("cifar_convnet", dict(minibatch_size=32, n_train=64, n_valid=32, layers_kwargs={"caffe_init": False}), (32, 32, 3)),
    ("mnist_lstm", dict(minibatch_size=50, n_train=200, n_valid=50), (784,)),
])
def test_native_runtime_matches_python_forward(tmp_path, model, kw, shape):
    from veles_b200.backends import B200Device
    lib, _ = _build()
    wf = _trained(model, B200Device(), **kw)
    pkg = str(tmp_path / (model + ".zip"))
    wf.package_export(pkg)
    # reference: the framework's own forward pass on one minibatch
    for u in [wf.loader] + list(wf.forwards):
        u.stopped = False
    wf.loader.run()
    for f in wf.forwards:
        f.forward_mode = True
        f.run()
    if hasattr(wf.forwards[-1], "defer_softmax") and wf.forwards[-1].defer_softmax:
        wf.forwards[-1].defer_softmax = False
        wf.forwards[-1].run()
    wf.loader.minibatch_data.map_read()
    wf.forwards[-1].output.map_read()
    x = np.ascontiguousarray(wf.loader.minibatch_data.mem.astype(np.float32))
    ref = wf.forwards[-1].output.mem.reshape(x.shape[0], -1).astype(np.float32)
    L = _native(lib)
    h = ctypes.c_void_p(L.veles_native_load(pkg.encode()))
    assert h.value, L.veles_native_last_error()
    dims = (ctypes.c_longlong * len(shape))(*shape)
    arena, naive = ctypes.c_longlong(), ctypes.c_longlong()
    assert L.veles_native_plan(h, x.shape[0], dims, len(shape), ctypes.byref(arena), ctypes.byref(naive)) == 0
    assert 0 < arena.value <= naive.value
    assert L.veles_native_initialize(h, x.shape[0], dims, len(shape), 0, 1) == 0, L.veles_native_last_error()
    n_out = L.veles_native_output_size(h)
    assert n_out == ref.shape[1]
    out = np.zeros((x.shape[0], n_out), dtype=np.float32)
    for _ in range(3):   # the third run replays the captured CUDA graph
        rc = L.veles_native_run(h, x.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p), x.shape[0])
        assert rc == 0, L.veles_native_last_error()
    assert L.veles_native_uses_graph(h) == 1
    assert np.abs(out.sum(axis=1) - 1.0).max() < 2e-2
    assert np.abs(out - ref).max() < 3e-2, np.abs(out - ref).max()
    assert (out.argmax(1) == ref.argmax(1)).mean() > 0.95
    L.veles_native_free(h)
    wf.launcher.stop()
