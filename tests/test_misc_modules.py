"""Small modules of the inventory: Uniform PRNG unit, audio / HDFS-text / MSE-image loaders, bboxer, txzmq, BLAS."""
import json
import os
import threading
import urllib.request
import wave

import numpy as np
import pytest


def test_uniform_unit_matches_oracle_and_is_reproducible():
    from veles_b200.backends import NumpyDevice
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.prng.uniform import Uniform
    from veles_b200.prng import random_generator as rg
    u = Uniform(DummyWorkflow(), num_states=8, output_bytes=8 * 16 * 8 * 3)
    u.initialize(device=NumpyDevice())
    s0 = u.states.mem.copy()
    u.run()
    ref = rg.xorshift1024star_fill(s0.copy(), 3)
    assert np.array_equal(u.output.mem, ref)
    assert len(np.unique(u.output.mem)) == u.output.mem.size
    first = u.output.mem.copy()
    u.run()
    assert not np.array_equal(first, u.output.mem)


@pytest.mark.gpu
def test_uniform_unit_device_bitstream_equals_numpy_oracle():
    from veles_b200.backends import B200Device
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.prng.uniform import Uniform
    from veles_b200.prng import random_generator as rg
    u = Uniform(DummyWorkflow(), num_states=256, output_bytes=256 * 128 * 2)
    u.initialize(device=B200Device())
    s0 = u.states.mem.copy()
    u.run()
    u.output.map_read()
    assert np.array_equal(u.output.mem, rg.xorshift1024star_fill(s0, 2))


def test_sndfile_reads_wav_and_loader_base_decodes(tmp_path):
    from veles_b200.loader.libsndfile_loader import SndFile, SndFileLoaderBase
    p = str(tmp_path / "a.wav")
    t = np.arange(800) / 8000.0
    pcm = (np.sin(2 * np.pi * 440 * t) * 20000).astype("<i2")
    with wave.open(p, "wb") as w:
        w.setnchannels(1)
        w.setsampwidth(2)
        w.setframerate(8000)
        w.writeframes(pcm.tobytes())
    data, rate = SndFile.read(p)
    assert rate == 8000 and data.shape == (800, 1)
    assert np.abs(data[:, 0] - pcm / 32768.0).max() < 1e-3
    d = SndFileLoaderBase.decode_file(p)
    assert d["channels"] == 1 and d["sampling_rate"] == 8000
    assert not SndFileLoaderBase.file_format_is_supported(__file__)


def test_hdfs_text_loader_streams_chunks_from_any_filesystem(tmp_path):
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.loader.hdfs_loader import HDFSTextLoader
    p = tmp_path / "t.txt"
    p.write_text("\n".join("line %d" % i for i in range(7)) + "\n")
    ld = HDFSTextLoader(DummyWorkflow(), file=str(p), chunk=3)
    ld.initialize()
    seen = []
    while not ld.finished:
        ld.run()
        seen += [l for l in ld.output if l]
    assert seen == ["line %d" % i for i in range(7)]
    ld.stop()


def test_bboxer_serves_and_saves_boxes(tmp_path):
    from veles_b200.scripts import bboxer
    (tmp_path / "sub").mkdir()
    (tmp_path / "sub" / "a.png").write_bytes(b"\x89PNG\r\n\x1a\n" + b"0" * 32)
    srv = bboxer.serve(str(tmp_path), port=0)
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    base = "http://127.0.0.1:%d/" % srv.server_address[1]
    try:
        assert json.loads(urllib.request.urlopen(base + "images").read()) == ["sub/a.png"]
        assert b"<canvas" in urllib.request.urlopen(base).read()
        req = urllib.request.Request(base + "boxes/0", data=json.dumps(
            [{"x": 1, "y": 2, "width": 30, "height": 40, "label": "cat"}]).encode(), method="POST")
        assert json.loads(urllib.request.urlopen(req).read())["ok"]
        saved = json.loads((tmp_path / "sub" / "a.json").read_text())
        assert saved[0]["label"] == "cat" and saved[0]["width"] == 30.0
        assert json.loads(urllib.request.urlopen(base + "boxes/0").read()) == saved
    finally:
        srv.shutdown()


def test_txzmq_connection_roundtrip_random_port():
    import zmq
    from veles_b200.txzmq import ZmqConnection, ZmqEndpoint
    a = ZmqConnection(zmq.PAIR, [ZmqEndpoint("bind", "rndtcp://127.0.0.1")])
    b = ZmqConnection(zmq.PAIR, [ZmqEndpoint("connect", a.endpoints[0].address)], codec="gzip")
    payload = {"w": np.arange(10000, dtype=np.float32)}
    b.send(payload)
    _, got = a.receive(5000)
    assert np.array_equal(got["w"], payload["w"])
    assert a.receive(10) is None
    a.shutdown()
    b.shutdown()


def test_dtype_tables_and_plugin_dir(tmp_path, monkeypatch):
    from veles_b200 import dtypes, dot_pip, portable
    assert dtypes.numpy_dtype_to_abi(np.float32) == 0 and dtypes.defines("float")["sizeof_dtype"] == 4
    monkeypatch.setenv("HOME", str(tmp_path))
    p = dot_pip.install_dot_pip(create=True)
    assert os.path.isdir(p)
    assert portable.clock() <= portable.clock()


@pytest.mark.gpu
def test_blas_facade_column_major_sgemm():
    import torch
    from veles_b200.blas import BLAS, OP_N, OP_T
    torch.manual_seed(0)
    m, n, k = 96, 200, 136
    # column-major A[m,k], B[k,n] stored as row-major transposes
    At = torch.randn(k, m, device="cuda").bfloat16()
    Bt = torch.randn(n, k, device="cuda").bfloat16()
    Ct = torch.zeros(n, m, device="cuda")
    BLAS().sgemm(OP_N, OP_N, m, n, k, 1.0, At, Bt, 0.0, Ct)
    ref = (At.float().t() @ Bt.float().t()).t()
    assert ((Ct - ref).abs().max() / ref.abs().max()).item() < 2e-3
    # op(A) = A^T: A stored column-major [k,m] == row-major [m,k]
    A2 = torch.randn(m, k, device="cuda").bfloat16()
    BLAS().sgemm(OP_T, OP_N, m, n, k, 1.0, A2, Bt, 0.0, Ct)
    ref = (A2.float() @ Bt.float().t()).t()
    assert ((Ct - ref).abs().max() / ref.abs().max()).item() < 2e-3


@pytest.mark.gpu
def test_blas_facade_fp32_sgemm_is_tf32_and_dgemm_refuses():
    import torch
    from veles_b200.blas import BLAS, OP_N
    torch.manual_seed(1)
    m, n, k = 128, 192, 256
    At = torch.randn(k, m, device="cuda")
    Bt = torch.randn(n, k, device="cuda")
    Ct = torch.zeros(n, m, device="cuda")
    BLAS().sgemm(OP_N, OP_N, m, n, k, 1.0, At, Bt, 0.0, Ct)
    ref = (At.t() @ Bt.t()).t()
    err = ((Ct - ref).abs().max() / ref.abs().max()).item()
    assert err < 2e-3            # tf32 products: ~1e-3; bf16 operands would sit near 1e-2 here
    with pytest.raises(NotImplementedError):
        BLAS().dgemm(OP_N, OP_N, m, n, k, 1.0, At.double(), Bt.double(), 0.0, Ct.double())


def test_blas_dgemm_refuses_without_a_device():
    from veles_b200.blas import BLAS
    with pytest.raises(NotImplementedError):
        BLAS().dgemm(0, 0, 1, 1, 1, 1.0, None, None, 0.0, None)


@pytest.mark.gpu
def test_sanitizer_smoke_program_runs():
    """The program compute-sanitizer is pointed at (scripts/sanitizer_smoke.py) must at least run clean on its own."""
    import runpy
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    mod = runpy.run_path(os.path.join(root, "scripts", "sanitizer_smoke.py"), run_name="smoke")
    mod["main"]()


def test_package_introspection(tmp_path):
    """veles_b200.__units__ / __modules__ / __loc__ / __plugins__ / allow_root (reference: veles/__init__.py:190-310)."""
    import os
    import veles_b200
    assert os.path.isdir(os.path.join(veles_b200.__root__, "veles_b200"))
    mods = veles_b200.__modules__
    names = {m.__name__ for m in mods}
    assert {"veles_b200.units", "veles_b200.znicz.conv", "veles_b200.loader.fullbatch", "veles_b200.forge.forge_server"} <= names
    units = [c.__name__ for c in veles_b200.__units__]
    assert {"Conv", "All2AllSoftmax", "LRNormalizerForward", "Shell", "Repeater"} <= set(units) and len(units) > 100
    loc = veles_b200.__loc__
    assert loc["core"] > 10000 and loc["cuda"] > 5000 and loc["c/c++"] > 1000 and loc["all"] == sum(
        v for k, v in loc.items() if k != "all")
    assert isinstance(veles_b200.__plugins__, set)
    veles_b200.allow_root = True
    assert veles_b200.allow_root is True
    veles_b200.allow_root = False
