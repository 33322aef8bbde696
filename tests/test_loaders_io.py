"""Image / file / pickles / saver / interactive / restful / ensemble / zeromq loaders and the RESTful API
(reference tests: test_minibatches_saver_loader.py, test_restful.py, test_zmq_loader.py, test_ensemble_loader.py)."""
import json
import os
import pickle
import threading
import urllib.request

import numpy as np
import pytest

from veles_b200.backends import NumpyDevice
from veles_b200.dummy import DummyWorkflow
from veles_b200.loader.base import TRAIN, VALID, TEST, UserLoaderRegistry

cv2 = pytest.importorskip("cv2")


@pytest.fixture
def image_tree(tmp_path):
    rng = np.random.RandomState(0)
    for split, n in (("train", 6), ("valid", 2)):
        for label in ("cat", "dog"):
            d = tmp_path / split / label
            d.mkdir(parents=True)
            for i in range(n):
                img = rng.randint(0, 256, size=(20, 24, 3)).astype(np.uint8)
                img[:, :, 0] = 200 if label == "cat" else 20
                cv2.imwrite(str(d / ("%d.png" % i)), img)
        (tmp_path / split / "notes.txt").write_text("not an image")
    return tmp_path


def test_full_batch_auto_label_image_loader(image_tree):
    cls = UserLoaderRegistry.loaders["full_batch_auto_label_file_image"]
    l = cls(DummyWorkflow(), train_paths=[str(image_tree / "train")], validation_paths=[str(image_tree / "valid")],
            minibatch_size=5, scale=(16, 16), crop=(12, 12), mirror=True, color_space="RGB")
    l.initialize(device=NumpyDevice())
    assert l.class_lengths == [0, 8, 24]          # mirror doubles the samples
    assert l.original_data.shape == (32, 12, 12, 3) and l.original_data.mem.dtype == np.uint8
    assert sorted(l.labels_mapping) == ["cat", "dog"]
    l.run()
    assert l.minibatch_data.mem.shape == (5, 12, 12, 3) and l.minibatch_class == VALID
    lab = l.minibatch_labels.mem[:5]
    red = l.minibatch_data.mem[:5, :, :, 2].mean(axis=(1, 2))   # BGR file -> RGB: the marker channel is last
    assert ((red > 100) == (lab == l.labels_mapping["cat"])).all()


def test_streaming_file_list_image_loader(image_tree):
    lst = image_tree / "train.txt"
    files = sorted((image_tree / "train").rglob("*.png"))
    lst.write_text("\n".join("%s %s" % (f, f.parent.name) for f in files))
    cls = UserLoaderRegistry.loaders["file_list_image"]
    l = cls(DummyWorkflow(), path_to_train_text_file=str(lst), minibatch_size=4, color_space="GRAY")
    l.initialize(device=NumpyDevice())
    assert l.class_lengths == [0, 0, 12]
    l.run()
    assert l.minibatch_data.mem.shape == (4, 20, 24, 1) and l.minibatch_size == 4
    assert set(l.minibatch_labels.mem[:4]) <= {0, 1}


def test_pickles_loader(tmp_path):
    rng = np.random.RandomState(1)
    for name, n in (("train", 30), ("test", 10)):
        with open(tmp_path / name, "wb") as f:
            pickle.dump({"data": rng.randint(0, 256, size=(n, 3 * 8 * 8)).astype(np.uint8),
                         "labels": rng.randint(0, 10, size=n).tolist()}, f)
    cls = UserLoaderRegistry.loaders["full_batch_pickles_image"]
    l = cls(DummyWorkflow(), train_pickles=[str(tmp_path / "train")], test_pickles=[str(tmp_path / "test")],
            image_shape=(8, 8, 3), minibatch_size=10)
    l.initialize(device=NumpyDevice())
    assert l.class_lengths == [10, 0, 30] and l.original_data.shape == (40, 8, 8, 3)


def test_hdf5_loader_accepts_npz(tmp_path):
    np.savez(tmp_path / "train.npz", data=np.random.rand(12, 5).astype(np.float32), label=np.arange(12) % 3)
    cls = UserLoaderRegistry.loaders["full_batch_hdf5"]
    l = cls(DummyWorkflow(), train_path=str(tmp_path / "train.npz"), minibatch_size=4)
    l.initialize(device=NumpyDevice())
    assert l.class_lengths == [0, 0, 12] and l.unique_labels_count == 3


def test_minibatches_saver_and_loader_roundtrip(tmp_path):
    from veles_b200.loader.saver import MinibatchesSaver, MinibatchesLoader
    from veles_b200.loader.synthetic import SyntheticFullBatchLoader
    wf = DummyWorkflow()
    src = SyntheticFullBatchLoader(wf, minibatch_size=16, n_train=40, n_valid=20, sample_shape=(6,))
    src.initialize(device=NumpyDevice())
    saver = MinibatchesSaver(wf, file_name=str(tmp_path / "mb.dat"), compression="xz")
    saver.link_attrs(src, "minibatch_data", "minibatch_labels", "minibatch_class", "class_lengths",
                     "max_minibatch_size", "minibatch_size")
    saver.initialize()
    served = []
    while True:
        src.run()
        saver.run()
        served.append((src.minibatch_class, src.minibatch_data.mem[:src.minibatch_size].copy(),
                       src.minibatch_labels.mem[:src.minibatch_size].copy()))
        if src.minibatch_class == TRAIN and src.last_minibatch:
            break
    saver.stop()
    l = MinibatchesLoader(DummyWorkflow(), file_name=str(tmp_path / "mb.dat"), minibatch_size=16)
    l.initialize(device=NumpyDevice())
    assert l.class_lengths == [0, 20, 40]
    for klass, data, labels in served:
        l.run()
        assert l.minibatch_class == klass and l.minibatch_size == len(data)
        assert np.allclose(l.minibatch_data.mem[:len(data)], data)
        assert np.array_equal(l.minibatch_labels.mem[:len(data)], labels)


def test_interactive_loader_feed():
    from veles_b200.loader.interactive import InteractiveLoader
    l = InteractiveLoader(DummyWorkflow(), sample_shape=(2, 3), minibatch_size=2)
    l.initialize(device=NumpyDevice())
    l.feed(np.arange(6))
    l.feed([[1, 1, 1], [2, 2, 2]])
    l.run()
    assert l.minibatch_size == 2 and l.minibatch_data.mem[0, 1, 2] == 5 and l.minibatch_data.mem[1, 1, 0] == 2
    l.feed(None)
    l.run()
    assert l.complete and l.minibatch_size == 0


def test_restful_api_serves_model_over_http():
    """POST /api -> RestfulLoader -> All2AllSoftmax -> answer (the reference's test_restful.py scenario)."""
    from veles_b200.loader.restful import RestfulLoader
    from veles_b200.restful_api import RESTfulAPI
    from veles_b200.znicz.all2all import All2AllSoftmax
    from veles_b200.plumbing import Repeater
    wf = DummyWorkflow()
    wf.end_point.unlink_all()
    rep = Repeater(wf)
    rep.link_from(wf.start_point)
    loader = RestfulLoader(wf, sample_shape=(4,), minibatch_size=3, max_response_time=0.05, timeout=5.0)
    loader.link_from(rep)
    fwd = All2AllSoftmax(wf, output_sample_shape=3)
    fwd.link_from(loader)
    fwd.link_attrs(loader, ("input", "minibatch_data"))
    api = RESTfulAPI(wf, port=0)
    api.link_from(fwd)
    api.feed = loader.feed
    api.link_attrs(loader, "requests", "minibatch_size")
    api.link_attrs(fwd, ("results", "output"))
    rep.link_from(api)
    rep.gate_block = loader.complete
    wf.end_point.link_from(api)
    wf.end_point.gate_block = ~loader.complete
    wf.initialize(device=NumpyDevice())
    t = threading.Thread(target=wf.run, daemon=True)
    t.start()
    answers = []
    for k in range(3):
        req = urllib.request.Request("http://127.0.0.1:%d/api" % api.port,
                                     data=json.dumps({"input": [0.1 * k, 0.2, 0.3, 0.4], "codec": "list"}).encode(),
                                     headers={"Content-Type": "application/json"})
        answers.append(json.loads(urllib.request.urlopen(req, timeout=10).read())["result"])
    assert all(len(a) == 3 and abs(sum(a) - 1.0) < 1e-4 for a in answers)
    bad = urllib.request.Request("http://127.0.0.1:%d/api" % api.port, data=b"{nonsense")
    with pytest.raises(urllib.error.HTTPError):
        urllib.request.urlopen(bad, timeout=10)
    loader.feed(None)
    t.join(10)
    assert not t.is_alive()
    wf.launcher.stop()


def test_ensemble_loader(tmp_path):
    rng = np.random.RandomState(2)
    data = {"models": [{"Output": rng.rand(20, 3).tolist()}, {"Output": rng.rand(20, 3).tolist()}],
            "labels": (np.arange(20) % 3).tolist()}
    path = tmp_path / "ens.json"
    path.write_text(json.dumps(data))
    cls = UserLoaderRegistry.loaders["ensemble"]
    l = cls(DummyWorkflow(), file=str(path), minibatch_size=5)
    l.initialize(device=NumpyDevice())
    assert l.original_data.shape == (20, 2, 3) and l.class_lengths == [0, 4, 16]


def test_zeromq_loader_accepts_external_work():
    import zmq
    from veles_b200.zmq_loader import ZeroMQLoader
    l = ZeroMQLoader(DummyWorkflow(), sample_shape=(3,), minibatch_size=1, timeout=5.0)
    l.initialize(device=NumpyDevice())
    assert l.generate_initial_data_for_master()["endpoints"]["tcp"].startswith("tcp://127.0.0.1:")
    sock = zmq.Context.instance().socket(zmq.DEALER)
    sock.connect(l.endpoints["tcp"])
    sock.send_multipart([b"cid-1", pickle.dumps(np.array([1.0, 2.0, 3.0]))])
    assert sock.poll(5000) and sock.recv_multipart()[-1] == b"ok"
    l.run()
    assert l.minibatch_size == 1 and np.allclose(l.minibatch_data.mem[0], [1, 2, 3])
    l.stop()
    sock.close()


def test_stream_loader_autoencoder_targets_train_on_cpu():
    """StreamLoader with targets="input" (autoencoder): Trainer.step on host-fed minibatches lowers the MSE."""
    import numpy as np
    from veles_b200.backends import NumpyDevice
    from veles_b200.trainer import Trainer
    tr = Trainer("wide_mlp_ae", minibatch_size=8, device=NumpyDevice(), streaming=True,
                 layers_kwargs={"width": 32, "depth": 2, "in_features": 4096})
    x = np.random.RandomState(0).rand(8, 4096).astype(np.float32)
    first = tr.step(x, None)
    for _ in range(15):
        last = tr.step(x, None)
    assert last < first
    tr.close()


def test_pipelined_step_returns_each_loss_one_call_late():
    """Trainer.step_pipelined: the loss of step i comes back from call i + 1 (flush_loss for the last one); the
    sequence equals what step() returns for the same minibatches."""
    import numpy as np
    from veles_b200.backends import NumpyDevice
    from veles_b200.trainer import Trainer
    xs = [np.random.RandomState(k).rand(8, 4096).astype(np.float32) for k in range(5)]
    seqs = []
    for pipelined in (False, True):
        tr = Trainer("wide_mlp_ae", minibatch_size=8, device=NumpyDevice(), streaming=True, seed=3,
                     layers_kwargs={"width": 32, "depth": 2, "in_features": 4096})
        if pipelined:
            got = [tr.step_pipelined(x, None) for x in xs]
            assert got[0] is None and tr._pending_loss is not None
            got = got[1:] + [tr.flush_loss()]
            assert tr.flush_loss() is None                     # nothing left
        else:
            got = [tr.step(x, None) for x in xs]
        seqs.append(got)
        tr.close()
    assert np.allclose(seqs[0], seqs[1], rtol=1e-6) and len(seqs[1]) == 5


def test_label_statistics_and_progress_spinner(caplog):
    """Loader start-up reports how evenly each class's labels are populated (reference: loader/base.py:947-980); the
    unit-run spinner writes only to terminals."""
    import io
    import logging
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.loader.base import Loader
    from veles_b200.units import spin, _SPIN
    ld = Loader.__new__(Loader)
    ld._logger_ = logging.getLogger("Loader")
    with caplog.at_level(logging.DEBUG):
        even = Loader._print_label_stats(ld, {0: 100, 1: 104, 2: 97}, "train")
        skew = Loader._print_label_stats(ld, {"cat": 5, "dog": 500}, "validation")
        none = Loader._print_label_stats(ld, {}, "test")
    assert 'min: 97 ("2"), max: 104 ("1")' in even and "UNBALANCED" not in even
    assert skew.endswith("UNBALANCED") and 'max: 500 ("dog")' in skew and none is None
    assert any("Detailed train label stats" in r.getMessage() for r in caplog.records)
    assert any(r.levelno == logging.WARNING and "UNBALANCED" in r.getMessage() for r in caplog.records)
    buf = io.StringIO()
    assert spin(buf) and len(buf.getvalue()) == 2 and buf.getvalue()[1] == "\b"
    _SPIN["enabled"] = None
    assert spin() is False and _SPIN["enabled"] is False        # pytest's stdout is not a terminal: run() never spins
