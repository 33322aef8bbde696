"""Thread pool behaviour, snapshot codecs, MSE image loader (reference: tests/test_thread_pool.py,
test_snapshotter-ish parts of test_workflow.py, loader/image_mse.py)."""
import os
import threading
import time

import numpy as np
import pytest


def test_thread_pool_pause_resume_failure_and_shutdown_callbacks():
    from veles_b200.thread_pool import ThreadPool
    pool = ThreadPool(minthreads=2, maxthreads=2, name="t")
    pool.silent = True
    pool.start()
    done = []
    pool.pause()
    assert pool.paused
    for i in range(4):
        pool.callInThread(done.append, i)
    time.sleep(0.1)
    assert done == []                       # nothing runs while paused
    pool.resume()
    assert pool.wait_idle(5)
    assert sorted(done) == [0, 1, 2, 3]
    failures = []
    pool.on_failures.append(failures.append)

    def boom():
        raise ValueError("x")
    pool.callInThread(boom)
    assert pool.wait_idle(5)
    assert len(failures) == 1 and isinstance(pool.failure, ValueError)
    called = []

    class Owner(object):
        def bye(self):
            called.append(threading.current_thread().name)
    o = Owner()
    pool.register_on_shutdown(o.bye)
    pool.shutdown()
    assert called and not pool.started
    pool.shutdown()   # idempotent


def test_thread_pool_stack_dump_lists_every_thread():
    from veles_b200.thread_pool import ThreadPool
    text = ThreadPool.dump_stacks()      # what SIGUSR1 prints
    assert "MainThread" in text and "test_thread_pool_stack_dump" in text


@pytest.mark.parametrize("compression", ["", "gz", "bz2", "xz"])
def test_snapshot_codecs_roundtrip_whole_workflow(tmp_path, compression):
    from veles_b200.backends import NumpyDevice
    from veles_b200.config import root
    from veles_b200.models import build_workflow
    from veles_b200.snapshotter import SnapshotterToFile
    root.common.disable.snapshotting = True
    wf = build_workflow("mnist_mlp", max_epochs=1, minibatch_size=50, n_train=100, n_valid=50)
    wf.initialize(device=NumpyDevice())
    wf.run()
    snap = SnapshotterToFile(wf, directory=str(tmp_path), prefix="t", compression=compression, time_interval=0)
    snap.suffix = "validation_1.00"
    snap.export()
    assert os.path.exists(snap.destination)
    assert snap.destination.endswith(".pickle" + ("." + compression if compression else ""))
    back = SnapshotterToFile.import_(snap.destination)
    assert back.restored_from_snapshot
    w0 = wf.forwards[0].weights.mem
    w1 = back.forwards[0].weights.mem
    assert np.array_equal(w0, w1)
    assert back.loader.epoch_number == wf.loader.epoch_number
    assert os.path.islink(os.path.join(str(tmp_path), os.path.basename(snap.destination).replace(
        "_validation_1.00", "_current")))
    wf.launcher.stop()


@pytest.mark.parametrize("compression", ["", "gz", "xz"])
def test_snapshot_to_database_roundtrip(tmp_path, compression):
    """SnapshotterToDB over the DB-API module that ships with Python (sqlite:<path>); pyodbc is optional."""
    from veles_b200.backends import NumpyDevice
    from veles_b200.config import root
    from veles_b200.models import build_workflow
    from veles_b200.snapshotter import Snapshotter, SnapshotterToDB
    root.common.disable.snapshotting = True
    wf = build_workflow("mnist_mlp", max_epochs=1, minibatch_size=50, n_train=100, n_valid=50)
    wf.initialize(device=NumpyDevice())
    wf.run()
    dsn = "sqlite:" + str(tmp_path / "snapshots.db")
    snap = Snapshotter(wf, odbc=dsn, table="runs", prefix="t", compression=compression, time_interval=0)
    assert isinstance(snap, SnapshotterToDB)
    snap.initialize()
    wf.launcher.id, wf.launcher.log_id = "job-1", "log-1"
    snap.suffix = "first"
    snap.export()
    first_w = wf.forwards[0].weights.mem.copy()
    wf.forwards[0].weights.mem[...] += 1.0
    snap.suffix = "second"
    snap.export()
    newest = SnapshotterToDB.import_(dsn, "runs", "job-1", "log-1")
    assert newest.restored_from_snapshot
    assert np.array_equal(newest.forwards[0].weights.mem, first_w + 1.0)
    named = Snapshotter.import_odbc(dsn, "runs", "job-1", "log-1", name=snap.destination.replace("second", "first"))
    assert np.array_equal(named.forwards[0].weights.mem, first_w)
    with pytest.raises(KeyError):
        SnapshotterToDB.import_(dsn, "runs", "nobody", "log-1")
    assert snap.get_metric_values()["Snapshot"]["table"] == "runs"
    # the CLI's odbc:// locator resolves to the same call
    from veles_b200.__main__ import Main
    m = Main.__new__(Main)
    m.error = lambda *a: None
    back = m._load_workflow("odbc://%s&runs&job-1&log-1&" % dsn + snap.destination)
    assert back is not None and np.array_equal(back.forwards[0].weights.mem, first_w + 1.0)
    assert m._load_workflow("odbc://%s&runs" % dsn) is None
    wf.launcher.stop()


def _png(path, rgb, size=8):
    from PIL import Image
    Image.fromarray(np.full((size, size, 3), rgb, dtype=np.uint8)).save(path)


def test_auto_label_image_loader_mse_serves_target_images(tmp_path):
    pytest.importorskip("PIL")
    from veles_b200.backends import NumpyDevice
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.loader.image_mse import AutoLabelFileImageLoaderMSE
    for label, rgb in (("cat", (200, 10, 10)), ("dog", (10, 10, 200))):
        os.makedirs(str(tmp_path / "train" / label))
        for i in range(3):
            _png(str(tmp_path / "train" / label / ("%d.png" % i)), rgb)
        os.makedirs(str(tmp_path / "targets" / label), exist_ok=True)
        _png(str(tmp_path / "targets" / label / "t.png"), tuple(255 - c for c in rgb))
    ld = AutoLabelFileImageLoaderMSE(
        DummyWorkflow(), train_paths=[str(tmp_path / "train")], target_paths=[str(tmp_path / "targets")],
        label_regexp=r".*/(\w+)/[^/]+$", minibatch_size=4, normalization_type="none")
    ld.initialize(device=NumpyDevice())
    ld.run()
    n = ld.minibatch_size
    assert ld.minibatch_targets.mem.shape[0] == 4 and n > 0
    for i in range(n):
        x, t = ld.minibatch_data.mem[i], ld.minibatch_targets.mem[i]
        # the target of every sample is the colour-inverted image of its label
        assert np.abs((255.0 - x) - t).max() < 1.0 or np.abs((1.0 - x) - t).max() < 1e-2


def test_manhole_repl_over_unix_socket(tmp_path):
    """``--manhole`` / SIGUSR2 (reference: veles/thread_pool.py:139-164, 527-533): a REPL on a UNIX socket whose
    namespace holds the workflow and the launcher."""
    import os
    import stat
    from veles_b200 import manhole

    class FakeWorkflow(object):
        name = "wf-under-inspection"

    path = str(tmp_path / "manhole.sock")
    assert manhole.install(workflow=FakeWorkflow(), launcher="the launcher", path=path) == path
    try:
        assert stat.S_IMODE(os.stat(path).st_mode) == 0o600
        out = manhole.connect(path, ["print(workflow.name)", "x = 6 * 7", "x", "launcher", "1/0"])
        assert "wf-under-inspection" in out and "42" in out and "the launcher" in out
        assert "ZeroDivisionError" in out          # errors are reported to the client, the server survives
        out2 = manhole.connect(path, ["dump_stacks is not None"])
        assert "True" in out2
        assert manhole.install(path=path, workflow=None) == path     # idempotent
    finally:
        manhole.uninstall(path)
    assert not os.path.exists(path)
