"""Master and slave agents in ONE process on localhost (pattern of veles/tests/test_network.py:111-138,
veles/tests/test_launcher.py:91-121): stub workflows record which IDistributable hooks fired; plus the
BASELINE config 1 end to end — MNIST-shape MLP, numpy backend, master + 1 slave."""
import threading
import time

import numpy as np
import pytest

from veles_b200.client import Client
from veles_b200.dummy import DummyLauncher
from veles_b200.network_common import dumps, loads, SharedIO, CODECS
from veles_b200.server import Server
from veles_b200.workflow import Workflow, NoMoreJobs


class StubWorkflow(Workflow):
    hide_from_registry = True

    def __init__(self, launcher, n_jobs=5, **kwargs):
        super(StubWorkflow, self).__init__(launcher, **kwargs)
        self.end_point.link_from(self.start_point)
        self.n_jobs = n_jobs
        self.served = 0
        self.flags = {k: 0 for k in ("job_requested", "job_done", "update_applied", "dropped", "initial")}
        self.lock = threading.Lock()

    @property
    def checksum(self):
        return "stub"

    def generate_initial_data_for_slave(self, slave=None):
        self.flags["initial"] += 1
        return None

    def generate_data_for_slave(self, slave=None):
        with self.lock:
            if self.served >= self.n_jobs:
                return None
            self.served += 1
            self.flags["job_requested"] += 1
            return [{"job": self.served, "blob": np.arange(30000)}]

    def apply_data_from_slave(self, data, slave=None):
        self.flags["update_applied"] += 1
        assert data[0]["echo"] > 0
        return True

    def drop_slave(self, slave=None):
        self.flags["dropped"] += 1

    def do_job(self, data, update, callback):
        self.flags["job_done"] += 1
        callback([{"echo": data[0]["job"], "blob": data[0]["blob"] * 2}])


def _run_pair(master_wf, slave_wf, port, **client_kw):
    server = Server("127.0.0.1:%d" % port, master_wf, job_timeout=5.0, heartbeat_timeout=3.0)
    client = Client("127.0.0.1:%d" % port, slave_wf, **client_kw)
    err = []

    def guarded(fn):
        def run():
            try:
                fn()
            except Exception as e:  # noqa: BLE001
                err.append(e)
        return run
    ts = threading.Thread(target=guarded(server.run), daemon=True)
    tc = threading.Thread(target=guarded(client.run), daemon=True)
    ts.start()
    tc.start()
    tc.join(30)
    ts.join(30)
    assert not ts.is_alive() and not tc.is_alive(), "agents did not finish"
    return server, client, err


def test_job_update_cycle_in_one_process():
    master = StubWorkflow(DummyLauncher(mode="master"), n_jobs=6)
    slave = StubWorkflow(DummyLauncher(mode="slave"))
    server, client, err = _run_pair(master, slave, 5071)
    assert not err, err
    assert master.flags["job_requested"] == 6 and master.flags["update_applied"] == 6
    assert slave.flags["job_done"] == 6 and master.flags["initial"] == 1
    assert client.state == "END"


def test_async_slave_overlaps_requests():
    master = StubWorkflow(DummyLauncher(mode="master"), n_jobs=8)
    slave = StubWorkflow(DummyLauncher(mode="slave"))
    server, client, err = _run_pair(master, slave, 5072, async_slave=True)
    assert not err and master.flags["update_applied"] == 8


def test_checksum_mismatch_is_rejected():
    class Other(StubWorkflow):
        @property
        def checksum(self):
            return "other"
    master = StubWorkflow(DummyLauncher(mode="master"), n_jobs=1)
    slave = Other(DummyLauncher(mode="slave"))
    server = Server("127.0.0.1:5073", master)
    ts = threading.Thread(target=server.run, daemon=True)
    ts.start()
    client = Client("127.0.0.1:5073", slave)
    from veles_b200.error import MasterSlaveCommunicationError
    with pytest.raises(MasterSlaveCommunicationError):
        client.run()
    server.stop()
    ts.join(5)


def test_dead_slave_is_dropped_and_job_requeued():
    master = StubWorkflow(DummyLauncher(mode="master"), n_jobs=3)
    slave = StubWorkflow(DummyLauncher(mode="slave"))
    server = Server("127.0.0.1:5074", master, job_timeout=0.5, heartbeat_timeout=0.7)
    ts = threading.Thread(target=server.run, daemon=True)
    ts.start()
    client = Client("127.0.0.1:5074", slave, death_probability=1.0)
    from veles_b200.error import Bug
    with pytest.raises(Bug):
        client.run()                      # dies on the first job (fault injection)
    deadline = time.time() + 10
    while master.flags["dropped"] == 0 and time.time() < deadline:
        time.sleep(0.1)
    assert master.flags["dropped"] == 1
    server.stop()
    ts.join(5)


@pytest.mark.parametrize("codec", [c for c in CODECS if c])
def test_payload_framing_roundtrip(codec):
    obj = {"a": np.arange(1000), "b": "text", "c": [1, 2, 3]}
    frames = dumps(obj, codec)
    assert frames[0].startswith(b"vpb") and frames[-1] == b"vpe"
    back = loads(frames)
    assert np.array_equal(back["a"], obj["a"]) and back["b"] == "text"


def test_shared_io_between_processes():
    import multiprocessing as mp
    payload = b"x" * 100000

    def child(name, q):
        q.put(SharedIO.get(name) == payload)
    SharedIO.put("veles-test-shm", payload)
    q = mp.Queue()
    p = mp.Process(target=child, args=("veles-test-shm", q))
    p.start()
    assert q.get(timeout=20)
    p.join()
    SharedIO.remove("veles-test-shm")


def test_mnist_mlp_master_plus_one_slave_numpy():
    """BASELINE config 1: 784-100-10 All2AllTanh+Softmax, GD, numpy backend, master + 1 slave on localhost."""
    from veles_b200 import prng
    from veles_b200.backends import NumpyDevice
    from veles_b200.config import root
    from veles_b200.models import build_workflow
    root.common.disable.snapshotting = True

    def make(mode):
        prng.get().seed(5)
        wf = build_workflow("mnist_mlp", launcher=DummyLauncher(mode=mode), minibatch_size=100, n_train=1000,
                            n_valid=200, max_epochs=3, loader_extra={"noise": 1.5}, layers_kwargs={"lr": 0.05})
        wf.initialize(device=NumpyDevice())
        return wf
    master, slave = make("master"), make("slave")
    server, client, err = _run_pair(master, slave, 5075)
    assert not err, err
    hist = master.decision.epoch_metrics_history
    assert len(hist) >= 3
    assert hist[-1]["n_err_pt"][1] < 20.0, hist
    assert master.decision.complete
    assert client.jobs_done >= 3 * 12
