"""Web status, publishing, forge, plotting, compare_snapshots, downloader (reference tests: test_web_status.py,
test_publisher.py, test_forge_client/server.py, test_plotting_units.py, test_downloader.py)."""
import json
import os
import pickle
import urllib.request

import numpy as np
import pytest

from veles_b200.config import root
from veles_b200.dummy import DummyWorkflow


def _post(url, obj):
    req = urllib.request.Request(url, data=json.dumps(obj).encode(), headers={"Content-Type": "application/json"})
    return json.loads(urllib.request.urlopen(req, timeout=10).read())


def test_web_status_update_and_service():
    from veles_b200.web_status import WebServer
    srv = WebServer(port=0, drop_time=3600).start()
    try:
        base = "http://127.0.0.1:%d" % srv.port
        assert _post(base + "/update", {"id": "m1", "name": "wf", "master": "host", "time": 3, "mode": "master",
                                        "slaves": {"s1": {"state": "WORK"}}, "graph": "digraph {}"})["ok"]
        wfs = _post(base + "/service", {"request": "workflows"})
        assert wfs["m1"]["name"] == "wf" and wfs["m1"]["slaves"]["s1"]["state"] == "WORK"
        assert _post(base + "/service", {"request": "graph", "id": "m1"})["graph"] == "digraph {}"
        assert b"wf" in urllib.request.urlopen(base + "/", timeout=10).read()
        srv.drop_time = -1
        _post(base + "/update", {"id": "m2", "name": "x"})
        assert "m1" not in _post(base + "/service", {"request": "workflows"})
    finally:
        srv.stop()


def test_launcher_status_heartbeat_payload():
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.launcher import Launcher, filter_argv
    assert filter_argv(["-l", "0.0.0.0:5000", "-v", "debug", "wf.py", "--nodes=a,b", "-b"],
                       "-l=", "--nodes=", "-b") == ["-v", "debug", "wf.py"]
    l = Launcher(True)
    wf = DummyWorkflow(launcher=l)
    st = l.status()
    assert st["mode"] == "standalone" and st["graph"].startswith("digraph") and st["name"] == wf.name


def test_plotters_and_publisher(tmp_path):
    from veles_b200.backends import NumpyDevice
    from veles_b200.models import build_workflow
    from veles_b200.plotting_units import AccumulatingPlotter, MatrixPlotter, AutoHistogramPlotter, ImagePlotter
    from veles_b200.publishing import Publisher
    root.common.disable.snapshotting = True
    root.common.disable.plotting = False
    root.common.disable.publishing = False
    try:
        wf = build_workflow("mnist_mlp", minibatch_size=50, n_train=300, n_valid=100, max_epochs=2)
        wf.plotters_are_enabled = True
        plots = str(tmp_path / "plots")
        acc = AccumulatingPlotter(wf, name="validation errors", input_field="epoch_n_err_pt", input_offset=1,
                                  output_dir=plots)
        acc.input = wf.decision
        acc.link_from(wf.decision)
        acc.gate_skip = ~wf.loader.epoch_ended
        cm = MatrixPlotter(wf, name="confusion", input_field="confusion_matrix", output_dir=plots)
        cm.input = wf.evaluator
        cm.link_from(wf.decision)
        hist = AutoHistogramPlotter(wf, name="weights", input_field="weights", output_dir=plots)
        hist.input = wf.forwards[0]
        hist.link_from(wf.decision)
        img = ImagePlotter(wf, name="samples", output_dir=plots, count=3)
        img.inputs, img.input_fields = [wf.loader], ["minibatch_data"]
        img.link_from(wf.decision)
        pub = Publisher(wf, backends={"markdown": {"directory": str(tmp_path)}, "html": {"directory": str(tmp_path)},
                                      "json": {"directory": str(tmp_path)}})
        pub.link_from(wf.decision)
        pub.gate_skip = ~wf.decision.complete
        wf.initialize(device=NumpyDevice())
        wf.run()
        wf.thread_pool.wait_idle(10)   # plotters hang off Decision in parallel branches
        assert len(acc.values) >= 2
        for p in (acc, cm, hist, img):
            assert p.last_file and os.path.getsize(p.last_file) > 200 and open(p.last_file).read().startswith("<svg")
        assert len(pub.reports) == 3
        md = open(pub.reports[0]).read()
        assert "EvaluationFitness" in md and "digraph Workflow" in md and "fc_tanh1" in md
        assert "<svg" in open(pub.reports[1]).read()
        assert json.load(open(pub.reports[2]))["results"]["Total epochs"] >= 1
        wf.launcher.stop()
    finally:
        root.common.disable.plotting = True


def test_graphics_server_publishes_to_client(tmp_path):
    from veles_b200.graphics_client import GraphicsClient
    from veles_b200.graphics_server import GraphicsServer
    from veles_b200.plotting_units import AccumulatingPlotter
    import time
    srv = GraphicsServer(publish=True, render_inline=False)
    try:
        client = GraphicsClient(srv.endpoints[0], output_dir=str(tmp_path))
        time.sleep(0.3)   # let the SUB socket connect
        p = AccumulatingPlotter(DummyWorkflow(), name="remote", input_field=None, output_dir=str(tmp_path / "x"))
        p.values = [1.0, 2.0, 1.5]
        for _ in range(5):
            srv.enqueue(p)
            path = client.poll_once(500)
            if path:
                break
        assert path and os.path.dirname(path) == str(tmp_path)
    finally:
        srv.shutdown()


def test_forge_roundtrip(tmp_path):
    from veles_b200.forge.forge_client import ForgeClient
    from veles_b200.forge.forge_server import ForgeServer
    model = tmp_path / "model"
    model.mkdir()
    (model / "wf.py").write_text("def run(load, main): pass\n")
    (model / "cfg.py").write_text("# config\n")
    (model / "manifest.json").write_text(json.dumps({
        "name": "tiny", "workflow": "wf.py", "configuration": "cfg.py", "short_description": "tiny model",
        "long_description": "a tiny test model", "author": "tester"}))
    srv = ForgeServer(str(tmp_path / "store")).start()
    try:
        c = ForgeClient(srv.url)
        with pytest.raises(PermissionError):
            c.upload(str(model))
        c.register("me@example.com")
        assert c.upload(str(model)) == {"name": "tiny", "version": "v1"}
        assert c.upload(str(model))["version"] == "v2"
        assert c.list()[0][:2] == ["tiny", "tiny model"]
        assert c.details("tiny")["versions"] == ["v1", "v2"]
        dest, manifest = c.fetch("tiny", str(tmp_path / "fetched"))
        assert manifest["author"] == "tester" and os.path.exists(os.path.join(dest, "wf.py"))
        other = ForgeClient(srv.url)
        other.register("evil@example.com")
        with pytest.raises(urllib.error.HTTPError):
            other.delete("tiny")
        c.delete("tiny")
        assert c.list() == []
    finally:
        srv.stop()


def test_compare_snapshots(tmp_path):
    from veles_b200.backends import NumpyDevice
    from veles_b200.models import build_workflow
    from veles_b200.scripts.compare_snapshots import compare
    root.common.disable.snapshotting = True
    wf = build_workflow("mnist_mlp", minibatch_size=50, n_train=100, n_valid=0, max_epochs=1)
    wf.initialize(device=NumpyDevice())
    a = pickle.loads(pickle.dumps(wf))
    wf.run()
    b = pickle.loads(pickle.dumps(wf))
    rows = dict((k, d) for k, _, d in compare(a, b))
    assert rows["fc_tanh1.weights"] > 0 and rows["fc_softmax2.bias"] > 0
    assert dict((k, d) for k, _, d in compare(a, a))["fc_tanh1.weights"] == 0
    wf.launcher.stop()


def test_downloader_unpacks_local_archive(tmp_path):
    import tarfile
    from veles_b200.downloader import Downloader
    src = tmp_path / "src"
    src.mkdir()
    (src / "data.txt").write_text("hello")
    arch = tmp_path / "ds.tar.gz"
    with tarfile.open(arch, "w:gz") as t:
        t.add(src / "data.txt", arcname="data.txt")
    d = Downloader(DummyWorkflow(), url=str(arch), directory=str(tmp_path / "out"), files=["data.txt"])
    d.initialize()
    assert (tmp_path / "out" / "data.txt").read_text() == "hello"
    d.initialize()   # second time: nothing to do


def test_frontend_generator_lists_every_flag():
    from veles_b200.scripts.generate_frontend import generate
    text = generate()
    for flag in ("--optimize", "--ensemble-train", "--snapshot", "--master-address", "--listen-address",
                 "--force-numpy", "--sync-run", "--train-ratio", "--job-timeout", "--slave-death-probability",
                 "--result-file", "--dry-run", "--backend", "--device", "--random-seed"):
        assert flag in text, flag
