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
        assert b"<td>wf</td>" in urllib.request.urlopen(base + "/plain", timeout=10).read()
        page = urllib.request.urlopen(base + "/", timeout=10).read()
        assert b"<script>" in page and b"request:'logs'" in page      # the polling UI
        srv.drop_time = -1
        _post(base + "/update", {"id": "m2", "name": "x"})
        assert "m1" not in _post(base + "/service", {"request": "workflows"})
    finally:
        srv.stop()


def test_web_status_queries_logs_and_events_of_other_processes(tmp_path):
    """Log / event queries by session, node, level, text (reference: MongoDB queries of veles/web_status.py:113-186)
    over this process' sinks and the JSON-lines files other processes wrote."""
    import logging
    from veles_b200.logger import Logger, events, install_log_sink, logs
    from veles_b200.web_status import WebServer
    other = tmp_path / "9.9-1.logs.jsonl"
    other.write_text("\n".join(json.dumps(r) for r in [
        {"session": "9.9", "node": "far:1", "time": 10.0, "level": "INFO", "levelno": 20, "name": "Loader",
         "message": "epoch 1 done"},
        {"session": "9.9", "node": "far:1", "time": 11.0, "level": "ERROR", "levelno": 40, "name": "Decision",
         "message": "diverged"}]) + "\n")
    (tmp_path / "9.9-1.events.jsonl").write_text(json.dumps(
        {"session": "9.9", "time": 10.5, "domain": "Workflow", "name": "run", "type": "begin"}) + "\n")
    install_log_sink(str(tmp_path / "own.logs.jsonl"))
    try:
        lg = logging.getLogger("QueryTest")
        lg.setLevel(logging.INFO)
        lg.warning("needle in %s", "haystack")
        Logger().event("probe", "single", height=3)
        srv = WebServer(port=0, log_dir=str(tmp_path)).start()
        try:
            url = "http://127.0.0.1:%d/service" % srv.port
            far = _post(url, {"request": "logs", "session": "9.9"})["logs"]
            assert [r["message"] for r in far] == ["epoch 1 done", "diverged"]
            assert [r["message"] for r in _post(url, {"request": "logs", "session": "9.9", "level": "warning"})["logs"]] \
                == ["diverged"]
            mine = _post(url, {"request": "logs", "text": "NEEDLE"})["logs"]
            assert len(mine) == 1 and mine[0]["session"] == events.session and mine[0]["level"] == "WARNING"
            assert _post(url, {"request": "logs", "node": "far:1", "since": 10.5, "limit": 5})["logs"][0]["time"] == 11.0
            ev = _post(url, {"request": "events", "session": "9.9"})["events"]
            assert len(ev) == 1 and ev[0]["name"] == "run"
            own = _post(url, {"request": "events", "name": "probe"})["events"]
            assert own and own[-1]["height"] == 3
            sessions = _post(url, {"request": "sessions"})["sessions"]
            assert "9.9" in sessions and events.session in sessions
        finally:
            srv.stop()
    finally:
        logging.getLogger().removeHandler(logs)


def test_frontend_composer_serves_the_form_and_returns_the_posted_command_line():
    """``--frontend`` (reference: veles/__main__.py:258-327): the form lists the parser's options; the argv the page
    posts comes back from serve()."""
    import threading
    from veles_b200.scripts.generate_frontend import generate, scan_workflows, serve
    page = generate(["wf.py", "cfg.py", "root.a.b=1"])
    assert "data-opt='--snapshot'" in page or "data-opt='-w'" in page
    offered = dict(scan_workflows())
    assert offered["veles_b200/znicz/samples/mnist.py"].endswith("mnist_config.py") and len(offered) >= 6
    assert not any("optimization_workflow" in w for w in offered)
    assert "<option value='veles_b200/znicz/samples/alexnet.py'>" in page
    assert "value='wf.py'" in page and "function compose()" in page and "--frontend" not in page.split("<table>")[1]
    got = {}

    def drive(port):
        def client():
            base = "http://127.0.0.1:%d" % port
            got["page"] = urllib.request.urlopen(base + "/", timeout=10).read()
            req = urllib.request.Request(base + "/cmdline", data=json.dumps(
                {"cmdline": "-v debug 'my wf.py' - root.x.y=3"}).encode())
            urllib.request.urlopen(req, timeout=10).read()
        threading.Thread(target=client, daemon=True).start()

    argv = serve(["wf.py"], open_browser=False, ready=drive, timeout=20)
    assert argv == ["-v", "debug", "my wf.py", "-", "root.x.y=3"]
    assert b"Run</button>" in got["page"]
    assert serve([], open_browser=False, timeout=0.2) is None       # nobody posted: cancelled


def test_yarn_node_discovery_against_a_local_resource_manager():
    """``--yarn-nodes`` (reference: launcher.py:887-906): RUNNING nodes of /ws/v1/cluster/nodes become node specs; an
    unreachable manager only warns."""
    import threading
    from http.server import BaseHTTPRequestHandler, HTTPServer
    from veles_b200.launcher import Launcher
    seen = []

    class RM(BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def do_GET(self):
            seen.append(self.path)
            body = json.dumps({"nodes": {"node": [
                {"nodeHostName": "gpu-a", "state": "RUNNING"}, {"nodeHostName": "gpu-b", "state": "LOST"},
                {"nodeHostName": "gpu-c", "state": "RUNNING"}, {"nodeHostName": "gpu-a", "state": "RUNNING"}]}}).encode()
            self.send_response(200)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

    srv = HTTPServer(("127.0.0.1", 0), RM)
    threading.Thread(target=srv.serve_forever, kwargs={"poll_interval": 0.05}, daemon=True).start()
    try:
        l = Launcher(True)
        nodes = l.discover_nodes_from_yarn("127.0.0.1:%d" % srv.server_address[1])
        assert nodes == ["gpu-a/auto:0", "gpu-c/auto:0"] and seen == ["/ws/v1/cluster/nodes"]
    finally:
        srv.shutdown()
        srv.server_close()
    assert Launcher(True).discover_nodes_from_yarn("127.0.0.1:9", timeout=0.5) == []


def test_launcher_status_heartbeat_payload():
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.launcher import Launcher, filter_argv
    assert filter_argv(["-l", "0.0.0.0:5000", "-v", "debug", "wf.py", "--nodes=a,b", "-b"],
                       "-l=", "--nodes=", "-b") == ["-v", "debug", "wf.py"]
    l = Launcher(True)
    wf = DummyWorkflow(launcher=l)
    st = l.status()
    assert st["mode"] == "standalone" and st["graph"].startswith("digraph") and st["name"] == wf.name
    assert st["graph_svg"].startswith("<svg") and st["session"]


def test_workflow_graph_picture_layers_units_and_marks_the_loop():
    """generate_graph_svg: one box per unit, forward edges between consecutive layers, the Repeater loop as a dashed
    back edge, gate-blocked edges in red; well-formed XML."""
    import xml.etree.ElementTree as ET
    from veles_b200.backends import NumpyDevice
    from veles_b200.models import build_workflow
    root.common.disable.snapshotting = True
    wf = build_workflow("mnist_mlp", minibatch_size=50, n_train=100, n_valid=50, max_epochs=1)
    wf.initialize(device=NumpyDevice())
    svg = wf.generate_graph_svg()
    tree = ET.fromstring(svg)
    ns = "{http://www.w3.org/2000/svg}"
    boxes = tree.findall(".//%srect" % ns)
    titles = [t.text for t in tree.findall(".//%stitle" % ns)]
    assert len(boxes) == len(titles) >= 8 and any("Repeater" in t for t in titles) and any("Decision" in t for t in titles)
    paths = tree.findall("%spath" % ns)
    dashed = [p for p in paths if p.get("stroke-dasharray")]
    assert dashed and len(dashed) < len(paths)                     # the training loop closes through back edges
    assert any(p.get("stroke") == "#c00" for p in paths)           # end point / gds are gate-blocked
    ys = sorted({float(b.get("y")) for b in boxes})
    assert len(ys) >= 5                                            # start -> repeater -> loader -> forwards -> ... layers
    wf.launcher.stop()


def test_status_and_forge_servers_run_as_processes(tmp_path):
    """Both servers have command-line entry points (the reference runs them as daemons): start, answer, stop on SIGINT."""
    import signal
    import socket
    import subprocess
    import sys
    import time

    def free_port():
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            return s.getsockname()[1]

    env = dict(os.environ, PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    p_status, p_forge = free_port(), free_port()
    procs = [subprocess.Popen([sys.executable, "-m", "veles_b200.web_status", "--port", str(p_status), "--log-dir",
                               str(tmp_path)], env=env, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL),
             subprocess.Popen([sys.executable, "-m", "veles_b200.forge.forge_server", "--root", str(tmp_path / "zoo"),
                               "--port", str(p_forge), "--confirm"], env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.DEVNULL)]
    try:
        assert b"status server on" in procs[0].stdout.readline()
        assert b"forge server on" in procs[1].stdout.readline()
        assert _post("http://127.0.0.1:%d/service" % p_status, {"request": "workflows"}) == {}
        base = "http://127.0.0.1:%d/forge" % p_forge
        assert json.loads(urllib.request.urlopen(base + "/service?query=list", timeout=10).read()) == []
        tok = json.loads(urllib.request.urlopen(base + "/service?query=register&email=a@b.c", timeout=10).read())["token"]
        assert tok and os.path.exists(tmp_path / "zoo" / "outbox" / "a@b.c.txt")     # --confirm: the code was "mailed"
    finally:
        for p in procs:
            p.send_signal(signal.SIGINT)
        assert [p.wait(timeout=30) for p in procs] == [0, 0]


def test_forge_keeps_a_git_history_of_every_model(tmp_path):
    """Every upload is a tagged commit in the model's git repository on the server (reference:
    forge_server.py:462-560): history and version-to-version diffs come from it."""
    import shutil
    if shutil.which("git") is None:
        pytest.skip("no git executable")
    from veles_b200.forge.forge_client import ForgeClient
    from veles_b200.forge.forge_server import ForgeServer
    model = tmp_path / "model"
    model.mkdir()
    (model / "wf.py").write_text("def run(load, main): pass\n")
    (model / "cfg.py").write_text("# config\n")
    manifest = {"name": "hist", "workflow": "wf.py", "configuration": "cfg.py", "short_description": "first cut",
                "long_description": "x", "author": "tester"}
    (model / "manifest.json").write_text(json.dumps(manifest))
    srv = ForgeServer(str(tmp_path / "store")).start()
    try:
        c = ForgeClient(srv.url)
        c.register("me@example.com")
        c.upload(str(model))
        (model / "cfg.py").write_text("# config\nroot.x = 1\n")
        (model / "extra.txt").write_text("notes\n")
        (model / "manifest.json").write_text(json.dumps(dict(manifest, short_description="second cut",
                                                             files=["extra.txt"])))
        c.upload(str(model))
        h = c.history("hist")
        assert [e["version"] for e in h] == ["v2", "v1"]
        assert h[0]["subject"] == "v2: second cut" and h[1]["author"] == "tester <me@example.com>"
        assert c.details("hist")["history"][0]["commit"] == h[0]["commit"]
        d = c.diff("hist", "v1", "v2")
        assert "cfg.py" in d and "extra.txt" in d and "wf.py" not in d
        # the browser pages: index with a link per model, details with history and downloads
        index = urllib.request.urlopen(srv.url, timeout=10).read().decode()
        assert "VelesForge: 1 model(s)" in index and "<a href='?name=hist'>hist</a>" in index and "second cut" in index
        page = urllib.request.urlopen(srv.url + "?name=hist", timeout=10).read().decode()
        assert "fetch?name=hist&version=v2" in page and "v1: first cut" in page and "&quot;workflow&quot;" in page
        with pytest.raises(urllib.error.HTTPError):
            urllib.request.urlopen(srv.url + "?name=nope", timeout=10)
        with pytest.raises(urllib.error.HTTPError):
            c.diff("hist", "v1", "v9")
    finally:
        srv.stop()


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
        # the browser view: index.html names the newest file of the plotter and reloads itself
        index = (tmp_path / "index.html").read_text()
        assert "http-equiv='refresh'" in index and os.path.basename(path) in index and "remote" in index
        report = client.write_report(str(tmp_path / "all.html"))
        assert "<svg" in open(report).read() and "<h3>remote</h3>" in open(report).read()
        assert not client.finished
    finally:
        srv.shutdown()
    # the server's goodbye ends the client's loop
    for _ in range(10):
        client.poll_once(200)
        if client.finished:
            break
    assert client.finished


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


def test_compare_snapshots(tmp_path, capsys):
    """Average relative / average / maximum difference per Array, sorting, the workflow identity check and the command
    line (reference: scripts/compare_snapshots.py:82-146)."""
    from veles_b200.backends import NumpyDevice
    from veles_b200.models import build_workflow
    from veles_b200.scripts.compare_snapshots import compare, format_table, main
    from veles_b200.snapshotter import SnapshotterToFile
    root.common.disable.snapshotting = True
    wf = build_workflow("mnist_mlp", minibatch_size=50, n_train=100, n_valid=0, max_epochs=1)
    wf.initialize(device=NumpyDevice())
    a = pickle.loads(pickle.dumps(wf))
    wf.run()
    b = pickle.loads(pickle.dumps(wf))
    rows = {(u, k): (rel, avg, mx) for _, u, k, rel, avg, mx in compare(a, b)}
    rel, avg, mx = rows[("fc_tanh1", "weights")]
    assert 0 < avg <= mx and 0 < rel < 2.0 and rows[("fc_softmax2", "bias")][2] > 0
    assert all(v == (0.0, 0.0, 0.0) for v in {(u, k): r[0:3] for (u, k), r in rows.items()}.values()) is False
    same = compare(a, a)
    assert same and all(r[3] == 0 and r[5] == 0 for r in same)
    deps = [r[0] for r in compare(a, b, ("dep", "avgreldiff"))]
    assert deps == sorted(deps)
    by_max = [r[5] for r in compare(a, b, ("maxdiff",))]
    assert by_max == sorted(by_max)
    table = format_table(compare(a, b))
    assert "Average Relative Diff" in table and "fc_tanh1" in table and table.count("+--") >= 3
    other = build_workflow("wide_mlp_ae", minibatch_size=8, n_train=16, n_valid=0, max_epochs=1)
    other.initialize(device=NumpyDevice())
    other_copy = pickle.loads(pickle.dumps(other))
    other_copy.__class__ = type("SomeOtherWorkflow", (type(other_copy),), {})
    with pytest.raises(ValueError):
        compare(a, other_copy)
    # command line on snapshot files
    pa, pb = str(tmp_path / "a.pickle"), str(tmp_path / "b.pickle")
    for path, obj in ((pa, a), (pb, b)):
        with open(path, "wb") as f:
            pickle.dump(obj, f)
    assert main(["-q", pa, pb, "-s", "maxdiff", "unit"]) == 0
    assert "fc_softmax2" in capsys.readouterr().out
    assert main([pa]) == 1
    wf.launcher.stop()
    other.launcher.stop()


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


def _report_info(tmp_path):
    plot = tmp_path / "curve.svg"
    plot.write_text("<svg xmlns='http://www.w3.org/2000/svg'></svg>")
    return {"name": "MnistMLP", "description": "two layers (test)", "id": "abc", "mode": "standalone", "device": "numpy",
            "run_time": 1.5, "epochs": 3, "seeds": [1, 2], "results": {"EvaluationFitness": 0.97, "Total epochs": 3},
            "unit_times": [("loader", 0.1), ("fc_tanh1", 0.25)], "config": {"loader": {"minibatch_size": 50}},
            "plots": [str(plot)], "graph": "digraph Workflow { a -> b }"}


def test_pdf_backend_writes_a_wellformed_document(tmp_path):
    import re
    from veles_b200.publishing.backend import BackendRegistry, write_text_pdf
    info = _report_info(tmp_path)
    info["unit_times"] = [("unit%d" % i, i * 0.01) for i in range(150)]       # forces a second page
    backend = BackendRegistry.backends["pdf"](directory=str(tmp_path))
    path = backend.render(info)
    data = open(path, "rb").read()
    assert data.startswith(b"%PDF-1.4") and data.rstrip().endswith(b"%%EOF")
    xref = int(re.search(rb"startxref\n(\d+)", data).group(1))
    assert data[xref:xref + 4] == b"xref"
    n = int(re.search(rb"xref\n0 (\d+)", data).group(1))
    offsets = [int(m) for m in re.findall(rb"(\d{10}) 00000 n", data)]
    assert len(offsets) == n - 1
    for i, off in enumerate(offsets, 1):
        assert data[off:off + len(b"%d 0 obj" % i)] == b"%d 0 obj" % i        # every xref entry points at its object
    assert len(re.findall(rb"/Type /Page ", data)) == getattr(backend, "pages", 2) >= 2
    assert b"EvaluationFitness" in data and b"unit149" in data
    assert write_text_pdf(str(tmp_path / "one.pdf"), "a (b) \\ c") == 1


def test_ipynb_and_latex_backends(tmp_path):
    """Notebook and LaTeX reports (the reference's jinja2 templates ipynb_template.ipynb / latex_template.tex)."""
    from veles_b200.publishing.backend import BackendRegistry
    info = _report_info(tmp_path)
    info["results"]["odd_key & 100%"] = "a_b{c}"
    nb = json.load(open(BackendRegistry.backends["ipynb"](directory=str(tmp_path)).render(info)))
    assert nb["nbformat"] == 4 and nb["cells"][0]["cell_type"] == "markdown"
    assert any("".join(c["source"]).startswith("## Results") for c in nb["cells"])
    code = [c for c in nb["cells"] if c["cell_type"] == "code"]
    scope = {}
    exec("".join(code[0]["source"]).rsplit("\n", 2)[0], scope)      # the data cell runs and reloads the numbers
    assert scope["results"]["odd_key & 100%"] == "a_b{c}" and scope["unit_times"][1] == ["fc_tanh1", 0.25]
    tex = open(BackendRegistry.backends["latex"](directory=str(tmp_path)).render(info)).read()
    assert tex.startswith("\\documentclass") and tex.rstrip().endswith("\\end{document}")
    assert "odd\\_key \\& 100\\%" in tex and "a\\_b\\{c\\}" in tex and "fc\\_tanh1 & 0.2500" in tex
    assert tex.count("\\begin{") == tex.count("\\end{")


def test_confluence_backend_against_a_local_xmlrpc_server(tmp_path):
    """The XML-RPC client and the backend against an in-process server speaking the confluence2 API."""
    import threading
    import xmlrpc.client
    from xmlrpc.server import SimpleXMLRPCServer
    from veles_b200.publishing.backend import BackendRegistry
    from veles_b200.publishing.confluence import Confluence, ConfluenceError

    pages, attachments, calls = {}, {}, []

    class API(object):
        def login(self, user, password):
            if password != "secret":
                raise xmlrpc.client.Fault(1, "bad credentials")
            return "tok-1"

        def getPageSummary(self, token, space, title):
            calls.append(("summary", title))
            if (space, title) not in pages:
                raise xmlrpc.client.Fault(2, "no such page")
            p = pages[(space, title)]
            return {"id": p["id"], "title": title, "space": space}

        def getPage(self, token, space, title):
            if (space, title) not in pages:
                raise xmlrpc.client.Fault(2, "no such page")
            return dict(pages[(space, title)])

        def storePage(self, token, data):
            key = (data["space"], data["title"])
            data = dict(data)
            data.setdefault("id", str(100 + len(pages)))
            data["url"] = "http://wiki/%s/%s" % key
            pages[key] = data
            return data

        def addAttachment(self, token, page_id, meta, blob):
            attachments.setdefault(page_id, {})[meta["fileName"]] = (meta["contentType"], blob.data)
            return meta

        def removeAttachment(self, token, page_id, name):
            if name not in attachments.get(page_id, {}):
                raise xmlrpc.client.Fault(3, "no attachment")
            del attachments[page_id][name]
            return True

        def getAttachments(self, token, page_id):
            return [{"fileName": n} for n in attachments.get(page_id, {})]

        def getSpaces(self, token):
            return [{"key": "ML"}]

    from xmlrpc.server import SimpleXMLRPCRequestHandler

    class Handler(SimpleXMLRPCRequestHandler):
        rpc_paths = ("/rpc/xmlrpc",)

    srv = SimpleXMLRPCServer(("127.0.0.1", 0), requestHandler=Handler, logRequests=False, allow_none=True)
    api = API()
    for name in dir(api):
        if not name.startswith("_"):
            srv.register_function(getattr(api, name), "confluence2." + name)
    port = srv.server_address[1]
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    try:
        url = "http://127.0.0.1:%d" % port
        with pytest.raises(ConfluenceError):
            Confluence(url, "bot", "wrong")
        conf = Confluence(url, "bot", "secret")
        assert conf.token == "tok-1" and conf.get_spaces() == [{"key": "ML"}]
        assert conf.get_page_summary("Reports", "ML") is None
        conf.store_page_content("Reports", "ML", "<p>index</p>", convert_wiki=False)
        kw = dict(server=url, username="bot", password="secret", space="ML", parent="Reports")
        info = _report_info(tmp_path)
        first = BackendRegistry.backends["confluence"](**kw)
        assert first.render(info) == "http://wiki/ML/MnistMLP" and first.page == "MnistMLP"
        second = BackendRegistry.backends["confluence"](**kw)
        second.render(info)
        assert second.page == "MnistMLP (1)"                       # an existing title is not overwritten
        stored = pages[("ML", "MnistMLP")]
        assert stored["parentId"] == pages[("ML", "Reports")]["id"]
        assert "EvaluationFitness" in stored["content"] and 'ri:filename="curve.svg"' in stored["content"]
        att = attachments[stored["id"]]
        assert set(att) == {"curve.svg", "workflow.dot"} and att["curve.svg"][0] == "image/svg+xml"
        assert att["workflow.dot"][1].startswith(b"digraph")
        conf.attach_file("MnistMLP", "ML", {"curve.svg": b"<svg/>"})             # replaces
        assert attachments[stored["id"]]["curve.svg"][1] == b"<svg/>"
        assert conf.remove_all_attachments("MnistMLP", "ML") == 2 and not attachments[stored["id"]]
        with pytest.raises(ConfluenceError):
            conf.store_page_content("X", "ML", "<p/>", convert_wiki=False, parent="missing")
        with pytest.raises(ValueError):
            BackendRegistry.backends["confluence"](server=url)
    finally:
        srv.shutdown()
        srv.server_close()


def test_update_forge_uploads_every_model_directory(tmp_path, capsys):
    """scripts/update_forge: walk a tree, upload every directory with a manifest (reference: scripts/update_forge.py)."""
    from veles_b200.forge.forge_client import ForgeClient
    from veles_b200.forge.forge_server import ForgeServer
    from veles_b200.scripts import update_forge
    for name in ("alpha", "nested/beta"):
        d = tmp_path / "models" / name
        d.mkdir(parents=True)
        (d / "wf.py").write_text("def run(load, main): pass\n")
        (d / "cfg.py").write_text("# config\n")
        (d / "manifest.json").write_text(json.dumps({
            "name": os.path.basename(name), "workflow": "wf.py", "configuration": "cfg.py",
            "short_description": name, "long_description": name, "author": "tester"}))
    (tmp_path / "models" / "not_a_model").mkdir()
    srv = ForgeServer(str(tmp_path / "store")).start()
    try:
        c = ForgeClient(srv.url)
        token = c.register("me@example.com")
        assert update_forge.main([]) == 1                       # usage
        token = token if isinstance(token, str) else getattr(c, "token", None)
        assert update_forge.main([str(tmp_path / "models"), srv.url, token]) == 0
        assert sorted(row[0] for row in c.list()) == ["alpha", "beta"]
        out = capsys.readouterr().out
        assert "alpha" in out and "beta" in out and "not_a_model" not in out
    finally:
        srv.stop()


def test_standard_workflow_service_links(tmp_path):
    """Every documented ``link_*`` of StandardWorkflow (manualrst_veles_workflow_creation.rst:454-472, 550-644): the
    plotters hang behind Decision, run at epoch ends and write their SVG files; ImageSaver dumps samples."""
    from veles_b200.backends import NumpyDevice
    from veles_b200.dummy import DummyLauncher
    from veles_b200.models.configs import MODEL_CONFIGS
    from veles_b200.loader import synthetic  # noqa: F401
    from veles_b200.znicz.standard_workflow import StandardWorkflow
    root.common.disable.snapshotting = True
    root.common.disable.plotting = False
    plots = str(tmp_path / "plots")

    class Wf(StandardWorkflow):
        def create_workflow(self):
            self.link_repeater(self.start_point)
            self.link_loader(self.repeater)
            self.link_forwards(("input", "minibatch_data"), self.loader)
            self.link_evaluator(self.forwards[-1])
            self.link_decision(self.evaluator)
            self.link_snapshotter(self.decision)
            last = self.link_image_saver(self.snapshotter)
            self.link_gds(last)
            self.link_loop(self.gds[0])
            last = self.link_error_plotter(last)      # a side branch behind ImageSaver: also runs on the last epoch
            last = self.link_conf_matrix_plotter(last)
            last = self.link_err_y_plotter(last)
            last = self.link_multi_hist_plotter(None, 4, last)
            last = self.link_weights_plotter(None, 16, "weights", last)
            last = self.link_similar_weights_plotter(None, 16, 0.2, 1.1, 0.5, last)
            last = self.link_image_plotter(last)
            last = self.link_table_plotter(last)
            self.link_end_point(last)

    try:
        cfg = MODEL_CONFIGS["mnist_mlp"]
        wf = Wf(DummyLauncher(), layers=cfg["layers"](), loader_name=cfg["loader_name"],
                loader_config={"minibatch_size": 50, "n_train": 200, "n_valid": 100, "sample_shape": (784,),
                               "n_classes": 10},
                loss_function="softmax", decision_config={"max_epochs": 2},
                image_saver_config={"out_dirs": [str(tmp_path / "img" / n) for n in ("test", "validation", "train")],
                                    "limit": 5, "only_errors": False})
        wf.plotters_are_enabled = True
        for p in wf.plotters:
            p.output_dir = plots
            p.redraw_threshold = 0
        wf.initialize(device=NumpyDevice())
        wf.run()
        wf.thread_pool.wait_idle(10)
        names = sorted(os.listdir(plots))
        assert len(wf.plotters) >= 2 + 2 + 2 + 2 + 2 + 2 + 1 + 1
        drawn = [p for p in wf.plotters if getattr(p, "last_file", None)]
        assert len(drawn) == len(wf.plotters), [p.name for p in wf.plotters if not getattr(p, "last_file", None)]
        for p in drawn:
            assert open(p.last_file).read().startswith("<svg"), p.name
        assert any("Confusion" in n for n in names) and any("weights" in n for n in names) and "Max__Min.svg" in names
        assert wf.image_saver.saved == 5
        assert sum(len(os.listdir(str(tmp_path / "img" / n))) for n in ("test", "validation", "train")) == 5
        wf.launcher.stop()
    finally:
        root.common.disable.plotting = True


def test_mse_workflow_plotter_links(tmp_path):
    from veles_b200.backends import NumpyDevice
    from veles_b200.dummy import DummyLauncher
    from veles_b200.loader import synthetic  # noqa: F401
    from veles_b200.znicz.standard_workflow import StandardWorkflow
    root.common.disable.snapshotting = True
    root.common.disable.plotting = False

    class Wf(StandardWorkflow):
        def create_workflow(self):
            self.link_repeater(self.start_point)
            self.link_loader(self.repeater)
            self.link_forwards(("input", "minibatch_data"), self.loader)
            self.link_evaluator(self.forwards[-1])
            self.link_decision(self.evaluator)
            self.link_snapshotter(self.decision)
            self.link_gds(self.snapshotter)
            self.link_loop(self.gds[0])
            last = self.link_mse_plotter(self.snapshotter)
            last = self.link_min_max_plotter(True, last)
            last = self.link_min_max_plotter(False, last)
            last = self.link_immediate_plotter(last)
            self.link_end_point(last)

    try:
        layers = [{"type": "all2all_tanh", "->": {"output_sample_shape": 16}, "<-": {"learning_rate": 0.01}},
                  {"type": "all2all", "->": {"output_sample_shape": 32}, "<-": {"learning_rate": 0.01}}]
        wf = Wf(DummyLauncher(), layers=layers, loader_name="synthetic_mse",
                loader_config={"minibatch_size": 20, "n_train": 80, "n_valid": 40, "sample_shape": (32,)},
                loss_function="mse", decision_config={"max_epochs": 2})
        wf.plotters_are_enabled = True
        for p in wf.plotters:
            p.output_dir = str(tmp_path)
            p.redraw_threshold = 0
        wf.initialize(device=NumpyDevice())
        wf.run()
        wf.thread_pool.wait_idle(10)
        assert len(wf.plotters) == 7
        for p in wf.plotters:
            assert p.last_file and open(p.last_file).read().startswith("<svg"), p.name
        assert len(wf.mse_plotters[0].values) >= 2 and wf.mse_plotters[0].values[-1] < 1e29
        wf.launcher.stop()
    finally:
        root.common.disable.plotting = True


def test_forge_confirmation_unregister_and_assist(tmp_path):
    """E-mail confirmed registration, ``unregister`` and ``assist`` (reference: forge_server.py:462-520,
    forge_client.py:514-621)."""
    import urllib.error
    from veles_b200.forge.forge_client import ForgeClient
    from veles_b200.forge.forge_server import ForgeServer
    sent = []
    srv = ForgeServer(str(tmp_path / "store"), require_confirmation=True,
                      mailer=lambda email, subject, body: sent.append((email, body))).start()
    try:
        wf_py = tmp_path / "my_model.py"
        wf_py.write_text("from veles_b200.znicz.standard_workflow import StandardWorkflow\n\n\n"
                         "class TinyWorkflow(StandardWorkflow):\n    pass\n\n\ndef run(load, main):\n    pass\n")
        (tmp_path / "my_model_config.py").write_text("# configuration\n")
        c = ForgeClient(srv.url)
        asked = []
        dest, meta = c.assist(str(wf_py), {"directory": str(tmp_path / "pkg"), "author": "tester"},
                              ask=lambda prompt: asked.append(prompt) or "")
        assert meta["name"] == "TinyWorkflow" and meta["configuration"] == "my_model_config.py"
        assert len(asked) == 3                       # name, short and long description; the author was given
        assert sorted(os.listdir(dest)) == ["manifest.json", "my_model.py", "my_model_config.py"]
        c.register("me@example.com")
        with pytest.raises(urllib.error.HTTPError):  # not confirmed yet
            c.upload(dest)
        assert sent and sent[0][0] == "me@example.com"
        code = sent[0][1].split("Confirmation code: ")[1].split()[0]
        with pytest.raises(urllib.error.HTTPError):
            c.confirm("wrong")
        assert c.confirm(code) is True
        assert c.upload(dest)["version"] == "v1"
        other = ForgeClient(srv.url)
        other.register("other@example.com")
        other.confirm(sent[1][1].split("Confirmation code: ")[1].split()[0])
        with pytest.raises(urllib.error.HTTPError):  # a taken name belongs to its first author
            other.upload(dest)
        token = c.token
        assert c.unregister() == "me@example.com" and c.token is None
        c.token = token
        with pytest.raises(urllib.error.HTTPError):  # the revoked token can no longer write
            c.upload(dest)
        assert [row[0] for row in c.list()] == ["TinyWorkflow"]      # what was published stays
        # the default mailer drops the message into the store's outbox
        srv2 = ForgeServer(str(tmp_path / "store2"), require_confirmation=True)
        srv2.register("x@example.com")
        assert "Confirmation code" in open(os.path.join(str(tmp_path / "store2"), "outbox", "x@example.com.txt")).read()
    finally:
        srv.stop()
