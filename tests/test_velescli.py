"""CLI: run a sample workflow, result file, snapshot + resume, dry runs, seeding, genetics, ensembles
(pattern of veles/tests/test_velescli.py:52-94)."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SAMPLES = os.path.join(ROOT, "veles_b200", "znicz", "samples")
ENV = dict(os.environ, PYTHONPATH=ROOT)


def cli(*args, timeout=300):
    cmd = [sys.executable, "-m", "veles_b200", "--no-logo", "-v", "warning", "-a", "numpy"] + list(args)
    return subprocess.run(cmd, cwd=ROOT, env=ENV, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=timeout)


def test_run_sample_with_result_file_snapshot_and_resume(tmp_path):
    res = tmp_path / "res.json"
    snaps = tmp_path / "snaps"
    r = cli("--result-file", str(res), "-r", "1234", os.path.join(SAMPLES, "mnist.py"), "-",
            "root.mnist.decision.max_epochs=2", 'root.common.dirs.snapshots="%s"' % snaps)
    assert r.returncode == 0, r.stdout.decode()[-2000:]
    data = json.loads(res.read_text())
    assert 0.8 < data["EvaluationFitness"] <= 1.0
    assert data["Total epochs"] >= 2 and "Snapshot" in data and len(data["Labels"]) == 10
    current = glob.glob(str(snaps / "mnist_current*"))
    assert current and os.path.islink(current[0])
    # resume: the restored workflow continues from epoch 2 and stops at max_epochs=3
    res2 = tmp_path / "res2.json"
    r = cli("--result-file", str(res2), "-w", current[0], os.path.join(SAMPLES, "mnist.py"), "-",
            "root.mnist.decision.max_epochs=3", 'root.common.dirs.snapshots="%s"' % snaps)
    assert r.returncode == 0, r.stdout.decode()[-2000:]
    d2 = json.loads(res2.read_text())
    assert d2["Total epochs"] >= data["Total epochs"]


def test_dry_run_and_graph_and_config_dump(tmp_path):
    graph = tmp_path / "wf.dot"
    r = cli("--dry-run", "init", "--workflow-graph", str(graph), os.path.join(SAMPLES, "mnist.py"), "-")
    assert r.returncode == 0, r.stdout.decode()[-2000:]
    text = graph.read_text()
    assert "digraph" in text and "All2AllTanh" in text and "GDSoftmax" in text
    pic = tmp_path / "wf.svg"
    r = cli("--dry-run", "init", "--workflow-graph", str(pic), os.path.join(SAMPLES, "mnist.py"), "-")
    assert r.returncode == 0 and pic.read_text().startswith("<svg") and "All2AllTanh" in pic.read_text()
    r = cli("--dump-config", os.path.join(SAMPLES, "mnist.py"), "-")
    assert r.returncode == 0 and b"loader_name" in r.stdout
    r = cli("--version")
    assert r.returncode == 0 and b"veles_b200" in r.stdout


def test_cli_run_leaves_logs_and_events_the_web_status_can_query(tmp_path):
    """``root.common.engine.log_sinks=True``: the run appends its log records and timeline events as JSON lines under
    ``dirs.logs``; a status server pointed at that directory answers queries about the session."""
    from veles_b200.web_status import WebServer
    logs = tmp_path / "logs"
    r = cli("-v", "info", os.path.join(SAMPLES, "mnist.py"), "-", "root.mnist.decision.max_epochs=1",
            "root.mnist.loader.n_train=200", "root.mnist.loader.n_valid=100", "root.common.disable.snapshotting=True",
            "root.common.engine.log_sinks=True", 'root.common.dirs.logs="%s"' % logs)
    assert r.returncode == 0, r.stdout.decode()[-2000:]
    files = sorted(os.listdir(logs))
    assert len(files) == 2 and files[0].endswith(".events.jsonl") and files[1].endswith(".logs.jsonl")
    srv = WebServer(port=0, log_dir=str(logs))
    session = files[1].rsplit("-", 1)[0]                 # <session>-<pid>.logs.jsonl
    assert session in srv.service({"request": "sessions"})["sessions"]     # (this process' own records may be there too)
    found = srv.service({"request": "logs", "session": session, "text": "samples:", "level": "info"})["logs"]
    assert found and found[0]["name"].endswith("Loader")
    ev = srv.service({"request": "events", "session": session, "name": "run"})["events"]
    assert {e["type"] for e in ev} >= {"begin", "end"}


def test_same_seed_same_result(tmp_path):
    outs = []
    for i in range(2):
        res = tmp_path / ("r%d.json" % i)
        r = cli("--result-file", str(res), "-r", "00ff00ff", os.path.join(SAMPLES, "mnist.py"), "-",
                "root.mnist.decision.max_epochs=1", "root.mnist.loader.n_train=500",
                "root.common.disable.snapshotting=True")
        assert r.returncode == 0, r.stdout.decode()[-2000:]
        outs.append(json.loads(res.read_text())["Errors"])
    assert outs[0] == outs[1]


def test_callable_module_runs_in_process(tmp_path):
    code = ("import veles_b200, json; l = veles_b200(%r, '-', 'root.mnist.decision.max_epochs=1', "
            "'root.mnist.loader.n_train=300', 'root.common.disable.snapshotting=True', backend='numpy', "
            "verbosity='error', no_logo=True); print('FIT', l.workflow.gather_results()['EvaluationFitness'])"
            % os.path.join(SAMPLES, "mnist.py"))
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=ENV, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, timeout=300)
    assert r.returncode == 0 and b"FIT" in r.stdout, r.stdout.decode()[-2000:]


def test_genetic_optimization_of_ranges(tmp_path):
    res = tmp_path / "gen.json"
    r = cli("--optimize", "3:2", "--result-file", str(res), os.path.join(SAMPLES, "mnist.py"), "-",
            "root.mnist.decision.max_epochs=1", "root.mnist.loader.n_train=400", "root.mnist.loader.n_valid=200",
            "root.common.disable.snapshotting=True", timeout=600)
    assert r.returncode == 0, r.stdout.decode()[-3000:]
    data = json.loads(res.read_text())
    assert data["Generations"] == 2 and data["EvaluationFitness"] > 0.1
    assert any("learning_rate" in k for k in data["BestGenes"])
    best = os.path.join(SAMPLES, "mnist_best_config.py")
    assert os.path.exists(best)
    os.remove(best)


def test_genetic_example_sample(tmp_path):
    """``veles_b200/samples/GeneticExample`` (reference: veles/samples/GeneticExample): a one-unit workflow whose
    fitness the optimizer raises by evolving two configuration values inside their ranges."""
    sample = os.path.join(ROOT, "veles_b200", "samples", "GeneticExample")
    manifest = json.load(open(os.path.join(sample, "manifest.json")))
    res = tmp_path / "gen.json"
    r = cli("--optimize", "6:3", "--result-file", str(res), os.path.join(sample, manifest["workflow"]), "-", timeout=600)
    assert r.returncode == 0, r.stdout.decode()[-3000:]
    data = json.loads(res.read_text())
    assert data["Generations"] == 3 and set(data["BestGenes"]) == {"root.test.x", "root.test.y"}
    x, y = data["BestGenes"]["root.test.x"], data["BestGenes"]["root.test.y"]
    assert -5.0 <= x <= 5.0 and -5.0 <= y <= 5.0
    assert abs(data["EvaluationFitness"] + ((x * x + y - 11.0) ** 2 + (x + y * y - 7.0) ** 2)) < 1e-6
    assert data["EvaluationFitness"] > -170.0          # better than the default genes (0, 0)
    best = os.path.join(sample, "genetics_best_config.py")
    assert os.path.exists(best)
    os.remove(best)


def test_ensemble_train_then_test(tmp_path):
    res = tmp_path / "ens.json"
    snaps = tmp_path / "snaps"
    r = cli("--ensemble-train", "2:0.8", "--result-file", str(res), os.path.join(SAMPLES, "mnist.py"), "-",
            "root.mnist.decision.max_epochs=1", "root.mnist.loader.n_train=400", "root.mnist.loader.n_valid=200",
            'root.common.dirs.snapshots="%s"' % snaps, timeout=600)
    assert r.returncode == 0, r.stdout.decode()[-3000:]
    data = json.loads(res.read_text())
    assert len(data["models"]) == 2 and all("EvaluationFitness" in m and "Snapshot" in m for m in data["models"])
    res2 = tmp_path / "ens_test.json"
    r = cli("--ensemble-test", str(res), "--result-file", str(res2), os.path.join(SAMPLES, "mnist.py"), "-",
            "root.mnist.loader.n_train=400", "root.mnist.loader.n_valid=200", timeout=600)
    assert r.returncode == 0, r.stdout.decode()[-3000:]
    d2 = json.loads(res2.read_text())
    assert len(d2["models"]) == 2 and all("EvaluationFitness" in m for m in d2["models"])
    # the members trained under this command line's overrides (one epoch => the snapshot is of a FINISHED run, which
    # the evaluation above still replays) and wrote distinct files
    names = sorted(os.path.basename(m["Snapshot"]) for m in data["models"])
    assert names[0].startswith("0000_") and names[1].startswith("0001_")
    assert all(m["Best epoch"] <= 1 for m in data["models"])
    # resuming a finished run without new limits returns at once instead of waiting on a shut gate
    r = cli("-w", data["models"][0]["Snapshot"], os.path.join(SAMPLES, "mnist.py"), "-",
            "root.mnist.loader.n_train=400", "root.mnist.loader.n_valid=200", timeout=120)
    assert r.returncode == 0 and b"already complete" in r.stdout, r.stdout.decode()[-2000:]


import pytest as _pytest


@_pytest.mark.parametrize("sample", ["mnist", "cifar", "mnist_ae", "alexnet", "sequences"])
def test_every_sample_loads_its_config_and_initialises(sample):
    """`python -m veles_b200 --dry-run init samples/<x>.py -` for every bundled sample (configs with Range genes too)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    wf = os.path.join(root, "veles_b200", "znicz", "samples", sample + ".py")
    r = subprocess.run([sys.executable, "-m", "veles_b200", "-a", "numpy", "--dry-run", "init", wf, "-"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=300, cwd=root,
                       env=dict(os.environ, VELES_ALLOW_ROOT="1"))
    assert r.returncode == 0, r.stdout.decode(errors="replace")[-2000:]


def test_range_arithmetic_keeps_a_tunable_interval():
    from veles_b200.genetics import Range
    r = 2 * Range(0.01, 0.0005, 0.05)
    assert isinstance(r, Range) and abs(r.default - 0.02) < 1e-12 and abs(r.max_value - 0.1) < 1e-12
    assert abs(float(Range(1.0, 0.0, 2.0) / 4) - 0.25) < 1e-12
    n = -Range(1.0, 0.5, 2.0)
    assert n.min_value == -2.0 and n.max_value == -0.5


def test_data_parallel_cli_under_the_rank_supervisor(tmp_path):
    """`python -m veles_b200 ...` started once per rank (WORLD_SIZE in the environment) trains replicas that exchange
    gradients (host mode over gloo on the numpy backend) and walks the same global minibatches as one process:
    same seed -> the same result as the single-process run, up to summation order."""
    from veles_b200.parallel.supervisor import Supervisor
    common = ["-r", "1234", os.path.join(SAMPLES, "mnist.py"), "-", "root.mnist.decision.max_epochs=2",
              "root.mnist.loader.n_train=600", "root.mnist.loader.n_valid=200",
              "root.common.disable.snapshotting=True"]
    single = tmp_path / "single.json"
    r = cli("--result-file", str(single), *common)
    assert r.returncode == 0, r.stdout.decode()[-2000:]
    multi = tmp_path / "multi.json"
    cmd = [sys.executable, "-m", "veles_b200", "--no-logo", "-v", "warning", "-a", "numpy",
           "--result-file", str(multi)] + common
    logs = []
    sup = Supervisor(cmd, 2, max_restarts=0, env={"PYTHONPATH": ROOT, "OMP_NUM_THREADS": "2"}, log=logs.append,
                     log_dir=str(tmp_path / "logs"))
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        rc = sup.run()
    finally:
        os.chdir(cwd)
    rank_logs = "".join(open(os.path.join(str(tmp_path / "logs"), f), errors="replace").read()[-1500:]
                        for f in sorted(os.listdir(str(tmp_path / "logs"))))
    assert rc == 0, rank_logs
    a, b = json.loads(single.read_text()), json.loads(multi.read_text())
    assert a["Total epochs"] == b["Total epochs"]
    assert abs(a["EvaluationFitness"] - b["EvaluationFitness"]) < 0.02, (a["EvaluationFitness"], b["EvaluationFitness"])


def test_supervised_cli_run_survives_a_rank_death(tmp_path):
    """Two data-parallel CLI ranks (numpy backend); rank 1 is killed in the second epoch; the supervisor restarts the
    group from the newest snapshot rank 0 wrote and the run completes all epochs."""
    from veles_b200.parallel.supervisor import Supervisor
    snaps = tmp_path / "snaps"
    res = tmp_path / "res.json"
    cmd = [sys.executable, "-m", "veles_b200", "--no-logo", "-v", "warning", "-a", "numpy", "--result-file", str(res),
           "-r", "1234", "-w", "{snapshot}", os.path.join(SAMPLES, "mnist.py"), "-",
           "root.mnist.decision.max_epochs=3", "root.mnist.loader.n_train=600", "root.mnist.loader.n_valid=200",
           'root.common.dirs.snapshots="%s"' % snaps]
    logs = []
    # 600 train + 200 validation samples in minibatches of 100: 8 loop iterations per epoch -> call 13 is in epoch 2
    sup = Supervisor(cmd, 2, max_restarts=2, snapshot_glob=str(snaps / "mnist_*.pickle.gz"), log=logs.append,
                     log_dir=str(tmp_path / "logs"),
                     env={"PYTHONPATH": ROOT, "OMP_NUM_THREADS": "2", "VELES_RANK_DEATH_STEP": "1:13"})
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        rc = sup.run()
    finally:
        os.chdir(cwd)
    rank_logs = "".join(open(os.path.join(str(tmp_path / "logs"), f), errors="replace").read()[-1200:]
                        for f in sorted(os.listdir(str(tmp_path / "logs"))))
    assert rc == 0, rank_logs + repr(sup.history)
    assert [h["outcome"] for h in sup.history] == ["exit 42", "ok"], sup.history
    assert sup.history[1]["snapshot"] and "mnist_" in sup.history[1]["snapshot"]
    data = json.loads(res.read_text())
    assert data["Total epochs"] >= 3 and data["EvaluationFitness"] > 0.8
