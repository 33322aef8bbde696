"""Rank supervisor (veles_b200.parallel.supervisor): crash / hang detection, group restart from the newest snapshot,
shrinking the rank group - exercised with gloo ranks on the CPU (reference behaviour: slave drop / respawn /
death-probability in veles/server.py:315-338, 619-655 and client.py:303-307)."""
import json
import os
import sys

import pytest

from veles_b200.parallel.supervisor import Supervisor, main as supervisor_main

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOY = os.path.join(ROOT, "tests", "supervised_toy.py")


def _sup(tmp_path, mode, nproc=2, steps=6, **kw):
    cmd = [sys.executable, TOY, str(tmp_path), str(steps), mode]
    logs = []
    sup = Supervisor(cmd, nproc, snapshot_glob=str(tmp_path / "snap_*.json"), grace=2.0, log=logs.append,
                     log_dir=str(tmp_path / "logs"), env={"OMP_NUM_THREADS": "1"}, **kw)
    return sup, logs


def test_clean_run_needs_no_restart(tmp_path):
    sup, logs = _sup(tmp_path, "ok")
    assert sup.run() == 0
    assert len(sup.history) == 1 and sup.history[0]["outcome"] == "ok"
    done = json.load(open(tmp_path / "done.json"))
    assert done["total"] == sum(range(1, 7)) and done["attempt"] == 0 and done["resumed_from"] is None


def test_crashed_rank_restarts_the_group_from_the_newest_snapshot(tmp_path):
    sup, logs = _sup(tmp_path, "crash")
    assert sup.run() == 0
    assert [h["outcome"] for h in sup.history] == ["exit 7", "ok"] and sup.history[0]["failed_ranks"] == [1]
    done = json.load(open(tmp_path / "done.json"))
    # the second attempt resumed from the last snapshot rank 0 wrote around the crash and finished the sum
    assert done["attempt"] == 1 and os.path.basename(done["resumed_from"]) in ("snap_003.json", "snap_004.json")
    assert done["total"] == sum(range(1, 7))
    assert any("resuming from" in m for m in logs)


def test_hung_rank_is_detected_by_its_heartbeat(tmp_path):
    sup, logs = _sup(tmp_path, "hang", heartbeat_timeout=3.0, startup_timeout=120.0)
    assert sup.run() == 0
    assert sup.history[0]["outcome"].startswith("no heartbeat") and sup.history[1]["outcome"] == "ok"
    assert json.load(open(tmp_path / "done.json"))["total"] == sum(range(1, 7))


def test_shrinking_and_giving_up(tmp_path):
    sup, logs = _sup(tmp_path, "flaky", nproc=3, shrink_on_failure=True, max_restarts=4)
    assert sup.run() == 0
    assert [h["world"] for h in sup.history] == [3, 2, 1] and sup.history[-1]["outcome"] == "ok"
    assert json.load(open(tmp_path / "done.json"))["world"] == 1
    other = tmp_path / "second"
    other.mkdir()
    sup, logs = _sup(other, "flaky", nproc=2, max_restarts=1)
    assert sup.run() == 1 and len(sup.history) == 2 and any("giving up" in m for m in logs)


def test_cli_and_snapshot_placeholder(tmp_path):
    sup = Supervisor(["prog", "-w", "{snapshot}", "x"], 1)
    assert sup.command_for(None) == ["prog", "x"]
    assert sup.command_for("/s.pickle") == ["prog", "-w", "/s.pickle", "x"]
    rc = supervisor_main(["--nproc", "2", "--max-restarts", "0", "--", sys.executable, "-c",
                          "import os,sys; sys.exit(0 if os.environ['WORLD_SIZE']=='2' and 'MASTER_PORT' in os.environ else 3)"])
    assert rc == 0
    with pytest.raises(SystemExit):
        supervisor_main(["--nproc", "2"])
