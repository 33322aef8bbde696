"""Rank supervisor for single-node multi-GPU jobs: spawn, watch, abort together, restart from the last snapshot.

The reference recovers at job granularity: the master drops a dead or hung slave, re-queues its minibatches and may
respawn it over SSH (``veles/server.py:315-338, 383-394, 619-655``; fault injection ``client.py:303-307``).  A
synchronous collective over NVLink cannot lose a participant mid-step, so the unit of recovery here is the whole rank
group:

* one process per GPU is started with the ``torch.distributed`` environment (``RANK``, ``LOCAL_RANK``, ``WORLD_SIZE``,
  ``MASTER_ADDR=127.0.0.1``, a free ``MASTER_PORT``), each in its own session so that it can be signalled as a group;
* **failure detection**: a rank exiting non-zero, or a rank whose heartbeat file (``VELES_HEARTBEAT_FILE``, touched by
  :func:`heartbeat` from the training loop) has not changed for ``heartbeat_timeout`` seconds (``startup_timeout`` before the first beat);
* on failure every surviving rank is terminated (SIGTERM to its process group, SIGKILL after ``grace`` seconds - exact
  PIDs only) and the group is started again, at most ``max_restarts`` times, with ``VELES_RESTART_COUNT`` incremented
  and the newest file matching ``snapshot_glob`` exported as ``VELES_RESUME_SNAPSHOT`` and substituted for
  ``{snapshot}`` in the command (``-w {snapshot}`` resumes a ``python -m veles_b200`` run);
* ``shrink_on_failure`` restarts with one rank less (weak scaling keeps the per-GPU batch), down to ``min_ranks`` -
  the "continue with fewer GPUs" policy;
* ``VELES_RANK_DEATH_PROBABILITY`` is the fault-injection knob (:func:`maybe_die`, the counterpart of
  ``--slave-death-probability``).

CLI: ``python -m veles_b200.parallel.supervisor --nproc 8 [options] -- python train.py ...``
"""
from __future__ import annotations

import argparse
import glob
import os
import random
import signal
import socket
import subprocess
import sys
import tempfile
import time


def heartbeat():
    """Called by the training loop (``Trainer.step``, ``Repeater``): tells the supervisor this rank is alive."""
    path = os.environ.get("VELES_HEARTBEAT_FILE")
    if not path:
        return
    now = time.time()
    if now - heartbeat.last < heartbeat.min_interval:
        return
    heartbeat.last = now
    try:
        with open(path, "w") as f:
            f.write("%f\n" % now)
    except OSError:
        pass


heartbeat.last = 0.0
heartbeat.min_interval = 0.5


def maybe_die(step=None):
    """Fault injection: with probability ``VELES_RANK_DEATH_PROBABILITY`` per call, exit abruptly (code 42).
    ``VELES_RANK_DEATH_STEP="<rank>:<n>"`` is the deterministic variant: that rank dies at its n-th call, in the first
    attempt only (what the tests use)."""
    spec = os.environ.get("VELES_RANK_DEATH_STEP")
    if spec and restart_count() == 0:
        rank, _, nth = spec.partition(":")
        if os.environ.get("RANK", "0") == rank:
            maybe_die.calls += 1
            if maybe_die.calls == int(nth or 1):
                sys.stderr.write("rank %s: injected death at call %d\n" % (rank, maybe_die.calls))
                sys.stderr.flush()
                os._exit(42)
    p = float(os.environ.get("VELES_RANK_DEATH_PROBABILITY", "0") or 0)
    if p > 0 and random.random() < p:
        sys.stderr.write("rank %s: injected death%s\n" % (os.environ.get("RANK", "?"),
                                                         "" if step is None else " at step %s" % step))
        sys.stderr.flush()
        os._exit(42)


maybe_die.calls = 0


def restart_count():
    return int(os.environ.get("VELES_RESTART_COUNT", "0") or 0)


def resume_snapshot():
    """Path of the snapshot the supervisor wants this attempt to resume from (or None)."""
    return os.environ.get("VELES_RESUME_SNAPSHOT") or None


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class Attempt(object):
    """One incarnation of the rank group."""

    def __init__(self, cmd, world, index, env_extra, hb_dir, log_dir=None):
        self.world = world
        self.index = index
        self.procs = []
        self.hb_files = []
        self.started = time.time()
        self.logs = []
        port = free_port()
        for rank in range(world):
            env = dict(os.environ)
            env.update(env_extra)
            hb = os.path.join(hb_dir, "attempt%d_rank%d.hb" % (index, rank))
            env.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world),
                        "LOCAL_WORLD_SIZE": str(world), "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                        "VELES_HEARTBEAT_FILE": hb, "VELES_RESTART_COUNT": str(index)})
            out = None
            if log_dir:
                os.makedirs(log_dir, exist_ok=True)
                out = open(os.path.join(log_dir, "attempt%d_rank%d.log" % (index, rank)), "wb")
                self.logs.append(out)
            self.procs.append(subprocess.Popen(cmd, env=env, start_new_session=True, stdout=out,
                                               stderr=subprocess.STDOUT if out else None))
            self.hb_files.append(hb)

    def poll(self):
        return [p.poll() for p in self.procs]

    def stalled(self, timeout, startup_timeout):
        """Ranks still running whose heartbeat is older than ``timeout`` (before the first beat - imports, process
        group, CUDA context, graph capture - ``startup_timeout`` since the start applies instead)."""
        if not timeout:
            return []
        now = time.time()
        out = []
        for rank, (p, hb) in enumerate(zip(self.procs, self.hb_files)):
            if p.poll() is not None:
                continue
            try:
                late = now - os.path.getmtime(hb) > timeout
            except OSError:
                late = now - self.started > startup_timeout
            if late:
                out.append(rank)
        return out

    def terminate(self, grace):
        """Stop every rank that is still running: SIGTERM to its own process group, SIGKILL after ``grace`` s."""
        alive = [p for p in self.procs if p.poll() is None]
        for p in alive:
            try:
                os.killpg(p.pid, signal.SIGTERM)      # the rank is its own session / group leader
            except (ProcessLookupError, PermissionError):
                pass
        deadline = time.time() + grace
        for p in alive:
            try:
                p.wait(max(0.0, deadline - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except (ProcessLookupError, PermissionError):
                    pass
                p.wait()
        for f in self.logs:
            f.close()


class Supervisor(object):
    def __init__(self, cmd, nproc, max_restarts=3, heartbeat_timeout=0.0, grace=5.0, snapshot_glob=None,
                 shrink_on_failure=False, min_ranks=1, poll_interval=0.2, env=None, log_dir=None, log=None,
                 startup_timeout=None):
        self.cmd = list(cmd)
        self.nproc = int(nproc)
        self.max_restarts = int(max_restarts)
        self.heartbeat_timeout = float(heartbeat_timeout or 0.0)
        self.startup_timeout = float(startup_timeout) if startup_timeout else max(300.0, self.heartbeat_timeout)
        self.grace = float(grace)
        self.snapshot_glob = snapshot_glob
        self.shrink_on_failure = bool(shrink_on_failure)
        self.min_ranks = max(1, int(min_ranks))
        self.poll_interval = float(poll_interval)
        self.env = dict(env or {})
        self.log_dir = log_dir
        self.log = log or (lambda msg: sys.stderr.write("[supervisor] %s\n" % msg))
        self.history = []          # one dict per attempt: world, outcome, failed_ranks, snapshot, seconds

    def newest_snapshot(self):
        if not self.snapshot_glob:
            return None
        files = [f for f in glob.glob(self.snapshot_glob) if os.path.isfile(f)]
        return max(files, key=os.path.getmtime) if files else None

    def command_for(self, snapshot):
        if snapshot is None:
            # drop "-w {snapshot}" style pairs when there is nothing to resume from
            out, skip = [], False
            for i, a in enumerate(self.cmd):
                if skip:
                    skip = False
                    continue
                if "{snapshot}" in a:
                    if out and out[-1] in ("-w", "--snapshot"):
                        out.pop()
                    continue
                out.append(a)
            return out
        return [a.replace("{snapshot}", snapshot) for a in self.cmd]

    def run(self):
        world = self.nproc
        hb_dir = tempfile.mkdtemp(prefix="veles_b200_hb_")
        attempt_index = 0
        while True:
            snapshot = self.newest_snapshot() if attempt_index else None
            env = dict(self.env)
            if snapshot:
                env["VELES_RESUME_SNAPSHOT"] = snapshot
            self.log("attempt %d: %d rank(s)%s" % (attempt_index, world,
                                                   ", resuming from %s" % snapshot if snapshot else ""))
            att = Attempt(self.command_for(snapshot), world, attempt_index, env, hb_dir, self.log_dir)
            outcome, failed = self.watch(att)
            att.terminate(self.grace)
            self.history.append({"attempt": attempt_index, "world": world, "outcome": outcome,
                                 "failed_ranks": failed, "snapshot": snapshot,
                                 "seconds": time.time() - att.started})
            if outcome == "ok":
                return 0
            self.log("attempt %d failed (%s, ranks %s)" % (attempt_index, outcome, failed))
            if attempt_index >= self.max_restarts:
                self.log("giving up after %d restart(s)" % attempt_index)
                return 1
            attempt_index += 1
            if self.shrink_on_failure and world > self.min_ranks:
                world -= 1

    def watch(self, att):
        while True:
            codes = att.poll()
            bad = [r for r, c in enumerate(codes) if c not in (None, 0)]
            if bad:
                return "exit %s" % ",".join(str(codes[r]) for r in bad), bad
            if all(c == 0 for c in codes):
                return "ok", []
            hung = att.stalled(self.heartbeat_timeout, self.startup_timeout)
            if hung:
                return "no heartbeat for %.0f s" % self.heartbeat_timeout, hung
            time.sleep(self.poll_interval)


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m veles_b200.parallel.supervisor", description=__doc__.split("\n")[0])
    ap.add_argument("--nproc", type=int, required=True, help="ranks (= GPUs) to start")
    ap.add_argument("--max-restarts", type=int, default=3)
    ap.add_argument("--heartbeat-timeout", type=float, default=0.0, help="seconds without a heartbeat = hung (0: off)")
    ap.add_argument("--startup-timeout", type=float, default=0.0,
                    help="seconds a rank may take to its first heartbeat (default: max(300, heartbeat timeout))")
    ap.add_argument("--grace", type=float, default=5.0, help="seconds between SIGTERM and SIGKILL")
    ap.add_argument("--snapshot-glob", default=None, help="newest match is passed to a restarted group")
    ap.add_argument("--shrink-on-failure", action="store_true", help="restart with one rank less after a failure")
    ap.add_argument("--min-ranks", type=int, default=1)
    ap.add_argument("--log-dir", default=None, help="per-rank stdout/stderr files instead of inheriting the terminal")
    ap.add_argument("cmd", nargs=argparse.REMAINDER, help="-- command to run as every rank")
    args = ap.parse_args(argv)
    cmd = args.cmd[1:] if args.cmd and args.cmd[0] == "--" else args.cmd
    if not cmd:
        ap.error("no command given")
    sup = Supervisor(cmd, args.nproc, args.max_restarts, args.heartbeat_timeout, args.grace, args.snapshot_glob,
                     args.shrink_on_failure, args.min_ranks, log_dir=args.log_dir,
                     startup_timeout=args.startup_timeout or None)
    return sup.run()


if __name__ == "__main__":
    sys.exit(main())
