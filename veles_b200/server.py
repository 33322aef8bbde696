"""Master agent of the job-control plane (ZeroMQ ROUTER).

Parity target: ``veles/server.py:62-762`` — handshake with workflow checksum check (:478-529),
job requests answered with ``workflow.generate_data_for_slave`` (:357-399), updates applied
with ``apply_data_from_slave`` (:401-430), at most 2 outstanding jobs per slave (:596-611),
sync-point queue when a unit has no data (:369-395), adaptive per-slave timeout
``max(mean + 3 sigma, --job-timeout)`` then drop + blacklist (:619-635), ``--respawn``
(:637-655), pause/resume (:734-745), slave table for the status page (:219-228).
One socket carries both control and data (the reference split them into Twisted TCP JSON
lines + ZeroMQ); same-host payloads travel through shared memory (``SharedIO``).
"""
from __future__ import annotations

import argparse
import threading
import time
import uuid

import zmq

from .cmdline import CommandLineArgumentsRegistry
from .logger import Logger
from .network_common import dumps, loads, machine_id, parse_address, default_codec, SharedIO
from . import pickle2
import pickle


class SlaveDescription(object):
    def __init__(self, sid, info):
        self.id = sid
        self.mid = info.get("mid")
        self.pid = info.get("pid")
        self.power = info.get("power", 0)
        self.host = info.get("host", "?")
        self.backend = info.get("backend")
        self.device = info.get("device")
        self.argv = info.get("argv")
        self.state = "WAIT"
        self.jobs_in_flight = 0
        self.jobs_done = 0
        self.job_times = []
        self.job_started = []
        self.last_seen = time.time()
        self.paused = False

    def timeout(self, floor):
        if len(self.job_times) < 3:
            return floor
        n = len(self.job_times)
        mean = sum(self.job_times) / n
        var = sum((t - mean) ** 2 for t in self.job_times) / n
        return max(mean + 3 * var ** 0.5, floor)


class Server(Logger, metaclass=CommandLineArgumentsRegistry):
    MAX_JOBS_IN_FLIGHT = 2
    SHM_THRESHOLD = 1 << 16

    def __init__(self, address, workflow, launcher=None, **kwargs):
        super(Server, self).__init__()
        parser = Server.init_parser(parser=argparse.ArgumentParser(add_help=False))
        import sys
        self.args, _ = parser.parse_known_args(kwargs.get("argv", sys.argv[1:]))
        self.job_timeout = kwargs.get("job_timeout", self.args.job_timeout * 60.0)
        self.respawn = kwargs.get("respawn", self.args.respawn)
        self.workflow = workflow
        self.launcher = launcher
        self.address = parse_address(address)
        self.ctx = zmq.Context.instance()
        self.sock = self.ctx.socket(zmq.ROUTER)
        self.sock.setsockopt(zmq.LINGER, 0)
        self.sock.bind(self.address)
        self.endpoint = self.sock.getsockopt_string(zmq.LAST_ENDPOINT)
        self.slaves = {}
        self.blacklist = set()
        self.job_requests = []      # sync-point queue of slave ids waiting for a job
        self.mid = machine_id()
        self.codec = default_codec()
        self._stop = threading.Event()
        self._no_more_jobs = False
        self.on_respawn = kwargs.get("on_respawn")
        self.heartbeat_timeout = kwargs.get("heartbeat_timeout", 10.0)

    @staticmethod
    def init_parser(**kwargs):
        parser = kwargs.get("parser", argparse.ArgumentParser())
        parser.add_argument("--job-timeout", type=float, default=2.0,
                            help="Minimal job timeout in minutes before a slave is considered hung.")
        parser.add_argument("--respawn", default=False, action="store_true",
                            help="Relaunch slaves which died or were dropped.")
        return parser

    # -- wire helpers ---------------------------------------------------------------------------------------
    def _send(self, sid, cmd, obj=None, slave=None):
        frames = [sid, cmd.encode()]
        if obj is not None:
            same_host = slave is not None and slave.mid == self.mid
            if same_host:
                payload = pickle.dumps(obj, protocol=pickle2.best_protocol)
                if len(payload) >= self.SHM_THRESHOLD:
                    name = "veles-job-%s" % sid.hex()[:16]
                    SharedIO.put(name, payload)
                    frames += [b"shm", name.encode()]
                    self.sock.send_multipart(frames)
                    return
            frames += dumps(obj, None if same_host else self.codec)
        self.sock.send_multipart(frames)

    @staticmethod
    def _decode(frames):
        if not frames:
            return None
        if frames[0] == b"shm":
            return pickle.loads(SharedIO.get(frames[1].decode()))
        return loads(frames)

    # -- main loop ----------------------------------------------------------------------------------------------
    def run(self):
        poller = zmq.Poller()
        poller.register(self.sock, zmq.POLLIN)
        self.info("Master is listening on %s", self.endpoint)
        wf = self.workflow
        wf._finished_ = False
        wf._is_running = True
        wf._sync_event_.clear()
        idle_since = None
        while not self._stop.is_set():
            events = dict(poller.poll(100))
            if self.sock in events:
                frames = self.sock.recv_multipart()
                try:
                    self._dispatch(frames)
                except Exception:  # noqa: BLE001 - a bad message must not kill the master
                    self.exception("Failed to process a message from %s", frames[0].hex()[:8])
            self._check_timeouts()
            self._serve_queued()
            if wf._finished_ or self._no_more_jobs:
                busy = any(s.jobs_in_flight for s in self.slaves.values())
                if not busy:
                    if idle_since is None:
                        idle_since = time.time()
                        for sid in list(self.slaves):
                            self._send(sid, "end")
                    if not wf._finished_:
                        wf.on_workflow_finished()
                    if time.time() - idle_since > 0.3:
                        break
        self.sock.close()

    def stop(self):
        self._stop.set()

    def _dispatch(self, frames):
        sid, cmd = frames[0], frames[1].decode()
        slave = self.slaves.get(sid)
        if slave is not None:
            slave.last_seen = time.time()
        if cmd == "handshake":
            return self._handshake(sid, self._decode(frames[2:]))
        if slave is None:
            self._send(sid, "error", {"error": "unknown slave: handshake first"})
            return
        if cmd == "job_request":
            self._job_request(slave)
        elif cmd == "update":
            self._update(slave, self._decode(frames[2:]))
        elif cmd == "power":
            slave.power = self._decode(frames[2:]).get("power", slave.power)
        elif cmd == "ping":
            pass
        elif cmd == "bye":
            self._drop(slave, "disconnected")
        else:
            self.warning("Unknown command %r from %s", cmd, slave.id)

    def _handshake(self, sid, info):
        if info.get("mid") in self.blacklist or (info.get("mid"), info.get("pid")) in self.blacklist:
            self._send(sid, "error", {"error": "blacklisted"})
            return
        if info.get("checksum") != self.workflow.checksum:
            self.warning("Rejected a slave with a different workflow checksum (%s != %s)",
                         info.get("checksum"), self.workflow.checksum)
            self._send(sid, "error", {"error": "workflow checksum mismatch"})
            return
        slave = SlaveDescription(sid.hex()[:8] + "-" + str(uuid.uuid4())[:8], info)
        self.slaves[sid] = slave
        self.workflow.apply_initial_data_from_slave(info.get("data"), slave)
        reply = {"id": slave.id, "log_id": getattr(self.launcher, "log_id", None), "mid": self.mid,
                 "data": self.workflow.generate_initial_data_for_slave(slave)}
        self._send(sid, "handshake_ok", reply, slave)
        self.info("New slave %s (%s, power %.1f)", slave.id, slave.host, slave.power or 0)
        self.event("slave", "single", id=slave.id, what="connected")

    def _sid_of(self, slave):
        for sid, s in self.slaves.items():
            if s is slave:
                return sid
        return None

    def _job_request(self, slave):
        if slave.paused:
            self.job_requests.append(slave)
            return
        if slave.jobs_in_flight >= self.MAX_JOBS_IN_FLIGHT:
            self.job_requests.append(slave)
            return
        self._serve(slave)

    def _serve(self, slave):
        sid = self._sid_of(slave)
        if sid is None:
            return
        if self._no_more_jobs or self.workflow._finished_:
            self._send(sid, "end")
            return
        slave.state = "GETTING_JOB"
        data = self.workflow.generate_data_for_slave(slave)
        if data is False:
            # sync point: some unit is waiting for outstanding updates
            if slave not in self.job_requests:
                self.job_requests.append(slave)
            slave.state = "WAIT"
            self._send(sid, "wait")
            return
        if data is None:
            self._no_more_jobs = True
            self._send(sid, "end")
            return
        slave.jobs_in_flight += 1
        slave.job_started.append(time.time())
        slave.state = "WORK"
        self._send(sid, "job", data, slave)

    def _serve_queued(self):
        if not self.job_requests:
            return
        pending, self.job_requests = self.job_requests, []
        for slave in pending:
            if self._sid_of(slave) is None:
                continue
            if slave.paused or slave.jobs_in_flight >= self.MAX_JOBS_IN_FLIGHT:
                self.job_requests.append(slave)
            else:
                self._serve(slave)

    def _update(self, slave, data):
        sid = self._sid_of(slave)
        ok = True
        try:
            self.workflow.apply_data_from_slave(data, slave)
        except Exception:  # noqa: BLE001
            self.exception("Failed to apply the update from %s", slave.id)
            ok = False
        if slave.job_started:
            slave.job_times.append(time.time() - slave.job_started.pop(0))
            slave.job_times = slave.job_times[-50:]
        slave.jobs_in_flight = max(0, slave.jobs_in_flight - 1)
        slave.jobs_done += 1
        slave.state = "WAIT" if slave.jobs_in_flight == 0 else "WORK"
        self._send(sid, "update_ack", {"ok": ok})

    # -- failure handling -----------------------------------------------------------------------------------------------
    def _check_timeouts(self):
        now = time.time()
        for sid, slave in list(self.slaves.items()):
            hung = slave.job_started and now - slave.job_started[0] > slave.timeout(self.job_timeout)
            dead = now - slave.last_seen > self.heartbeat_timeout and slave.jobs_in_flight > 0
            if hung or dead:
                self.warning("Slave %s %s: dropping it", slave.id, "timed out" if hung else "stopped responding")
                if hung:
                    self.blacklist.add((slave.mid, slave.pid))
                self._drop(slave, "timeout" if hung else "dead")

    def _drop(self, slave, why):
        sid = self._sid_of(slave)
        if sid is None:
            return
        del self.slaves[sid]
        if slave in self.job_requests:
            self.job_requests.remove(slave)
        self.event("drop_slave", "single", id=slave.id, why=why)
        try:
            self.workflow.drop_slave(slave)
        finally:
            if self.respawn and self.on_respawn is not None and why != "disconnected":
                self.on_respawn(slave)

    # -- control ------------------------------------------------------------------------------------------------------------
    def pause(self, slave_id):
        for sid, s in self.slaves.items():
            if s.id == slave_id:
                s.paused = True
                self._send(sid, "pause")

    def resume(self, slave_id):
        for sid, s in self.slaves.items():
            if s.id == slave_id:
                s.paused = False
                self._send(sid, "resume")

    def slaves_status(self):
        return {s.id: {"host": s.host, "power": s.power, "state": s.state, "jobs": s.jobs_done,
                       "backend": s.backend, "device": s.device} for s in self.slaves.values()}
