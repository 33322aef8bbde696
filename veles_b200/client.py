"""Slave agent of the job-control plane (ZeroMQ DEALER).

Parity target: ``veles/client.py:63-517``: handshake (checksum, power, mid, pid, argv),
FSM INIT -> WAIT -> GETTING_JOB -> BUSY -> WAIT | END (:177-195), ``workflow.do_job``
(:278-318), update upload (:333-342), ``--async-slave`` (request the next job while the update
is in flight, :293-295), ``--slave-death-probability`` fault injection (:303-307), reconnect
attempts (:488-507), periodic power reports (:309-313).
"""
from __future__ import annotations

import argparse
import os
import pickle
import random
import socket
import sys
import threading
import time

import zmq

from . import pickle2
from .cmdline import CommandLineArgumentsRegistry
from .error import Bug, MasterSlaveCommunicationError
from .logger import Logger
from .network_common import dumps, loads, machine_id, parse_address, default_codec, SharedIO


class Client(Logger, metaclass=CommandLineArgumentsRegistry):
    SHM_THRESHOLD = 1 << 16

    def __init__(self, address, workflow, launcher=None, **kwargs):
        super(Client, self).__init__()
        parser = Client.init_parser(parser=argparse.ArgumentParser(add_help=False))
        self.args, _ = parser.parse_known_args(kwargs.get("argv", sys.argv[1:]))
        self.async_slave = kwargs.get("async_slave", self.args.async_slave)
        self.death_probability = kwargs.get("death_probability", self.args.slave_death_probability)
        self.workflow = workflow
        self.launcher = launcher
        self.address = parse_address(address)
        self.ctx = zmq.Context.instance()
        self.sock = None
        self.id = None
        self.master_mid = None
        self.mid = machine_id()
        self.codec = default_codec()
        self.state = "INIT"
        self._stop = threading.Event()
        self.jobs_done = 0
        self.reconnect_attempts = kwargs.get("reconnect_attempts", 60)
        self.handshake_timeout = kwargs.get("handshake_timeout", 1.0)
        self.timings = {"job": 0.0, "update": 0.0, "wait": 0.0}
        self._job_result = None
        self._job_event = threading.Event()

    @staticmethod
    def init_parser(**kwargs):
        parser = kwargs.get("parser", argparse.ArgumentParser())
        parser.add_argument("--async-slave", default=False, action="store_true",
                            help="Request the next job before the previous update is acknowledged.")
        parser.add_argument("--slave-death-probability", type=float, default=0.0,
                            help="Fault injection: probability of dying on each received job.")
        return parser

    # -- wire -------------------------------------------------------------------------------------------------------
    def _connect(self):
        self.sock = self.ctx.socket(zmq.DEALER)
        self.sock.setsockopt(zmq.LINGER, 0)
        self.sock.connect(self.address)

    def _send(self, cmd, obj=None):
        frames = [cmd.encode()]
        if obj is not None:
            same_host = self.master_mid == self.mid
            if same_host:
                payload = pickle.dumps(obj, protocol=pickle2.best_protocol)
                if len(payload) >= self.SHM_THRESHOLD:
                    name = "veles-update-%s" % (self.id or "x").replace("-", "")[:16]
                    SharedIO.put(name, payload)
                    self.sock.send_multipart(frames + [b"shm", name.encode()])
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

    def _recv(self, timeout_ms):
        if self.sock.poll(timeout_ms):
            frames = self.sock.recv_multipart()
            return frames[0].decode(), self._decode(frames[1:])
        return None, None

    # -- main loop ------------------------------------------------------------------------------------------------------
    def handshake(self):
        wf = self.workflow
        info = {"checksum": wf.checksum, "power": getattr(wf, "computing_power", 0), "mid": self.mid,
                "pid": os.getpid(), "host": socket.gethostname(), "argv": sys.argv,
                "backend": getattr(getattr(self.launcher, "device", None), "backend_name", None),
                "device": getattr(getattr(self.launcher, "device", None), "id", None),
                "executable": sys.executable, "cwd": os.getcwd(),
                "data": wf.generate_initial_data_for_master()}
        for attempt in range(self.reconnect_attempts):
            if self._stop.is_set():
                return False
            self._connect()
            self._send("handshake", info)
            cmd, reply = self._recv(int(self.handshake_timeout * 1000))
            if cmd == "handshake_ok":
                self.id = reply["id"]
                self.master_mid = reply.get("mid")
                if self.launcher is not None:
                    self.launcher.id = self.id
                    self.launcher.log_id = reply.get("log_id")
                wf.apply_initial_data_from_master(reply.get("data"))
                self.state = "WAIT"
                self.info("Connected to the master as %s", self.id)
                return True
            if cmd == "error":
                raise MasterSlaveCommunicationError("Master rejected the handshake: %s" % reply.get("error"))
            self.sock.close()
            self.debug("No answer from the master, retry %d/%d", attempt + 1, self.reconnect_attempts)
        raise MasterSlaveCommunicationError("Could not reach the master at %s" % self.address)

    def run(self):
        if not self.handshake():
            return
        wf = self.workflow
        last_ping = time.time()
        self._send("job_request")
        self.state = "GETTING_JOB"
        awaiting_ack = 0
        while not self._stop.is_set():
            cmd, data = self._recv(200)
            if time.time() - last_ping > 2.0:
                self._send("ping")
                last_ping = time.time()
            if cmd is None:
                continue
            if cmd == "job":
                if self.death_probability > 0 and random.random() < self.death_probability:
                    self.error("Fault injection: dying on job receipt (--slave-death-probability)")
                    self.sock.close()
                    raise Bug("slave death was injected")
                self.state = "BUSY"
                t0 = time.time()
                update = self._do_job(data)
                self.timings["job"] += time.time() - t0
                self.jobs_done += 1
                if self.async_slave:
                    self._send("job_request")     # overlap the next job's generation with our upload
                t0 = time.time()
                self._send("update", update)
                self.timings["update"] += time.time() - t0
                awaiting_ack += 1
                self.state = "WAIT"
            elif cmd == "update_ack":
                awaiting_ack -= 1
                if not self.async_slave:
                    self._send("job_request")
                    self.state = "GETTING_JOB"
            elif cmd == "wait":
                self.state = "POSTPONED"      # the master will push a job when the sync point clears
            elif cmd == "end":
                self.state = "END"
                break
            elif cmd == "pause":
                if self.launcher is not None:
                    self.launcher.pause()
            elif cmd == "resume":
                if self.launcher is not None:
                    self.launcher.resume()
            elif cmd == "error":
                raise MasterSlaveCommunicationError(str(data))
        try:
            self._send("bye")
            self.sock.close()
        except zmq.ZMQError:
            pass
        if not wf._finished_ and wf.is_running:
            wf.on_workflow_finished()
        if self.launcher is not None:
            self.launcher.on_workflow_finished()

    def _do_job(self, job):
        """Run one workflow pass on the job; returns the update for the master."""
        wf = self.workflow
        self._job_event.clear()
        self._job_result = None

        def callback(update):
            self._job_result = update
            self._job_event.set()
        wf.do_job(job, None, callback)
        while not self._job_event.wait(0.2):
            pool = wf.thread_pool
            if pool is not None and pool.failure is not None:
                raise pool.failure
        return self._job_result

    def stop(self):
        self._stop.set()
