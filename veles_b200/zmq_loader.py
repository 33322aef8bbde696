"""ZeroMQLoader: a slave-side ROUTER socket through which external producers (the reference's Java
"Mastodon" bridge) push work items; the bound endpoints are reported to the master in the handshake
(reference: ``veles/zmq_loader.py:47-138``, ``negotiates_on_connect = True`` at ``:85``)."""
from __future__ import annotations

import pickle
import queue
import threading

import numpy
import zmq

from .loader.base import Loader, TEST
from .loader.interactive import InteractiveLoader
from .network_common import machine_id


class ZeroMQLoader(InteractiveLoader):
    """Samples arrive as pickled messages on a ZeroMQ socket; one job per message."""
    MAPPING = "zeromq"
    hide_from_registry = False
    negotiates_on_connect = True

    def __init__(self, workflow, **kwargs):
        super(ZeroMQLoader, self).__init__(workflow, **kwargs)
        self.bind_address = kwargs.get("bind_address", "tcp://127.0.0.1")
        self.endpoints = {}

    def init_unpickled(self):
        super(ZeroMQLoader, self).init_unpickled()
        self.sock_ = None
        self.thread_ = None
        self.stop_ = threading.Event()
        self.cids_ = queue.Queue()

    def initialize(self, device=None, **kwargs):
        res = super(ZeroMQLoader, self).initialize(device=device, **kwargs)
        if self.sock_ is None:
            ctx = zmq.Context.instance()
            self.sock_ = ctx.socket(zmq.ROUTER)
            self.sock_.setsockopt(zmq.LINGER, 0)
            port = self.sock_.bind_to_random_port(self.bind_address)
            self.endpoints = {"tcp": "%s:%d" % (self.bind_address, port), "mid": machine_id()}
            self.thread_ = threading.Thread(target=self._serve, name="zmq-loader", daemon=True)
            self.thread_.start()
        return res

    def _serve(self):
        poller = zmq.Poller()
        poller.register(self.sock_, zmq.POLLIN)
        while not self.stop_.is_set():
            if not dict(poller.poll(100)):
                continue
            frames = self.sock_.recv_multipart()
            routing, cid, payload = frames[0], frames[1], frames[-1]
            try:
                obj = pickle.loads(payload)
            except Exception:  # noqa: BLE001 - raw float32 bytes are accepted too
                obj = numpy.frombuffer(payload, dtype=numpy.float32)
            self.feed(obj)
            self.sock_.send_multipart([routing, cid, b"ok"])

    def stop(self):
        self.stop_.set()

    def generate_initial_data_for_master(self):
        return {"endpoints": self.endpoints}

    def apply_initial_data_from_slave(self, data, slave):
        self.info("Slave %s accepts external work at %s", getattr(slave, "id", slave), data.get("endpoints"))
        self.__dict__.setdefault("slave_endpoints", {})[getattr(slave, "id", slave)] = data.get("endpoints")
