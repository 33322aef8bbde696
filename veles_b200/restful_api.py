"""RESTfulAPI: serves a trained model as ``POST /api``.

Parity target: ``veles/restful_api.py:54-217``: request JSON ``{"input": ..., "codec": "list"|"base64",
"shape": [...], "type": "float32"}``; the sample goes to ``loader.feed``; the unit, linked after the last
forward unit, answers every request of the finished minibatch with ``{"result": ...}``.  Twisted's web server
is replaced by ``http.server`` in a thread."""
from __future__ import annotations

import base64
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import numpy

from .config import root, get
from .distributable import TriviallyDistributable
from .units import Unit


class _Pending(object):
    def __init__(self):
        self.event = threading.Event()
        self.result = None


class RESTfulAPI(Unit, TriviallyDistributable):
    """HTTP endpoint that turns a posted sample into a request for the RestfulLoader and answers with the
    model's output."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(RESTfulAPI, self).__init__(workflow, **kwargs)
        self.port = kwargs.get("port", get(root.common.api.port, 0))
        self.path = kwargs.get("path", get(root.common.api.path, "/api"))
        self.response_timeout = kwargs.get("response_timeout", 10.0)
        self.demand("feed", "requests", "results", "minibatch_size")

    def init_unpickled(self):
        super(RESTfulAPI, self).init_unpickled()
        self.server_ = None
        self.thread_ = None

    def initialize(self, **kwargs):
        if self.server_ is not None:
            return
        api = self

        class Handler(BaseHTTPRequestHandler):
            def log_message(self, fmt, *args):
                api.debug(fmt, *args)

            def _fail(self, code, msg):
                body = json.dumps({"error": msg}).encode()
                self.send_response(code)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def do_POST(self):
                if self.path != api.path:
                    return self._fail(404, "unknown path")
                try:
                    req = json.loads(self.rfile.read(int(self.headers.get("Content-Length", 0))))
                    sample = api.decode_input(req)
                except Exception as e:  # noqa: BLE001
                    return self._fail(400, "invalid request: %s" % e)
                pending = _Pending()
                api.feed(sample, pending)
                if not pending.event.wait(api.response_timeout):
                    return self._fail(504, "the model did not answer in time")
                body = json.dumps({"result": pending.result}).encode()
                self.send_response(200)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

        self.server_ = ThreadingHTTPServer(("127.0.0.1", self.port), Handler)
        self.port = self.server_.server_address[1]
        self.thread_ = threading.Thread(target=self.server_.serve_forever, name="restful-api", daemon=True)
        self.thread_.start()
        self.info("Listening on http://127.0.0.1:%d%s", self.port, self.path)

    @staticmethod
    def decode_input(req):
        codec = req.get("codec", "list")
        if codec == "list":
            return numpy.asarray(req["input"], dtype=numpy.float32)
        if codec == "base64":
            arr = numpy.frombuffer(base64.b64decode(req["input"]), dtype=numpy.dtype(req.get("type", "float32")))
            return arr.reshape(req["shape"]).astype(numpy.float32)
        raise ValueError("unknown codec %r" % codec)

    def run(self):
        """Answer the requests of the minibatch that just went through the network."""
        res = self.results
        if hasattr(res, "map_read"):
            res.map_read()
            res = res.mem
        for i, pending in enumerate(self.requests[:int(self.minibatch_size)]):
            if pending is None:
                continue
            row = numpy.asarray(res[i])
            pending.result = row.tolist()
            pending.event.set()

    def stop(self):
        if self.server_ is not None:
            self.server_.shutdown()
            self.server_.server_close()
            self.server_ = None
