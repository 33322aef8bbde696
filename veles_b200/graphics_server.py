"""GraphicsServer: ships filled plotters to a renderer.

Parity target: ``veles/graphics_server.py:73-245``: ``enqueue(plotter)`` -> stripped pickle -> compression ->
ZeroMQ PUB on inproc/ipc/tcp endpoints; a separate client process renders.  Here the default renderer is an
in-process background thread writing SVG files (headless); the PUB socket is still offered so that
``python -m veles_b200.graphics_client <endpoint>`` can render elsewhere."""
from __future__ import annotations

import pickle
import queue
import threading
import zlib

from .config import root, get
from .logger import Logger


class GraphicsServer(Logger):
    _instance = None

    def __init__(self, publish=True, render_inline=True):
        super(GraphicsServer, self).__init__()
        self.queue = queue.Queue()
        self.endpoints = []
        self.render_inline = render_inline
        self.rendered = []
        self._sock = None
        if publish:
            try:
                import zmq
                self._sock = zmq.Context.instance().socket(zmq.PUB)
                self._sock.setsockopt(zmq.LINGER, 0)
                port = self._sock.bind_to_random_port("tcp://127.0.0.1")
                self.endpoints.append("tcp://127.0.0.1:%d" % port)
            except Exception as e:  # noqa: BLE001
                self.debug("plot publishing is off: %s", e)
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._loop, name="graphics-server", daemon=True)
        self._thread.start()

    @staticmethod
    def launch(launcher=None, **kwargs):
        if GraphicsServer._instance is None:
            GraphicsServer._instance = GraphicsServer(**kwargs)
        return GraphicsServer._instance

    def enqueue(self, plotter):
        plotter.stripped_pickle = True
        try:
            blob = pickle.dumps(plotter, protocol=pickle.HIGHEST_PROTOCOL)
        finally:
            plotter.stripped_pickle = False
        self.queue.put(blob)

    def _loop(self):
        while not self._stop.is_set():
            try:
                blob = self.queue.get(timeout=0.2)
            except queue.Empty:
                continue
            if self._sock is not None:
                try:
                    self._sock.send_multipart([b"graphics", zlib.compress(blob, 1)])
                except Exception:  # noqa: BLE001
                    pass
            if self.render_inline:
                try:
                    plotter = pickle.loads(blob)
                    path = plotter.redraw()
                    if path:
                        self.rendered.append(path)
                except Exception:  # noqa: BLE001
                    self.exception("failed to render a plot")
            self.queue.task_done()

    def flush(self, timeout=5.0):
        import time
        t0 = time.time()
        while not self.queue.empty() and time.time() - t0 < timeout:
            time.sleep(0.05)
        time.sleep(0.05)

    def shutdown(self):
        self._stop.set()
        if self._sock is not None:
            try:     # tell remote clients that the stream is over (the reference pickles None for that too)
                self._sock.send_multipart([b"graphics", zlib.compress(pickle.dumps(None), 1)])
            except Exception:  # noqa: BLE001
                pass
            self._sock.close()
        GraphicsServer._instance = None
