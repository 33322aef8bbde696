"""GraphicsClient: subscribes to a GraphicsServer endpoint and renders the received plotters
(reference: ``veles/graphics_client.py:84-417``; GUI backends are unavailable here — files are written).
Run as ``python -m veles_b200.graphics_client tcp://host:port [output_dir]``."""
from __future__ import annotations

import pickle
import sys
import zlib


class GraphicsClient(object):
    def __init__(self, endpoint, output_dir=None):
        import zmq
        self.sock = zmq.Context.instance().socket(zmq.SUB)
        self.sock.connect(endpoint)
        self.sock.setsockopt(zmq.SUBSCRIBE, b"graphics")
        self.output_dir = output_dir
        self.rendered = []

    def poll_once(self, timeout_ms=1000):
        if not self.sock.poll(timeout_ms):
            return None
        topic, blob = self.sock.recv_multipart()
        plotter = pickle.loads(zlib.decompress(blob))
        if self.output_dir:
            plotter.output_dir = self.output_dir
        path = plotter.redraw()
        self.rendered.append(path)
        return path

    def run(self):
        while True:
            self.poll_once()


def main(argv=None):
    """Console entry point (``veles_graphics_client`` in pyproject.toml): endpoint [output_dir]."""
    argv = sys.argv[1:] if argv is None else list(argv)
    if not argv or argv[0] in ("-h", "--help"):
        print("usage: veles_graphics_client tcp://host:port [output_dir]")
        return 0 if argv else 2
    GraphicsClient(argv[0], argv[1] if len(argv) > 1 else None).run()
    return 0


if __name__ == "__main__":
    sys.exit(main())
