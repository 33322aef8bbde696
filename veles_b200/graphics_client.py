"""GraphicsClient: subscribes to a GraphicsServer endpoint and renders the plotters it receives
(reference: ``veles/graphics_client.py:84-417``: a separate process that unpickles the stripped plotters coming over
ZeroMQ, calls ``redraw()`` under a matplotlib GUI / WebAgg backend, dumps all figures to PDF on SIGUSR2 and exits when
the server sends ``None``).  No display exists on this box: plots are rendered to SVG files, and the browser view of the
reference's WebAgg mode is an ``index.html`` gallery that refreshes itself and always shows the newest file of every
plotter.  SIGUSR2 (or ``--report``) writes the gallery as one self-contained HTML file with the SVGs inlined.
Run as ``python -m veles_b200.graphics_client tcp://host:port [output_dir] [--once] [--refresh SECONDS]``."""
from __future__ import annotations

import html
import os
import pickle
import signal
import sys
import time
import zlib


class GraphicsClient(object):
    def __init__(self, endpoint, output_dir=None, refresh=2, index=True):
        import zmq
        self.sock = zmq.Context.instance().socket(zmq.SUB)
        self.sock.connect(endpoint)
        self.sock.setsockopt(zmq.SUBSCRIBE, b"graphics")
        self.endpoint = endpoint
        self.output_dir = output_dir
        self.refresh = refresh
        self.index = index and bool(output_dir)
        self.rendered = []          # every file written, in order
        self.latest = {}            # plotter name -> (path, time)
        self.finished = False       # the server said goodbye
        self._report_requested = False

    # -- receiving --------------------------------------------------------------------------------------------------
    def poll_once(self, timeout_ms=1000):
        """Render one pending plotter; returns the written path, None when nothing arrived (or the stream ended:
        ``finished`` is then set)."""
        if not self.sock.poll(timeout_ms):
            return None
        topic, blob = self.sock.recv_multipart()
        plotter = pickle.loads(zlib.decompress(blob))
        if plotter is None:
            self.finished = True
            return None
        if self.output_dir:
            plotter.output_dir = self.output_dir
        path = plotter.redraw()
        if path:
            self.rendered.append(path)
            self.latest[getattr(plotter, "name", None) or type(plotter).__name__] = (path, time.time())
            if self.index:
                self.write_index()
        return path

    def run(self, once=False):
        try:
            signal.signal(signal.SIGUSR2, self._on_sigusr2)
        except (ValueError, AttributeError):      # not the main thread / no such signal
            pass
        while not self.finished:
            path = self.poll_once()
            if self._report_requested:
                self._report_requested = False
                self.write_report()
            if once and path:
                break
        return self.rendered

    def _on_sigusr2(self, signum, frame):
        self._report_requested = True

    # -- browser views ----------------------------------------------------------------------------------------------
    def _entries(self):
        return sorted(self.latest.items(), key=lambda kv: kv[0])

    def write_index(self):
        """``index.html`` in the output directory: the newest file of every plotter, reloading every ``refresh`` s."""
        rows = []
        for name, (path, t) in self._entries():
            rel = os.path.relpath(path, self.output_dir)
            rows.append("<figure><figcaption>%s <small>(%s)</small></figcaption><img src='%s?t=%d' alt='%s'></figure>"
                        % (html.escape(name), time.strftime("%H:%M:%S", time.localtime(t)), html.escape(rel), int(t),
                           html.escape(name)))
        page = ("<!doctype html><html><head><meta charset='utf-8'><meta http-equiv='refresh' content='%d'>"
                "<title>veles_b200 plots</title><style>body{font-family:sans-serif}figure{display:inline-block;"
                "margin:8px;vertical-align:top}img{max-width:640px;border:1px solid #ccc}</style></head><body>"
                "<h3>%s</h3>%s</body></html>" % (max(1, int(self.refresh)), html.escape(self.endpoint), "\n".join(rows)))
        path = os.path.join(self.output_dir, "index.html")
        tmp = path + ".part"
        with open(tmp, "w") as f:
            f.write(page)
        os.replace(tmp, path)
        return path

    def write_report(self, path=None):
        """One self-contained HTML file with the newest plot of every plotter inlined (the reference writes all open
        figures into a PDF on SIGUSR2)."""
        path = path or os.path.join(self.output_dir or ".", "plots_%s.html" % time.strftime("%Y%m%d_%H%M%S"))
        parts = ["<!doctype html><html><head><meta charset='utf-8'><title>veles_b200 plots</title></head><body>"]
        for name, (src, t) in self._entries():
            parts.append("<h3>%s</h3>" % html.escape(name))
            if src.endswith(".svg") and os.path.exists(src):
                with open(src) as f:
                    parts.append(f.read())
            else:
                parts.append("<p>%s</p>" % html.escape(src))
        parts.append("</body></html>")
        with open(path, "w") as f:
            f.write("\n".join(parts))
        return path


def main(argv=None):
    """Console entry point (``veles_graphics_client`` in pyproject.toml)."""
    import argparse
    ap = argparse.ArgumentParser(prog="veles_graphics_client", description=__doc__.split("\n\n")[0])
    ap.add_argument("endpoint", help="the server's PUB endpoint, e.g. tcp://127.0.0.1:5555")
    ap.add_argument("output_dir", nargs="?", default=None, help="where the plots are written (default: the plotters' own)")
    ap.add_argument("--once", action="store_true", help="exit after the first rendered plot")
    ap.add_argument("--refresh", type=int, default=2, help="reload period of index.html, seconds")
    ap.add_argument("--no-index", action="store_true", help="do not maintain index.html")
    ap.add_argument("--report", default=None, help="on exit, write all newest plots into this single HTML file")
    ap.add_argument("-b", "--backend", default=None, help="accepted for compatibility (rendering is headless SVG)")
    argv = sys.argv[1:] if argv is None else list(argv)
    if not argv:
        ap.print_usage()
        return 2
    a = ap.parse_args(argv)
    client = GraphicsClient(a.endpoint, a.output_dir, refresh=a.refresh, index=not a.no_index)
    try:
        client.run(once=a.once)
    except KeyboardInterrupt:
        pass
    if a.report:
        client.write_report(a.report)
    return 0


if __name__ == "__main__":
    sys.exit(main())
