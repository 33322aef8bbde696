"""Web status server: receives master heartbeats on ``POST /update`` and serves the list of running workflows,
their slaves, graph (DOT) and recent log/events on ``/service`` + a minimal HTML page on ``/``.

Parity target: ``veles/web_status.py:113-314`` (Tornado + MongoDB there; ``http.server`` + the in-process
event sink here), garbage collection of sessions that stopped reporting (``drop_time``)."""
from __future__ import annotations

import json
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from .config import root, get
from .logger import Logger, events


class WebServer(Logger):
    def __init__(self, host=None, port=None, drop_time=None):
        super(WebServer, self).__init__()
        self.host = host or get(root.common.web.host, "127.0.0.1")
        self.port = port if port is not None else get(root.common.web.port, 8090)
        self.drop_time = drop_time or get(root.common.web.drop_time, 30 * 24 * 3600)
        self.masters = {}
        self.lock = threading.Lock()
        self.httpd = None
        self.thread = None

    def receive_update(self, data):
        with self.lock:
            data["last_update"] = time.time()
            self.masters[data["id"]] = data
            self._gc()

    def _gc(self):
        now = time.time()
        for mid in [m for m, d in self.masters.items() if now - d["last_update"] > self.drop_time]:
            del self.masters[mid]

    def service(self, request):
        kind = request.get("request", "workflows")
        with self.lock:
            if kind == "workflows":
                args = request.get("args", [])
                fields = args or ["name", "master", "time", "user", "mode", "slaves", "description"]
                return {mid: {k: d.get(k) for k in fields} for mid, d in self.masters.items()}
            if kind == "graph":
                d = self.masters.get(request.get("id"))
                return {"graph": d.get("graph") if d else None}
        if kind == "events":
            session = request.get("session")
            recs = [r for r in events.records if session in (None, r.get("session"))]
            return {"events": recs[-int(request.get("limit", 200)):]}
        return {"error": "unknown request %r" % kind}

    def page(self):
        rows = []
        with self.lock:
            for mid, d in sorted(self.masters.items(), key=lambda kv: -kv[1]["last_update"]):
                rows.append("<tr><td>%s</td><td>%s</td><td>%s</td><td>%.0f s</td><td>%d</td></tr>"
                            % (d.get("name"), d.get("master"), d.get("mode"), d.get("time", 0),
                               len(d.get("slaves") or {})))
        return ("<html><head><title>veles_b200 status</title></head><body><h2>Running workflows</h2>"
                "<table border=1><tr><th>workflow</th><th>host</th><th>mode</th><th>uptime</th><th>slaves</th></tr>"
                "%s</table></body></html>" % "".join(rows))

    def start(self):
        server = self

        class Handler(BaseHTTPRequestHandler):
            def log_message(self, fmt, *args):
                server.debug(fmt, *args)

            def _json(self, obj, code=200):
                body = json.dumps(obj, default=str).encode()
                self.send_response(code)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def do_POST(self):
                try:
                    data = json.loads(self.rfile.read(int(self.headers.get("Content-Length", 0))) or b"{}")
                except ValueError:
                    return self._json({"error": "bad json"}, 400)
                if self.path == "/update":
                    if "id" not in data:
                        return self._json({"error": "id is required"}, 400)
                    server.receive_update(data)
                    return self._json({"ok": True})
                if self.path == "/service":
                    return self._json(server.service(data))
                self._json({"error": "not found"}, 404)

            def do_GET(self):
                body = server.page().encode()
                self.send_response(200)
                self.send_header("Content-Type", "text/html")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

        self.httpd = ThreadingHTTPServer((self.host, self.port), Handler)
        self.port = self.httpd.server_address[1]
        self.thread = threading.Thread(target=self.httpd.serve_forever, name="web-status", daemon=True)
        self.thread.start()
        self.info("Web status is listening on http://%s:%d", self.host, self.port)
        return self

    def stop(self):
        if self.httpd is not None:
            self.httpd.shutdown()
            self.httpd.server_close()
            self.httpd = None


def main():
    srv = WebServer().start()
    try:
        while True:
            time.sleep(1)
    except KeyboardInterrupt:
        srv.stop()


if __name__ == "__main__":
    main()
