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
from .logger import Logger, events, logs, query_records, read_jsonl_dir

PAGE = """<!doctype html><html><head><meta charset="utf-8"><title>veles_b200 status</title>
<style>body{font-family:sans-serif;margin:1.2em}table{border-collapse:collapse}td,th{border:1px solid #999;padding:3px 8px}
tr.sel{background:#def}pre{background:#f4f4f4;padding:6px;max-height:18em;overflow:auto}.ERROR,.CRITICAL{color:#b00}
.WARNING{color:#a60}input,select{margin-right:8px}</style></head><body>
<h2>Running workflows</h2><table id="wf"><thead><tr><th>workflow</th><th>host</th><th>mode</th><th>uptime</th>
<th>slaves</th><th>user</th></tr></thead><tbody></tbody></table>
<h3 id="title"></h3><div id="slaves"></div><div id="graphsvg" style="overflow:auto;max-height:32em"></div>
<details><summary>graph (DOT)</summary><pre id="graph"></pre></details>
<h3>Logs</h3>level <select id="level"><option>DEBUG</option><option selected>INFO</option><option>WARNING</option>
<option>ERROR</option></select>text <input id="text" size="24"><label><input type="checkbox" id="all"> all sessions</label>
<table id="logs"><thead><tr><th>time</th><th>node</th><th>level</th><th>logger</th><th>message</th></tr></thead><tbody></tbody></table>
<h3>Timeline</h3><div id="timeline" style="overflow:auto"></div>
<h3>Events</h3><table id="events"><thead><tr><th>time</th><th>domain</th><th>name</th><th>type</th></tr></thead><tbody></tbody></table>
<script>
let selected = null;
async function ask(req){const r = await fetch('/service',{method:'POST',body:JSON.stringify(req)});return r.json();}
function esc(s){return String(s==null?'':s).replace(/[&<>]/g,c=>({'&':'&amp;','<':'&lt;','>':'&gt;'}[c]));}
function ts(t){return new Date(1000*t).toISOString().substr(11,12);}
function timeline(evs){
  // begin / end pairs of (domain, name, instance) become bars on one row per (domain, name); singles become ticks
  const open={}, rows={}, bars=[], ticks=[];
  let t0=Infinity, t1=-Infinity;
  for(const e of evs){ t0=Math.min(t0,e.time); t1=Math.max(t1,e.time);
    const key=e.domain+'.'+e.name; if(!(key in rows)) rows[key]=Object.keys(rows).length;
    const k=key+'#'+e.instance;
    if(e.type==='begin') open[k]=e.time;
    else if(e.type==='end' && k in open){ bars.push([rows[key],open[k],e.time,key]); delete open[k]; }
    else if(e.type==='single') ticks.push([rows[key],e.time,key]); }
  for(const k in open){ const key=k.split('#')[0]; bars.push([rows[key],open[k],t1,key+' (running)']); }
  const n=Object.keys(rows).length; if(!n||!(t1>t0)) return '';
  const W=900,L=170,H=16, sx=t=>L+(W-L-10)*(t-t0)/(t1-t0);
  let out=`<svg xmlns="http://www.w3.org/2000/svg" width="${W}" height="${n*H+18}" font-size="10" font-family="sans-serif">`;
  for(const [key,r] of Object.entries(rows)) out+=`<text x="2" y="${r*H+12}">${esc(key)}</text>`;
  for(const [r,a,b,label] of bars) out+=`<rect x="${sx(a).toFixed(1)}" y="${r*H+3}" width="${Math.max(1,sx(b)-sx(a)).toFixed(1)}" height="${H-5}" fill="#69c"><title>${esc(label)} ${(b-a).toFixed(3)} s</title></rect>`;
  for(const [r,t,label] of ticks) out+=`<line x1="${sx(t).toFixed(1)}" x2="${sx(t).toFixed(1)}" y1="${r*H+2}" y2="${r*H+H-2}" stroke="#c33"><title>${esc(label)}</title></line>`;
  out+=`<text x="${L}" y="${n*H+14}">0 s</text><text x="${W-60}" y="${n*H+14}">${(t1-t0).toFixed(1)} s</text></svg>`;
  return out;
}
async function refresh(){
  const w = await ask({request:'workflows'});
  const tb = document.querySelector('#wf tbody'); tb.innerHTML='';
  for (const [id,d] of Object.entries(w)){
    const tr=document.createElement('tr'); if(id===selected) tr.className='sel';
    tr.innerHTML=`<td>${esc(d.name)}</td><td>${esc(d.master)}</td><td>${esc(d.mode)}</td><td>${Math.round(d.time||0)} s</td>`+
      `<td>${Object.keys(d.slaves||{}).length}</td><td>${esc(d.user)}</td>`;
    tr.onclick=()=>{selected=id;refresh();}; tb.appendChild(tr);
    if(selected===null) selected=id;
  }
  if(selected && w[selected]){
    document.getElementById('title').textContent=w[selected].name+' ('+selected+')';
    const sl=w[selected].slaves||{};
    document.getElementById('slaves').innerHTML=Object.entries(sl).map(([k,v])=>esc(k)+': '+esc(JSON.stringify(v))).join('<br>');
    const g=await ask({request:'graph',id:selected}); document.getElementById('graph').textContent=g.graph||'';
    if(g.svg!==document.getElementById('graphsvg').dataset.src){document.getElementById('graphsvg').innerHTML=g.svg||'';
      document.getElementById('graphsvg').dataset.src=g.svg||'';}
  }
  const sess = document.getElementById('all').checked ? null : (selected && w[selected] ? w[selected].session : null);
  const lg = await ask({request:'logs',session:sess,level:document.getElementById('level').value,
                        text:document.getElementById('text').value,limit:100});
  document.querySelector('#logs tbody').innerHTML=(lg.logs||[]).reverse().map(r=>`<tr class="${esc(r.level)}"><td>${ts(r.time)}</td>`+
    `<td>${esc(r.node)}</td><td>${esc(r.level)}</td><td>${esc(r.name)}</td><td>${esc(r.message)}</td></tr>`).join('');
  const tl = await ask({request:'events',session:sess,limit:2000});
  document.getElementById('timeline').innerHTML = timeline(tl.events||[]);
  const ev = await ask({request:'events',session:sess,limit:60});
  document.querySelector('#events tbody').innerHTML=(ev.events||[]).reverse().map(r=>`<tr><td>${ts(r.time)}</td>`+
    `<td>${esc(r.domain)}</td><td>${esc(r.name)}</td><td>${esc(r.type)}</td></tr>`).join('');
}
refresh(); setInterval(refresh, 2000);
</script></body></html>"""


class WebServer(Logger):
    def __init__(self, host=None, port=None, drop_time=None, log_dir=None):
        """``log_dir``: directory with the ``*.logs.jsonl`` / ``*.events.jsonl`` sinks of the monitored processes
        (``root.common.dirs.logs``); the server's own in-process sinks are always searched too."""
        super(WebServer, self).__init__()
        self.log_dir = log_dir if log_dir is not None else get(root.common.dirs.logs, None)
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
                return {"graph": d.get("graph") if d else None, "svg": d.get("graph_svg") if d else None}
        if kind in ("events", "logs"):
            # the reference queried MongoDB collections here (veles/web_status.py:113-186): session / node / level /
            # logger or event name / text / since filters over the JSON-lines sinks
            own = events.records if kind == "events" else logs.records
            recs = read_jsonl_dir(self.log_dir, ".%s.jsonl" % kind) + list(own)
            seen, uniq = set(), []
            for r in recs:
                key = (r.get("session"), r.get("time"), r.get("name"), r.get("message"), r.get("type"))
                if key not in seen:
                    seen.add(key)
                    uniq.append(r)
            flt = {k: request.get(k) for k in ("session", "node", "level", "name", "domain", "text", "since")}
            return {kind: query_records(uniq, limit=request.get("limit", 200), **flt)}
        if kind == "sessions":
            recs = read_jsonl_dir(self.log_dir, ".logs.jsonl") + list(logs.records)
            return {"sessions": sorted({r.get("session") for r in recs if r.get("session")})}
        return {"error": "unknown request %r" % kind}

    def page(self):
        return PAGE

    def static_page(self):
        """Script-free table of the running workflows (``GET /plain``)."""
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
                body = (server.static_page() if self.path.startswith("/plain") else server.page()).encode()
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


def main(argv=None):
    """``python -m veles_b200.web_status [--host H] [--port P] [--log-dir DIR] [--drop-time S]``: the status server
    as a process of its own (reference: ``veles/web_status.py`` runs as a daemon the launchers POST to)."""
    import argparse
    ap = argparse.ArgumentParser(prog="veles_b200.web_status", description=main.__doc__)
    ap.add_argument("--host", default=None, help="interface to listen on (default: root.common.web.host)")
    ap.add_argument("--port", type=int, default=None, help="port (default: root.common.web.port)")
    ap.add_argument("--log-dir", default=None, help="directory with the *.logs.jsonl / *.events.jsonl sinks of the runs")
    ap.add_argument("--drop-time", type=float, default=None, help="forget a workflow this long after its last heartbeat, s")
    a = ap.parse_args(argv)
    srv = WebServer(host=a.host, port=a.port, drop_time=a.drop_time, log_dir=a.log_dir).start()
    print("status server on http://%s:%d/ (POST /update, POST /service)" % (srv.host, srv.port), flush=True)
    try:
        while True:
            time.sleep(1)
    except KeyboardInterrupt:
        pass
    finally:
        srv.stop()
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
