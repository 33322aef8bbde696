"""The command line composer (the reference's ``--frontend``: ``veles/scripts/generate_frontend.py`` builds the form,
``veles/__main__.py:258-327`` serves it and waits for the composed command line).

``generate()`` renders every option of the argument parser as a form control (checkbox for flags, select for choices,
text otherwise) plus the positional workflow / config / overrides; a few lines of JavaScript assemble the command line
live.  ``serve()`` runs a local HTTP server with that page, returns the argv the user posted ("Run") or None
("Cancel" / interrupt).  ``python -m veles_b200.scripts.generate_frontend out.html`` writes the static page."""
from __future__ import annotations

import argparse
import html
import json
import shlex
import sys
import threading

SCRIPT = """
function quote(s){return /^[A-Za-z0-9_.,:=\\/@%+-]+$/.test(s)?s:"'"+s.replace(/'/g,"'\\\\''")+"'";}
function compose(){
  const parts=[];
  document.querySelectorAll('[data-opt]').forEach(el=>{
    const o=el.dataset.opt;
    if(el.type==='checkbox'){if(el.checked!==(el.dataset.def==='1'))parts.push(o);}
    else if(el.value!==''&&el.value!==el.dataset.def){parts.push(o);parts.push(quote(el.value));}
  });
  ['workflow','config'].forEach(id=>{const v=document.getElementById(id).value.trim();if(v)parts.push(quote(v));});
  const ov=document.getElementById('overrides').value.trim();
  if(ov)ov.split(/\\s+/).forEach(x=>parts.push(quote(x)));
  document.getElementById('cmdline').value=parts.join(' ');
}
function send(run){
  fetch('/cmdline',{method:'POST',body:JSON.stringify({cmdline:run?document.getElementById('cmdline').value:null})})
    .then(()=>{document.body.innerHTML='<h3>'+(run?'Started.':'Cancelled.')+' This page can be closed.</h3>';});
}
document.addEventListener('input',compose);document.addEventListener('DOMContentLoaded',compose);
"""


def scan_workflows():
    """Workflow files that ship with the package (modules defining ``run(load, main)``) with their ``*_config.py``
    when there is one: the composer offers them in the workflow field (reference: ``scan_workflows``)."""
    import os
    import re
    base = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    found = []
    for root_dir, dirs, files in os.walk(base):
        dirs[:] = sorted(d for d in dirs if d not in ("__pycache__", "csrc", "lib", "tests"))
        for fn in sorted(files):
            if not fn.endswith(".py") or fn.endswith("_config.py") or fn.startswith("__"):
                continue
            path = os.path.join(root_dir, fn)
            try:
                with open(path, errors="replace") as f:
                    text = f.read()
            except OSError:
                continue
            if re.search(r"^def run\(load, main\):", text, re.M):
                cfg = path[:-3] + "_config.py"
                found.append((os.path.relpath(path, os.path.dirname(base)),
                              os.path.relpath(cfg, os.path.dirname(base)) if os.path.exists(cfg) else "-"))
    return found


def _control(act):
    opt = act.option_strings[-1]
    q = html.escape(opt, quote=True)
    if isinstance(act, (argparse._StoreTrueAction, argparse._StoreFalseAction, argparse._CountAction)) or act.nargs == 0:
        return "<input type='checkbox' data-opt='%s' data-def='0'>" % q
    default = "" if act.default in (None, argparse.SUPPRESS) else str(act.default)
    d = html.escape(default, quote=True)
    if act.choices:
        opts = "".join("<option%s>%s</option>" % (" selected" if str(c) == default else "", html.escape(str(c)))
                       for c in ([""] if default == "" else []) + list(act.choices))
        return "<select data-opt='%s' data-def='%s'>%s</select>" % (q, d, opts)
    return "<input type='text' size='28' data-opt='%s' data-def='%s' value='%s'>" % (q, d, d)


def generate(cmdline=""):
    from ..__main__ import Main
    parser = Main.init_parser(ignore_conflicts=True)
    given = shlex.split(cmdline) if isinstance(cmdline, str) else list(cmdline)
    rows = []
    for act in parser._actions:
        if not act.option_strings or act.option_strings[-1] in ("--frontend", "--help", "-h"):
            continue
        rows.append("<tr><td><code>%s</code></td><td>%s</td><td>%s</td></tr>"
                    % (html.escape(", ".join(act.option_strings)), _control(act), html.escape(act.help or "")))
    positional = [a for a in given if not a.startswith("-")]
    wf = html.escape(positional[0], quote=True) if positional else ""
    cfg = html.escape(positional[1], quote=True) if len(positional) > 1 else "-"
    ov = html.escape(" ".join(positional[2:]), quote=True)
    return ("<!doctype html><html><head><meta charset='utf-8'><title>veles_b200 command line</title>"
            "<style>body{font-family:sans-serif;margin:1.2em}table{border-collapse:collapse}"
            "td,th{border:1px solid #999;padding:3px 8px}#cmdline{width:100%%;font-family:monospace}</style>"
            "<script>%s</script></head><body>"
            "<h2>veles_b200 [options] workflow.py [config.py] [root.x.y=value ...]</h2>"
            "<p>workflow <input id='workflow' size='48' value='%s' list='workflows'> config "
            "<input id='config' size='32' value='%s'><datalist id='workflows'>%s</datalist>"
            "<br>overrides <input id='overrides' size='90' value='%s' placeholder='root.loader.minibatch_size=128'></p>"
            "<table><tr><th>option</th><th>value</th><th>meaning</th></tr>%s</table>"
            "<p><input id='cmdline' readonly></p>"
            "<p><button onclick='send(true)'>Run</button> <button onclick='send(false)'>Cancel</button></p>"
            "</body></html>" % (SCRIPT, wf, cfg, "".join(
                "<option value='%s'>" % html.escape(w, quote=True) for w, _c in scan_workflows()), ov, "\n".join(rows)))


def serve(argv=(), host="127.0.0.1", port=0, open_browser=True, ready=None, timeout=None):
    """Serve the composer until a command line is posted; returns it as an argv list (None when cancelled).
    ``ready(port)`` is called once the socket listens (tests drive the server through it)."""
    from http.server import BaseHTTPRequestHandler, HTTPServer
    result = {"cmdline": None}
    page = generate(list(argv)).encode()
    done = threading.Event()

    class Handler(BaseHTTPRequestHandler):
        def log_message(self, *a):  # noqa: D102 - quiet
            pass

        def do_GET(self):
            self.send_response(200)
            self.send_header("Content-Type", "text/html; charset=utf-8")
            self.send_header("Content-Length", str(len(page)))
            self.end_headers()
            self.wfile.write(page)

        def do_POST(self):
            n = int(self.headers.get("Content-Length", 0))
            try:
                result["cmdline"] = json.loads(self.rfile.read(n) or b"{}").get("cmdline")
                code = 200
            except ValueError:
                code = 400
            self.send_response(code)
            self.send_header("Content-Length", "0")
            self.end_headers()
            if code == 200:
                done.set()

    srv = HTTPServer((host, port), Handler)
    url = "http://%s:%d/" % (host, srv.server_address[1])
    th = threading.Thread(target=srv.serve_forever, kwargs={"poll_interval": 0.05}, daemon=True)
    th.start()
    try:
        print("Compose the command line at %s" % url)
        if ready is not None:
            ready(srv.server_address[1])
        if open_browser:
            try:
                import webbrowser
                webbrowser.open(url)
            except Exception:  # noqa: BLE001 - headless boxes have no browser; the URL was printed
                pass
        try:
            done.wait(timeout)
        except KeyboardInterrupt:
            pass
    finally:
        srv.shutdown()
        srv.server_close()
    return shlex.split(result["cmdline"]) if result["cmdline"] else None


def main(argv=None):
    argv = argv if argv is not None else sys.argv[1:]
    text = generate()
    if argv:
        with open(argv[0], "w") as f:
            f.write(text)
    else:
        print(text)
    return 0


if __name__ == "__main__":
    sys.exit(main())
