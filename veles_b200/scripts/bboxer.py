"""Bounding-box labelling web tool (reference: ``veles/scripts/bboxer.py``, 263 lines of tornado + JS).

``python -m veles_b200.scripts.bboxer --root <image dir> [--port 8090]`` serves:
  GET  /            one-page UI (drag to draw boxes on the current image, label them, save)
  GET  /images      JSON list of image paths under the root
  GET  /image/<i>   the image bytes
  GET  /boxes/<i>   saved boxes of image i
  POST /boxes/<i>   JSON [{x, y, width, height, label}] -> stored next to the image as <name>.json

The saved ``<image>.json`` files are what ``loader/image.py`` reads through ``bbox`` crop settings.
Built on the standard-library HTTP server (no tornado on this image).
"""
from __future__ import annotations

import argparse
import json
import mimetypes
import os
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

PAGE = """<!doctype html><html><head><meta charset="utf-8"><title>bboxer</title>
<style>body{font-family:sans-serif;margin:1em}#c{border:1px solid #888;cursor:crosshair}</style></head><body>
<div><button onclick="go(-1)">&larr;</button> <span id="name"></span> <button onclick="go(1)">&rarr;</button>
 label <input id="label" value="object"> <button onclick="save()">save</button>
 <button onclick="boxes=[];draw()">clear</button> <span id="st"></span></div>
<canvas id="c"></canvas>
<script>
let imgs=[],i=0,boxes=[],img=new Image(),c=document.getElementById('c'),g=c.getContext('2d'),d=null;
fetch('images').then(r=>r.json()).then(l=>{imgs=l;load()});
function load(){document.getElementById('name').textContent=(i+1)+'/'+imgs.length+' '+imgs[i];
 img.onload=()=>{c.width=img.width;c.height=img.height;fetch('boxes/'+i).then(r=>r.json()).then(b=>{boxes=b;draw()})};
 img.src='image/'+i}
function go(s){i=(i+s+imgs.length)%imgs.length;load()}
function draw(){g.drawImage(img,0,0);g.strokeStyle='#f00';g.fillStyle='#f00';g.lineWidth=2;
 for(const b of boxes){g.strokeRect(b.x,b.y,b.width,b.height);g.fillText(b.label,b.x+3,b.y+12)}}
c.onmousedown=e=>{d={x:e.offsetX,y:e.offsetY}};
c.onmousemove=e=>{if(!d)return;draw();g.strokeRect(d.x,d.y,e.offsetX-d.x,e.offsetY-d.y)};
c.onmouseup=e=>{if(!d)return;const w=e.offsetX-d.x,h=e.offsetY-d.y;
 if(Math.abs(w)>3&&Math.abs(h)>3)boxes.push({x:Math.min(d.x,e.offsetX),y:Math.min(d.y,e.offsetY),
 width:Math.abs(w),height:Math.abs(h),label:document.getElementById('label').value});d=null;draw()};
function save(){fetch('boxes/'+i,{method:'POST',body:JSON.stringify(boxes)}).then(r=>{
 document.getElementById('st').textContent=r.ok?'saved':'error'})}
</script></body></html>"""


def scan(root):
    out = []
    for d, _, files in sorted(os.walk(root)):
        for f in sorted(files):
            mt = mimetypes.guess_type(f)[0] or ""
            if mt.startswith("image/"):
                out.append(os.path.relpath(os.path.join(d, f), root))
    return out


def boxes_path(root, rel):
    return os.path.join(root, os.path.splitext(rel)[0] + ".json")


def make_handler(root):
    images = scan(root)

    class Handler(BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def _send(self, code, body, ctype="application/json"):
            self.send_response(code)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def _index(self, prefix):
            try:
                i = int(self.path[len(prefix):])
                return i if 0 <= i < len(images) else None
            except ValueError:
                return None

        def do_GET(self):
            if self.path in ("/", "/index.html"):
                return self._send(200, PAGE.encode(), "text/html; charset=utf-8")
            if self.path == "/images":
                return self._send(200, json.dumps(images).encode())
            if self.path.startswith("/image/"):
                i = self._index("/image/")
                if i is None:
                    return self._send(404, b"{}")
                with open(os.path.join(root, images[i]), "rb") as f:
                    return self._send(200, f.read(), mimetypes.guess_type(images[i])[0] or "application/octet-stream")
            if self.path.startswith("/boxes/"):
                i = self._index("/boxes/")
                if i is None:
                    return self._send(404, b"{}")
                p = boxes_path(root, images[i])
                data = open(p, "rb").read() if os.path.exists(p) else b"[]"
                return self._send(200, data)
            self._send(404, b"{}")

        def do_POST(self):
            i = self._index("/boxes/") if self.path.startswith("/boxes/") else None
            if i is None:
                return self._send(404, b"{}")
            body = self.rfile.read(int(self.headers.get("Content-Length", "0")))
            try:
                boxes = json.loads(body.decode())
                for b in boxes:
                    for k in ("x", "y", "width", "height"):
                        b[k] = float(b[k])
                    b["label"] = str(b.get("label", ""))
            except (ValueError, KeyError, TypeError):
                return self._send(400, b'{"error": "bad boxes"}')
            with open(boxes_path(root, images[i]), "w") as f:
                json.dump(boxes, f)
            self._send(200, b'{"ok": true}')

    return Handler


def serve(root, port=8090, host="127.0.0.1"):
    srv = ThreadingHTTPServer((host, port), make_handler(os.path.abspath(root)))
    return srv


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--root", required=True)
    ap.add_argument("--port", type=int, default=8090)
    ap.add_argument("--host", default="127.0.0.1")
    a = ap.parse_args(argv)
    srv = serve(a.root, a.port, a.host)
    print("bboxer on http://%s:%d/ (%s)" % (a.host, srv.server_address[1], a.root))
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass


if __name__ == "__main__":
    main()
