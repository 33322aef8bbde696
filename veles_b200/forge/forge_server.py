"""Forge server: stores versioned model packages and answers ``list``, ``details``, ``fetch``, ``upload``, ``delete``.

Parity target: ``veles/forge/forge_server.py:462-915`` (Tornado, git-backed storage, e-mail confirmed tokens).
Here: ``http.server`` + a directory per model with numbered versions; uploads need a token from ``register``."""
from __future__ import annotations

import json
import os
import shutil
import threading
import time
import uuid
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from urllib.parse import urlparse, parse_qs

from ..logger import Logger
from .forge_common import unpack, validate_manifest


class ForgeServer(Logger):
    def __init__(self, root_dir, host="127.0.0.1", port=0, require_confirmation=False, mailer=None):
        """``require_confirmation``: a new token stays pending until ``confirm(token, code)`` presents the code that
        was mailed to the address (reference: ``forge_server.py:462-520``).  ``mailer(email, subject, body)`` delivers
        it; the default writes ``outbox/<email>.txt`` under the storage root (no SMTP relay is assumed)."""
        super(ForgeServer, self).__init__()
        self.require_confirmation = bool(require_confirmation)
        self.mailer = mailer or self._outbox_mailer
        self.root_dir = os.path.abspath(root_dir)
        os.makedirs(self.root_dir, exist_ok=True)
        self.host, self.port = host, port
        self.tokens_file = os.path.join(self.root_dir, "tokens.json")
        self.tokens = {}
        if os.path.exists(self.tokens_file):
            with open(self.tokens_file) as f:
                self.tokens = json.load(f)
        self.lock = threading.Lock()
        self.httpd = None

    # -- storage ----------------------------------------------------------------------------------------------
    def model_dir(self, name):
        if not name or "/" in name or name.startswith("."):
            raise ValueError("invalid model name %r" % name)
        return os.path.join(self.root_dir, "models", name)

    def versions(self, name):
        d = self.model_dir(name)
        return sorted((v for v in os.listdir(d) if v.isdigit()), key=int) if os.path.isdir(d) else []

    # -- git history of a model (the reference keeps every model as a git repository: forge_server.py:462-560) -----
    def _git(self, name, *args, check=True):
        import subprocess
        repo = os.path.join(self.model_dir(name), "history")
        env = dict(os.environ, GIT_CONFIG_GLOBAL=os.devnull, GIT_CONFIG_SYSTEM=os.devnull)
        r = subprocess.run(["git", "-C", repo] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
        if check and r.returncode != 0:
            raise RuntimeError("git %s failed: %s" % (" ".join(args), r.stderr.decode(errors="replace")[-300:]))
        return r.stdout.decode(errors="replace")

    def _record_in_git(self, name, version, src_dir, email, manifest):
        """Mirror version ``version`` (already stored as a numbered directory) into the model's git repository:
        one commit per upload, tagged ``v<N>``.  Without a ``git`` executable the numbered directories are the store."""
        if shutil.which("git") is None:
            return False
        repo = os.path.join(self.model_dir(name), "history")
        try:
            if not os.path.isdir(os.path.join(repo, ".git")):
                os.makedirs(repo, exist_ok=True)
                self._git(name, "init", "-q")
            for entry in os.listdir(repo):                      # the tree of a version replaces the previous one
                if entry != ".git":
                    path = os.path.join(repo, entry)
                    shutil.rmtree(path) if os.path.isdir(path) else os.remove(path)
            for entry in os.listdir(src_dir):
                if entry in ("package.tar.gz", "owner"):
                    continue
                path = os.path.join(src_dir, entry)
                dst = os.path.join(repo, entry)
                shutil.copytree(path, dst) if os.path.isdir(path) else shutil.copy2(path, dst)
            self._git(name, "add", "-A")
            who = "%s <%s>" % (manifest.get("author") or email.split("@")[0], email)
            self._git(name, "-c", "user.name=forge", "-c", "user.email=forge@localhost", "commit", "-q", "--allow-empty",
                      "--author", who, "-m", "v%s: %s" % (version, manifest.get("short_description", "")))
            self._git(name, "tag", "-f", "v%s" % version)
            return True
        except Exception as e:  # noqa: BLE001 - the history is a mirror: the upload itself has succeeded
            self.warning("git history of %s was not updated: %s", name, e)
            return False

    def history(self, name):
        """[{version, commit, date, author, subject}] newest first (empty without git)."""
        if not self.versions(name):
            raise KeyError(name)
        if shutil.which("git") is None or not os.path.isdir(os.path.join(self.model_dir(name), "history", ".git")):
            return []
        out = []
        log = self._git(name, "log", "--format=%H%x1f%aI%x1f%an <%ae>%x1f%s%x1f%D", check=False)
        for line in log.splitlines():
            h, date, author, subject, refs = (line.split("\x1f") + [""] * 5)[:5]
            tags = [r.strip()[5:] for r in refs.split(",") if r.strip().startswith("tag: ")]
            out.append({"version": tags[0] if tags else None, "commit": h, "date": date, "author": author,
                        "subject": subject})
        return out

    def diff(self, name, a, b):
        """``git diff --stat`` between two versions of a model."""
        vs = ["v" + v for v in self.versions(name)]
        if a not in vs or b not in vs:
            raise KeyError("%s %s..%s" % (name, a, b))
        return self._git(name, "diff", "--stat", a, b)

    def list(self):
        base = os.path.join(self.root_dir, "models")
        out = []
        for name in sorted(os.listdir(base)) if os.path.isdir(base) else []:
            vs = self.versions(name)
            if vs:
                with open(os.path.join(base, name, vs[-1], "manifest.json")) as f:
                    m = json.load(f)
                out.append([name, m.get("short_description", ""), m.get("author", ""), "v%s" % vs[-1]])
        return out

    def details(self, name):
        vs = self.versions(name)
        if not vs:
            raise KeyError(name)
        with open(os.path.join(self.model_dir(name), vs[-1], "manifest.json")) as f:
            m = json.load(f)
        return {"name": name, "description": m.get("long_description", ""), "versions": ["v" + v for v in vs],
                "manifest": m, "history": self.history(name)}

    def _outbox_mailer(self, email, subject, body):
        box = os.path.join(self.root_dir, "outbox")
        os.makedirs(box, exist_ok=True)
        safe = "".join(c if c.isalnum() or c in "@._-" else "_" for c in email)
        with open(os.path.join(box, safe + ".txt"), "a") as f:
            f.write("Subject: %s\n\n%s\n" % (subject, body))

    def _save_tokens(self):
        with open(self.tokens_file, "w") as f:
            json.dump(self.tokens, f)

    def register(self, email):
        if not email or "@" not in email:
            raise ValueError("a valid e-mail address is required")
        with self.lock:
            if any(t["email"] == email and not t.get("pending") for t in self.tokens.values()):
                raise PermissionError("%s is already registered; unregister first" % email)
            token = str(uuid.uuid4())
            rec = {"email": email, "time": time.time()}
            if self.require_confirmation:
                rec["pending"] = uuid.uuid4().hex[:8]
            self.tokens[token] = rec
            self._save_tokens()
        if self.require_confirmation:
            self.mailer(email, "VelesForge registration",
                        "Confirmation code: %s\n(forge confirm -t %s -c %s)" % (rec["pending"], token, rec["pending"]))
        return token

    def confirm(self, token, code):
        with self.lock:
            rec = self.tokens.get(token)
            if rec is None:
                raise PermissionError("unknown token")
            if rec.get("pending") and rec["pending"] != code:
                raise PermissionError("wrong confirmation code")
            rec.pop("pending", None)
            self._save_tokens()
        return True

    def unregister(self, token):
        """Revoke write access: the token dies; packages it uploaded stay published but can no longer be deleted or
        superseded by it (reference: ``forge_client.py:601-621``)."""
        with self.lock:
            if token not in self.tokens:
                raise PermissionError("unknown token")
            rec = self.tokens.pop(token)
            self._save_tokens()
        return rec["email"]

    def _active(self, token):
        rec = self.tokens.get(token)
        if rec is None:
            raise PermissionError("unknown token")
        if rec.get("pending"):
            raise PermissionError("registration of %s is not confirmed yet" % rec["email"])
        return rec

    def upload(self, token, blob):
        self._active(token)
        tmp = os.path.join(self.root_dir, "incoming-%s" % uuid.uuid4().hex)
        try:
            manifest = validate_manifest(unpack(blob, tmp))
            with self.lock:
                vs = self.versions(manifest["name"])
                if vs:
                    with open(os.path.join(self.model_dir(manifest["name"]), vs[-1], "owner")) as f:
                        if f.read() != self.tokens[token]["email"]:
                            raise PermissionError("%s belongs to another author" % manifest["name"])
                version = str(int(vs[-1]) + 1 if vs else 1)
                dest = os.path.join(self.model_dir(manifest["name"]), version)
                os.makedirs(os.path.dirname(dest), exist_ok=True)
                shutil.move(tmp, dest)
                with open(os.path.join(dest, "package.tar.gz"), "wb") as f:
                    f.write(blob)
                with open(os.path.join(dest, "owner"), "w") as f:
                    f.write(self.tokens[token]["email"])
                self._record_in_git(manifest["name"], version, dest, self.tokens[token]["email"], manifest)
            return {"name": manifest["name"], "version": "v" + version}
        finally:
            shutil.rmtree(tmp, ignore_errors=True)

    def fetch(self, name, version=None):
        vs = self.versions(name)
        if not vs:
            raise KeyError(name)
        v = (version or vs[-1]).lstrip("v")
        with open(os.path.join(self.model_dir(name), v, "package.tar.gz"), "rb") as f:
            return f.read()

    def delete(self, token, name):
        d = self.model_dir(name)
        vs = self.versions(name)
        if not vs:
            raise KeyError(name)
        with open(os.path.join(d, vs[-1], "owner")) as f:
            owner = f.read()
        if self._active(token)["email"] != owner:
            raise PermissionError("only the uploader may delete a model")
        shutil.rmtree(d)

    # -- html ---------------------------------------------------------------------------------------------------
    _PAGE = ("<!doctype html><html><head><meta charset='utf-8'><title>%s</title><style>body{font-family:sans-serif;"
             "margin:1.5em}table{border-collapse:collapse}td,th{border:1px solid #999;padding:4px 10px;text-align:left}"
             "pre{background:#f4f4f4;padding:8px}</style></head><body>%s</body></html>")

    def index_page(self):
        import html
        rows = ["<tr><td><a href='?name=%s'>%s</a></td><td>%s</td><td>%s</td><td>%s</td></tr>"
                % (html.escape(n, quote=True), html.escape(n), html.escape(d), html.escape(a), html.escape(v))
                for n, d, a, v in self.list()]
        body = ("<h2>VelesForge: %d model(s)</h2><table><tr><th>name</th><th>description</th><th>author</th>"
                "<th>latest</th></tr>%s</table><p><code>veles forge fetch -n NAME -s %s</code></p>"
                % (len(rows), "".join(rows), html.escape(self.url)))
        return self._PAGE % ("VelesForge", body)

    def details_page(self, name):
        import html
        d = self.details(name)
        m = d["manifest"]
        hist = "".join("<tr><td>%s</td><td>%s</td><td>%s</td><td>%s</td></tr>"
                       % (html.escape(str(h["version"] or "")), html.escape(h["date"]), html.escape(h["author"]),
                          html.escape(h["subject"])) for h in d["history"])
        links = " ".join("<a href='fetch?name=%s&version=%s'>%s</a>" % (html.escape(name, quote=True), v, v)
                         for v in reversed(d["versions"]))
        body = ("<p><a href='?'>all models</a></p><h2>%s</h2><p>%s</p><p>author: %s; workflow: <code>%s</code>; "
                "configuration: <code>%s</code></p><p>download: %s</p>%s<h3>manifest</h3><pre>%s</pre>"
                % (html.escape(name), html.escape(d["description"]), html.escape(str(m.get("author", ""))),
                   html.escape(str(m.get("workflow", ""))), html.escape(str(m.get("configuration", ""))), links,
                   ("<h3>history</h3><table><tr><th>version</th><th>date</th><th>author</th><th>subject</th></tr>%s"
                    "</table>" % hist) if hist else "", html.escape(json.dumps(m, indent=1))))
        return self._PAGE % (html.escape(name), body)

    # -- http -----------------------------------------------------------------------------------------------------
    def start(self):
        srv = self

        class Handler(BaseHTTPRequestHandler):
            def log_message(self, fmt, *args):
                srv.debug(fmt, *args)

            def _send(self, code, body, ctype="application/json"):
                if not isinstance(body, bytes):
                    body = json.dumps(body).encode()
                self.send_response(code)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def _guard(self, fn):
                try:
                    return fn()
                except KeyError as e:
                    self._send(404, {"error": "not found: %s" % e})
                except PermissionError as e:
                    self._send(403, {"error": str(e)})
                except Exception as e:  # noqa: BLE001
                    self._send(400, {"error": str(e)})

            def do_GET(self):
                url = urlparse(self.path)
                q = {k: v[0] for k, v in parse_qs(url.query).items()}
                if url.path.endswith("/service"):
                    kind = q.get("query")
                    if kind == "list":
                        return self._guard(lambda: self._send(200, srv.list()))
                    if kind == "details":
                        return self._guard(lambda: self._send(200, srv.details(q.get("name"))))
                    if kind == "history":
                        return self._guard(lambda: self._send(200, srv.history(q.get("name"))))
                    if kind == "diff":
                        return self._guard(lambda: self._send(200, {"diff": srv.diff(q.get("name"), q.get("a"),
                                                                                      q.get("b"))}))
                    if kind == "register":
                        return self._guard(lambda: self._send(200, {"token": srv.register(q.get("email", "?"))}))
                    if kind == "confirm":
                        return self._guard(lambda: self._send(200, {"ok": srv.confirm(q.get("token"), q.get("code"))}))
                    if kind == "unregister":
                        return self._guard(lambda: self._send(200, {"email": srv.unregister(q.get("token"))}))
                    if kind == "delete":
                        return self._guard(lambda: (srv.delete(q.get("token"), q.get("name")),
                                                    self._send(200, {"ok": True})))
                    return self._send(400, {"error": "unknown query"})
                if url.path.endswith("/fetch"):
                    return self._guard(lambda: self._send(200, srv.fetch(q.get("name"), q.get("version")),
                                                          "application/x-gzip"))
                if url.path.rstrip("/").endswith("/forge") or url.path in ("", "/"):
                    # the browser face of the zoo (reference: forge_server.py serves forge.html / details pages)
                    name = q.get("name")
                    return self._guard(lambda: self._send(
                        200, (srv.details_page(name) if name else srv.index_page()).encode(),
                        "text/html; charset=utf-8"))
                self._send(404, {"error": "not found"})

            def do_POST(self):
                url = urlparse(self.path)
                q = {k: v[0] for k, v in parse_qs(url.query).items()}
                if url.path.endswith("/upload"):
                    blob = self.rfile.read(int(self.headers.get("Content-Length", 0)))
                    return self._guard(lambda: self._send(200, srv.upload(q.get("token"), blob)))
                self._send(404, {"error": "not found"})

        self.httpd = ThreadingHTTPServer((self.host, self.port), Handler)
        self.port = self.httpd.server_address[1]
        threading.Thread(target=self.httpd.serve_forever, name="forge-server", daemon=True).start()
        return self

    @property
    def url(self):
        return "http://%s:%d/forge" % (self.host, self.port)

    def stop(self):
        if self.httpd is not None:
            self.httpd.shutdown()
            self.httpd.server_close()


def smtp_mailer(host, port=25, sender="forge@localhost", user=None, password=None, tls=False):
    """A ``mailer(email, subject, body)`` that hands confirmation mails to an SMTP relay (the default mailer only
    writes them into the outbox directory)."""
    def send(email, subject, body):
        import smtplib
        from email.message import EmailMessage
        msg = EmailMessage()
        msg["From"], msg["To"], msg["Subject"] = sender, email, subject
        msg.set_content(body)
        with smtplib.SMTP(host, port, timeout=30) as s:
            if tls:
                s.starttls()
            if user:
                s.login(user, password or "")
            s.send_message(msg)
    return send


def main(argv=None):
    """``python -m veles_b200.forge.forge_server --root DIR [--host H] [--port P] [--confirm] [--smtp host[:port]]``:
    the model zoo server as a process of its own (reference: ``veles/forge/forge_server.py`` daemon)."""
    import argparse
    import time as _time
    ap = argparse.ArgumentParser(prog="veles_b200.forge.forge_server", description=main.__doc__)
    ap.add_argument("--root", required=True, help="storage directory (models, tokens, outbox)")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8180)
    ap.add_argument("--confirm", action="store_true", help="new tokens stay pending until the mailed code is presented")
    ap.add_argument("--smtp", default=None, help="SMTP relay host[:port] for the confirmation mails (default: outbox files)")
    ap.add_argument("--sender", default="forge@localhost")
    a = ap.parse_args(argv)
    mailer = None
    if a.smtp:
        h, _, p = a.smtp.partition(":")
        mailer = smtp_mailer(h, int(p or 25), sender=a.sender)
    srv = ForgeServer(a.root, host=a.host, port=a.port, require_confirmation=a.confirm, mailer=mailer).start()
    print("forge server on %s (storage %s)" % (srv.url, srv.root_dir), flush=True)
    try:
        while True:
            _time.sleep(1)
    except KeyboardInterrupt:
        pass
    finally:
        srv.stop()
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
