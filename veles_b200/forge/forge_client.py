"""Forge client: ``veles_b200 forge <action>`` with actions list | details | history | diff | fetch | upload | delete | register |
confirm | unregister | assist (reference: ``veles/forge/forge_client.py:91-799``)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import urllib.parse
import urllib.request

from ..config import root, get
from ..logger import Logger
from .forge_common import pack, unpack


class ForgeClient(Logger):
    def __init__(self, server=None, token=None):
        super(ForgeClient, self).__init__()
        self.server = (server or get(root.common.forge.server, "http://127.0.0.1:8080/forge")).rstrip("/")
        self.token = token or os.environ.get("VELES_FORGE_TOKEN")

    def _get(self, path, **query):
        url = "%s/%s?%s" % (self.server, path, urllib.parse.urlencode({k: v for k, v in query.items() if v is not None}))
        with urllib.request.urlopen(url, timeout=30) as r:
            return r.read()

    def list(self):
        return json.loads(self._get("service", query="list"))

    def details(self, name):
        return json.loads(self._get("service", query="details", name=name))

    def history(self, name):
        """Upload history of a model from the server's git repository: [{version, commit, date, author, subject}]."""
        return json.loads(self._get("service", query="history", name=name))

    def diff(self, name, a, b):
        """File-level summary of what changed between two versions (``git diff --stat``)."""
        return json.loads(self._get("service", query="diff", name=name, a=a, b=b))["diff"]

    def register(self, email):
        self.token = json.loads(self._get("service", query="register", email=email))["token"]
        return self.token

    def confirm(self, code):
        return json.loads(self._get("service", query="confirm", token=self.token, code=code))["ok"]

    def unregister(self):
        """Revoke this token's write access (the packages it uploaded stay published)."""
        res = json.loads(self._get("service", query="unregister", token=self.token))
        self.token = None
        return res["email"]

    def assist(self, workflow_path, answers=None, ask=input):
        """Create ``<name>/manifest.json`` next to a copy of the workflow (and its ``*_config.py``) so that the
        directory can be uploaded (reference: ``forge_client.py:514-599``).  ``answers`` pre-fills fields; whatever is
        missing is asked interactively through ``ask``."""
        import shutil
        answers = dict(answers or {})
        workflow_path = os.path.abspath(workflow_path)
        if not os.path.isfile(workflow_path):
            raise FileNotFoundError(workflow_path)
        base, modname = os.path.dirname(workflow_path), os.path.splitext(os.path.basename(workflow_path))[0]
        # the Workflow subclass defined in the file names the package by default
        default_name = modname
        try:
            import ast
            with open(workflow_path) as f:
                tree = ast.parse(f.read())
            classes = [n.name for n in tree.body if isinstance(n, ast.ClassDef)
                       and any("Workflow" in ast.unparse(b) for b in n.bases)]
            if classes:
                default_name = classes[-1]
        except SyntaxError:
            pass
        fields = (("name", default_name), ("short_description", ""), ("long_description", ""),
                  ("author", os.environ.get("USER", "")))
        meta = {}
        for key, default in fields:
            val = answers.get(key)
            if val is None:
                val = ask("%s [%s]: " % (key.replace("_", " "), default)).strip() or default
            meta[key] = val
        cfg_src = answers.get("configuration") or os.path.join(base, modname + "_config.py")
        dest = os.path.join(answers.get("directory", "."), meta["name"])
        os.makedirs(dest, exist_ok=True)
        shutil.copy(workflow_path, os.path.join(dest, os.path.basename(workflow_path)))
        cfg_name = os.path.basename(cfg_src)
        if os.path.isfile(cfg_src):
            shutil.copy(cfg_src, os.path.join(dest, cfg_name))
        else:
            open(os.path.join(dest, cfg_name), "w").close()      # an empty configuration, as the reference does
        meta.update(workflow=os.path.basename(workflow_path), configuration=cfg_name,
                    requires=answers.get("requires", ["veles_b200"]))
        with open(os.path.join(dest, "manifest.json"), "w") as f:
            json.dump(meta, f, sort_keys=True, indent=4)
        return dest, meta

    def fetch(self, name, directory=".", version=None):
        blob = self._get("fetch", name=name, version=version)
        dest = os.path.join(directory, name)
        manifest = unpack(blob, dest)
        self.info("Fetched %s into %s", name, dest)
        return dest, manifest

    def upload(self, directory):
        if not self.token:
            raise PermissionError("upload needs a token: run `forge register -e you@example.com` first")
        manifest, blob = pack(directory)
        req = urllib.request.Request("%s/upload?%s" % (self.server, urllib.parse.urlencode({"token": self.token})),
                                     data=blob, headers={"Content-Type": "application/x-gzip"})
        with urllib.request.urlopen(req, timeout=60) as r:
            return json.loads(r.read())

    def delete(self, name):
        return json.loads(self._get("service", query="delete", name=name, token=self.token))


def main(argv=None):
    p = argparse.ArgumentParser(prog="veles_b200 forge")
    p.add_argument("action", choices=["list", "details", "fetch", "upload", "delete", "register", "confirm", "unregister",
                                      "assist", "history", "diff"])
    p.add_argument("--against", default=None, help="the other version of a diff (diff -n NAME -v v2 --against v1)")
    p.add_argument("-c", "--code", default=None, help="confirmation code (confirm)")
    p.add_argument("-w", "--workflow", default=None, help="path to the workflow file (assist)")
    p.add_argument("-s", "--server", default=None)
    p.add_argument("-n", "--name", default=None)
    p.add_argument("-d", "--directory", default=".")
    p.add_argument("-v", "--version", default=None)
    p.add_argument("-e", "--email", default=None)
    p.add_argument("-t", "--token", default=None)
    a = p.parse_args(argv)
    c = ForgeClient(a.server, a.token)
    if a.action == "list":
        for row in c.list():
            print("%-24s %-40s %-16s %s" % tuple(row))
    elif a.action == "details":
        print(json.dumps(c.details(a.name), indent=1))
    elif a.action == "history":
        for h in c.history(a.name):
            print("%-6s %s  %-32s %s" % (h["version"] or "", h["date"], h["author"], h["subject"]))
    elif a.action == "diff":
        print(c.diff(a.name, a.against, a.version))
    elif a.action == "fetch":
        print(c.fetch(a.name, a.directory, a.version)[0])
    elif a.action == "upload":
        print(json.dumps(c.upload(a.directory)))
    elif a.action == "delete":
        print(json.dumps(c.delete(a.name)))
    elif a.action == "register":
        print(c.register(a.email))
    elif a.action == "confirm":
        print(c.confirm(a.code))
    elif a.action == "unregister":
        print(c.unregister())
    elif a.action == "assist":
        print(c.assist(a.workflow, {"name": a.name, "directory": a.directory})[0])
    return 0


if __name__ == "__main__":
    sys.exit(main())
