"""Forge client: ``veles_b200 forge <action>`` with actions list | details | fetch | upload | delete | register
(reference: ``veles/forge/forge_client.py:91-799``)."""
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

    def register(self, email):
        self.token = json.loads(self._get("service", query="register", email=email))["token"]
        return self.token

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
    p.add_argument("action", choices=["list", "details", "fetch", "upload", "delete", "register"])
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
    elif a.action == "fetch":
        print(c.fetch(a.name, a.directory, a.version)[0])
    elif a.action == "upload":
        print(json.dumps(c.upload(a.directory)))
    elif a.action == "delete":
        print(json.dumps(c.delete(a.name)))
    elif a.action == "register":
        print(c.register(a.email))
    return 0


if __name__ == "__main__":
    sys.exit(main())
