"""Downloader unit: fetches and unpacks a dataset archive at initialize time unless the files are already there
(reference: ``veles/downloader.py:56-125``).  ``file://`` URLs and local paths work offline."""
from __future__ import annotations

import os
import shutil
import tarfile
import urllib.request
import zipfile

from .config import root, get
from .distributable import TriviallyDistributable
from .units import Unit


class Downloader(Unit, TriviallyDistributable):
    """Fetches and unpacks a dataset archive (URL or local path) into a directory unless the listed files are
    already there."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(Downloader, self).__init__(workflow, **kwargs)
        self.url = kwargs["url"]
        self.directory = kwargs.get("directory", get(root.common.dirs.datasets, "."))
        self.files = list(kwargs.get("files", []))

    def initialize(self, **kwargs):
        if self.files and all(os.path.exists(os.path.join(self.directory, f)) for f in self.files):
            self.debug("All the files are already present in %s", self.directory)
            return
        os.makedirs(self.directory, exist_ok=True)
        local = os.path.join(self.directory, os.path.basename(self.url.split("?")[0]) or "download")
        if os.path.exists(self.url):
            shutil.copy(self.url, local)
        else:
            self.info("Downloading %s...", self.url)
            urllib.request.urlretrieve(self.url, local)
        if zipfile.is_zipfile(local):
            with zipfile.ZipFile(local) as z:
                z.extractall(self.directory)
        elif tarfile.is_tarfile(local):
            with tarfile.open(local) as t:
                t.extractall(self.directory, filter="data")
        missing = [f for f in self.files if not os.path.exists(os.path.join(self.directory, f))]
        if missing:
            raise FileNotFoundError("The archive %s does not contain %s" % (self.url, missing))

    def run(self):
        pass
