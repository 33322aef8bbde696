"""File scanning loaders: lists of paths per class, MIME/extension filters, labels from path regex.

Parity target: ``veles/loader/file_loader.py:54-297``: ``FileFilter`` (ignored/included name patterns,
MIME check), ``FileLoaderBase`` (``test_paths`` / ``validation_paths`` / ``train_paths`` scanned
recursively), ``FileListLoaderBase`` (text/json file lists ``path label`` per line), ``AutoLabelFileLoader``
(label = regex group over the path, default: the parent directory name)."""
from __future__ import annotations

import fnmatch
import json
import mimetypes
import os
import re

from .base import TEST, VALID, TRAIN, LoaderError


class FileFilter(object):
    def __init__(self, **kwargs):
        self.ignored_files = list(kwargs.get("ignored_files", []))
        self.included_files = list(kwargs.get("included_files", ["*"]))
        self.filename_types = list(kwargs.get("filename_types", []))   # e.g. ["png", "jpeg"]
        self.mime_prefix = kwargs.get("mime_prefix", None)               # e.g. "image"

    def is_valid_filename(self, filename):
        base = os.path.basename(filename)
        if any(fnmatch.fnmatch(base, pat) for pat in self.ignored_files):
            return False
        if not any(fnmatch.fnmatch(base, pat) for pat in self.included_files):
            return False
        ext = os.path.splitext(base)[1][1:].lower()
        if self.filename_types and ext not in [t.lower() for t in self.filename_types]:
            return False
        if self.mime_prefix:
            mime = mimetypes.guess_type(filename)[0]
            if mime is None or not mime.startswith(self.mime_prefix):
                return False
        return True


class FileLoaderBase(FileFilter):
    """Mixin: scans ``*_paths`` into ``class_keys`` = [test files, validation files, train files]."""

    def __init__(self, **kwargs):
        FileFilter.__init__(self, **kwargs)
        self.test_paths = list(kwargs.get("test_paths", []))
        self.validation_paths = list(kwargs.get("validation_paths", []))
        self.train_paths = list(kwargs.get("train_paths", []))
        self.class_keys = [[], [], []]

    def scan_files(self, pathname):
        files = []
        if os.path.isfile(pathname):
            return [pathname] if self.is_valid_filename(pathname) else []
        if not os.path.isdir(pathname):
            raise LoaderError("Path %s does not exist" % pathname)
        for basedir, _, names in sorted(os.walk(pathname, followlinks=True)):
            for name in sorted(names):
                full = os.path.join(basedir, name)
                if self.is_valid_filename(full):
                    files.append(full)
        if not files:
            raise LoaderError("No files were found in %s" % pathname)
        return files

    def collect_keys(self):
        for ci, paths in ((TEST, self.test_paths), (VALID, self.validation_paths), (TRAIN, self.train_paths)):
            keys = []
            for p in paths:
                keys.extend(self.scan_files(p))
            self.class_keys[ci] = keys
        return self.class_keys


class FileListLoaderBase(FileFilter):
    """Mixin: ``path_to_*_text_file`` list files with ``<path> [label]`` lines (or a json {path: label})."""

    def __init__(self, **kwargs):
        FileFilter.__init__(self, **kwargs)
        self.path_to_test_text_file = kwargs.get("path_to_test_text_file", "")
        self.path_to_val_text_file = kwargs.get("path_to_val_text_file", "")
        self.path_to_train_text_file = kwargs.get("path_to_train_text_file", "")
        self.base_directory = kwargs.get("base_directory", "")
        self.class_keys = [[], [], []]
        self.labels_by_key = {}

    def _read_list(self, fname):
        if not fname:
            return []
        keys = []
        if fname.endswith(".json"):
            with open(fname) as f:
                for path, label in json.load(f).items():
                    full = os.path.join(self.base_directory, path)
                    keys.append(full)
                    self.labels_by_key[full] = label
            return keys
        with open(fname) as f:
            for line in f:
                line = line.strip()
                if not line or line.startswith("#"):
                    continue
                parts = line.rsplit(None, 1)
                full = os.path.join(self.base_directory, parts[0])
                keys.append(full)
                if len(parts) > 1:
                    self.labels_by_key[full] = parts[1]
        return keys

    def collect_keys(self):
        self.class_keys = [self._read_list(self.path_to_test_text_file),
                           self._read_list(self.path_to_val_text_file),
                           self._read_list(self.path_to_train_text_file)]
        return self.class_keys

    def get_label_from_filename(self, filename):
        return self.labels_by_key.get(filename)


class AutoLabelFileLoader(FileLoaderBase):
    """Mixin: label = first group of ``label_regexp`` over the path (default: parent directory name)."""

    def __init__(self, **kwargs):
        FileLoaderBase.__init__(self, **kwargs)
        self.label_regexp = re.compile(kwargs.get("label_regexp", r".*%s([^%s]+)%s[^%s]+$" % ((re.escape(os.sep),) * 4)))

    def get_label_from_filename(self, filename):
        m = self.label_regexp.search(filename)
        if m is None:
            raise LoaderError("%s does not match the label regular expression" % filename)
        return m.group(1)
