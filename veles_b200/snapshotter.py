"""Snapshotter: pickles the whole workflow object graph.

Parity target: ``veles/snapshotter.py:84-535``: runs every ``interval`` calls and at most once
per ``time_interval`` seconds (:159-174); file name
``<prefix>_<suffix>.<protocol>.pickle[.gz|.bz2|.xz]`` plus a ``<prefix>_current...`` symlink
(:387-409); codecs (:365-380); ``import_`` marks the object as restored (:411-420); >200 MB
size warning with a per-unit breakdown (:203-225); master waits for slaves before snapshotting.
``snappy`` is optional (not installed here) and ``SnapshotterToDB`` needs ``pyodbc``: both are
gated.  In data-parallel runs only rank 0 writes (all ranks hold identical weights after the
fused all-reduce+GD step).
"""
from __future__ import annotations

import bz2
import gzip
import lzma
import os
import pickle
import time

from .config import root, get
from .distributable import TriviallyDistributable
from .mapped_object_registry import MappedObjectsRegistry
from .pickle2 import best_protocol
from .unit_registry import MappedUnitRegistry
from .units import Unit


class SnapshotterRegistry(MappedUnitRegistry):
    mapping = "snapshotters"
    snapshotters = {}


def _snappy():
    try:
        import snappy
        return snappy
    except ImportError:
        return None


class SnapshotterBase(Unit, TriviallyDistributable, metaclass=SnapshotterRegistry):
    hide_from_registry = True
    VIEW_GROUP = "SERVICE"
    SIZE_WARNING_THRESHOLD = 200 * 1000 * 1000
    CODECS = {"": lambda f, lvl: open(f, "wb"),
              "gz": lambda f, lvl: gzip.GzipFile(f, "wb", compresslevel=lvl),
              "bz2": lambda f, lvl: bz2.BZ2File(f, "wb", compresslevel=lvl),
              "xz": lambda f, lvl: lzma.LZMAFile(f, "wb", preset=lvl)}
    READ_CODECS = {"pickle": lambda f: open(f, "rb"), "gz": lambda f: gzip.GzipFile(f, "rb"),
                   "bz2": lambda f: bz2.BZ2File(f, "rb"), "xz": lambda f: lzma.LZMAFile(f, "rb")}

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(SnapshotterBase, self).__init__(workflow, **kwargs)
        self.verbose = kwargs.get("verbose", True)
        self.prefix = kwargs.get("prefix", "")
        # members of an ensemble train side by side (one per GPU) and can reach the same metric in the same epoch:
        # their files are told apart by the model index the manager passes (root.common.ensemble.model_index)
        model_index = kwargs.get("model_index", get(root.common.ensemble.model_index, None))
        if model_index is not None:
            self.prefix = ("%04d_" % int(model_index)) + self.prefix
        self.compression = kwargs.get("compression", "gz")
        self.compression_level = kwargs.get("compression_level", 6)
        self.interval = kwargs.get("interval", 1)
        self.time_interval = kwargs.get("time_interval", 15)
        self.time = 0
        self._skipped_counter = 0
        self.skip = kwargs.get("skip", False)
        self.suffix = None
        self.destination = ""
        self._slaves_ = {}

    def init_unpickled(self):
        super(SnapshotterBase, self).init_unpickled()
        self._slaves_ = {}

    def initialize(self, **kwargs):
        self.time = time.time()
        self.debug("Compression is set to %s, interval %d, time interval %s s",
                   self.compression, self.interval, self.time_interval)

    def run(self):
        if self.is_slave or get(root.common.disable.snapshotting, False) or self.skip:
            return
        dp = getattr(self.workflow, "dp_group_", None)
        if dp is not None and getattr(dp, "rank", 0) != 0:
            return  # data-parallel group: identical state everywhere, rank 0 writes
        self._skipped_counter += 1
        if self._skipped_counter < self.interval:
            return
        delta = time.time() - self.time
        if delta < self.time_interval:
            self.debug("%f < %f, dropped", delta, self.time_interval)
            return
        self._skipped_counter = 0
        self.export()
        self.time = time.time()

    def export(self):
        raise NotImplementedError

    def check_snapshot_size(self, size):
        if size > self.SIZE_WARNING_THRESHOLD:
            self.warning("Snapshot size looks too big: %d bytes", size)
            sizes = []
            for u in self.workflow.units:
                try:
                    sizes.append((len(pickle.dumps(u, protocol=best_protocol)), u.name))
                except Exception:  # noqa: BLE001
                    pass
            for sz, name in sorted(sizes, reverse=True)[:5]:
                self.warning("  %-32s %d bytes (with everything it references)", name, sz)

    def get_metric_names(self):
        return {"Snapshot"}

    def get_metric_values(self):
        return {"Snapshot": self.destination}

    def _slave_ended(self, slave):
        self._slaves_.pop(getattr(slave, "id", slave), None)


class SnapshotterToFile(SnapshotterBase):
    """Writes snapshots to ``directory``."""
    MAPPING = "file"
    hide_from_registry = False
    WRITE_CODECS = SnapshotterBase.CODECS

    def __init__(self, workflow, **kwargs):
        super(SnapshotterToFile, self).__init__(workflow, **kwargs)
        self.directory = kwargs.get("directory", get(root.common.dirs.snapshots, "."))

    def export(self):
        ext = ("." + self.compression) if self.compression else ""
        rel = "%s_%s.%d.pickle%s" % (self.prefix, self.suffix, best_protocol, ext)
        os.makedirs(self.directory, exist_ok=True)
        self.destination = os.path.abspath(os.path.join(self.directory, rel))
        t0 = time.time()
        # written under a temporary name and renamed: a crash (or a failing pickle) in the middle of a snapshot must
        # not leave a truncated file that a restart would then pick up as "the newest snapshot"
        tmp = "%s.%d.part" % (self.destination, os.getpid())
        try:
            if self.compression == "snappy":
                sn = _snappy()
                if sn is None:
                    raise RuntimeError("snappy compression requested but python-snappy is not installed")
                payload = pickle.dumps(self.workflow, protocol=best_protocol)
                with open(tmp, "wb") as f:
                    f.write(sn.compress(payload))
            else:
                with self.WRITE_CODECS[self.compression](tmp, self.compression_level) as fout:
                    pickle.dump(self.workflow, fout, protocol=best_protocol)
            os.replace(tmp, self.destination)
        finally:
            if os.path.exists(tmp):
                os.remove(tmp)
        size = os.path.getsize(self.destination)
        self.check_snapshot_size(size)
        if self.verbose:
            self.info("Snapshotted to %s (%d bytes, %.2f s)", self.destination, size, time.time() - t0)
        link = os.path.join(self.directory, "%s_current.%d.pickle%s" % (self.prefix, best_protocol, ext))
        try:
            if os.path.lexists(link):
                os.remove(link)
            os.symlink(os.path.basename(self.destination), link)
        except OSError:
            pass

    @staticmethod
    def import_(file_name):
        """Load a workflow snapshot; the result has ``restored_from_snapshot == True``."""
        file_name = file_name.strip()
        if not os.path.exists(file_name):
            raise FileNotFoundError(file_name)
        ext = os.path.splitext(file_name)[1][1:]
        if ext == "snappy":
            sn = _snappy()
            if sn is None:
                raise RuntimeError("python-snappy is not installed")
            with open(file_name, "rb") as f:
                return SnapshotterToFile._mark(pickle.loads(sn.decompress(f.read())))
        opener = SnapshotterBase.READ_CODECS.get(ext, SnapshotterBase.READ_CODECS["pickle"])
        with opener(file_name) as fin:
            return SnapshotterToFile._mark(pickle.load(fin))

    @staticmethod
    def _mark(obj):
        try:
            obj._restored_from_snapshot_ = True
        except AttributeError:
            pass
        return obj


class SnapshotterToDB(SnapshotterBase):
    """Snapshots as blobs in a database table (reference ``snapshotter.py:428-518``, ODBC only there).

    ``odbc`` is an ODBC connection string (needs ``pyodbc``) or ``sqlite:<path>`` for the DB-API module that ships
    with Python, which creates the table on first use.  Columns: timestamp, id, log_id, workflow, name, codec, data.
    ``SnapshotterToDB.import_(odbc, table, id_, log_id, name=None)`` restores the named (or the newest) snapshot.
    """
    MAPPING = "odbc"
    hide_from_registry = False
    CODECS = {None: (lambda b, l: b, lambda b: b, "pickle"), "": (lambda b, l: b, lambda b: b, "pickle"),
              "gz": (lambda b, l: gzip.compress(b, l), gzip.decompress, "gz"),
              "bz2": (lambda b, l: __import__("bz2").compress(b, l), lambda b: __import__("bz2").decompress(b), "bz2"),
              "xz": (lambda b, l: __import__("lzma").compress(b, preset=l), lambda b: __import__("lzma").decompress(b),
                     "xz")}

    def __init__(self, workflow, **kwargs):
        super(SnapshotterToDB, self).__init__(workflow, **kwargs)
        self.odbc = kwargs.get("odbc")
        self.table = kwargs.get("table", "veles")
        if self.compression not in self.CODECS:
            raise ValueError("SnapshotterToDB: unsupported compression %r" % (self.compression,))

    @staticmethod
    def connect(dsn, table=None):
        if dsn is None:
            raise ValueError("SnapshotterToDB needs odbc=<connection string>")
        if str(dsn).startswith("sqlite:"):
            import sqlite3
            conn = sqlite3.connect(str(dsn)[len("sqlite:"):])
            if table is not None:
                conn.execute("create table if not exists %s(timestamp text, id text, log_id text, workflow text, "
                             "name text, codec text, data blob)" % table)
            return conn
        try:
            import pyodbc
        except ImportError:
            raise RuntimeError("SnapshotterToDB: %r needs pyodbc, which is not installed (sqlite:<path> works "
                               "without it)" % dsn)
        return pyodbc.connect(dsn)

    def initialize(self, **kwargs):
        super(SnapshotterToDB, self).initialize(**kwargs)
        self.connect(self.odbc, self.table).close()      # fail now, not at the first snapshot

    def export(self):
        enc, _, codec = self.CODECS[self.compression]
        payload = enc(pickle.dumps(self.workflow, protocol=best_protocol), self.compression_level)
        name = ".".join((self.prefix, self.suffix or "", str(best_protocol)))
        conn = self.connect(self.odbc, self.table)
        try:
            cur = conn.cursor()
            now = time.strftime("%Y-%m-%d %H:%M:%S") + (".%06d" % int((time.time() % 1) * 1e6))
            cur.execute("insert into %s(timestamp, id, log_id, workflow, name, codec, data) "
                        "values (?, ?, ?, ?, ?, ?, ?);" % self.table,
                        (now, str(getattr(self.launcher, "id", "")), str(getattr(self.launcher, "log_id", "")),
                         str(self.workflow.name), name, codec, payload))
            conn.commit()
            self.destination = name
            self.info("Wrote %d bytes as %s into table %s", len(payload), name, self.table)
        finally:
            conn.close()

    @staticmethod
    def import_(odbc, table, id_, log_id, name=None):
        conn = SnapshotterToDB.connect(odbc)
        try:
            cur = conn.cursor()
            query = "select codec, data from %s where id=? and log_id=?" % table
            args = [str(id_), str(log_id)]
            if name is not None:
                query += " and name=?"
                args.append(name)
            query += " order by timestamp desc"
            cur.execute(query, args)
            row = cur.fetchone()
        finally:
            conn.close()
        if row is None:
            raise KeyError("no snapshot id=%s log_id=%s name=%s in %s" % (id_, log_id, name, table))
        decoders = {c[2]: c[1] for c in SnapshotterToDB.CODECS.values()}
        wf = pickle.loads(decoders[row[0]](bytes(row[1])))
        try:
            wf._restored_from_snapshot_ = True
        except AttributeError:
            pass
        return wf

    def get_metric_values(self):
        return {"Snapshot": {"odbc": self.odbc, "table": self.table, "name": self.destination}}


class Snapshotter(SnapshotterToFile):
    """Default snapshotter: files, or the database variant when constructed with ``odbc=...`` (reference :522-535)."""
    MAPPING = "default"

    def __new__(cls, workflow=None, **kwargs):
        if cls is Snapshotter and kwargs.get("odbc"):
            return SnapshotterToDB(workflow, **kwargs)
        return super(Snapshotter, cls).__new__(cls)

    def __init__(self, workflow, **kwargs):
        kwargs.get("odbc"), kwargs.get("table")      # accepted: they select the database variant in __new__
        super(Snapshotter, self).__init__(workflow, **kwargs)

    @staticmethod
    def import_file(file_name):
        return SnapshotterToFile.import_(file_name)

    @staticmethod
    def import_odbc(odbc, table, id_, log_id, name=None):
        return SnapshotterToDB.import_(odbc, table, id_, log_id, name)
