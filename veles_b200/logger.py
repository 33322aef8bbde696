"""Logging mixin with a timeline-event sink.

Parity target: ``veles/logger.py:59-290`` (``Logger`` mixin with colour formatter, file
redirection and ``event(name, "begin"|"end"|"single", **info)`` timeline records which the
reference sends to MongoDB ``veles.events``).  MongoDB is not available on the B200 boxes, so
events go to an in-memory ring plus an optional JSON-lines file; a Mongo handler is attached
only when ``pymongo`` imports.
"""
from __future__ import annotations

import json
import logging
import os
import sys
import threading
import time


class ColorFormatter(logging.Formatter):
    COLORS = {"DEBUG": "\033[36m", "INFO": "\033[92m", "WARNING": "\033[93m", "ERROR": "\033[91m",
              "CRITICAL": "\033[95m"}
    RESET = "\033[0m"

    def __init__(self, fmt=None, use_color=None):
        super().__init__(fmt or "%(asctime)s %(levelname)s:%(name)s:%(message)s", "%Y-%m-%d %H:%M:%S")
        self.use_color = sys.stderr.isatty() if use_color is None else use_color

    def format(self, record):
        msg = super().format(record)
        if self.use_color:
            return self.COLORS.get(record.levelname, "") + msg + self.RESET
        return msg


class EventSink(object):
    """Process-wide sink for timeline events."""

    def __init__(self, capacity=65536):
        self._lock = threading.Lock()
        self.records = []
        self.capacity = capacity
        self._file = None
        self.session = "%d.%d" % (os.getpid(), int(time.time()))

    def open(self, path):
        with self._lock:
            if self._file is not None:
                self._file.close()
            os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
            self._file = open(path, "a")

    def close(self):
        with self._lock:
            if self._file is not None:
                self._file.close()
                self._file = None

    def emit(self, record):
        with self._lock:
            self.records.append(record)
            if len(self.records) > self.capacity:
                del self.records[:len(self.records) - self.capacity]
            if self._file is not None:
                self._file.write(json.dumps(record, default=str) + "\n")
                self._file.flush()


events = EventSink()


class LogSink(logging.Handler):
    """Log records as JSON lines (ring buffer + optional file): what the reference kept in MongoDB
    (``veles/logger.py:226-262``) and the web status queries by session / node / level / text."""

    def __init__(self, capacity=65536):
        logging.Handler.__init__(self)
        self.records = []
        self.capacity = capacity
        self._file = None
        self.node = "%s:%d" % (os.uname().nodename, os.getpid())

    def open(self, path):
        self.acquire()
        try:
            if self._file is not None:
                self._file.close()
            os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
            self._file = open(path, "a")
        finally:
            self.release()

    def emit(self, record):
        try:
            rec = {"session": events.session, "node": self.node, "time": record.created, "level": record.levelname,
                   "levelno": record.levelno, "name": record.name, "message": record.getMessage(),
                   "pathname": record.pathname, "lineno": record.lineno}
            self.records.append(rec)
            if len(self.records) > self.capacity:
                del self.records[:len(self.records) - self.capacity]
            if self._file is not None:
                self._file.write(json.dumps(rec, default=str) + "\n")
                self._file.flush()
        except Exception:  # noqa: BLE001 - logging must never raise
            pass


logs = LogSink()


def install_log_sink(path=None):
    """Route every log record into ``logs`` (idempotent); ``path`` additionally appends JSON lines to a file."""
    rootlog = logging.getLogger()
    if logs not in rootlog.handlers:
        rootlog.addHandler(logs)
    if path:
        logs.open(path)
    return logs


def query_records(records, session=None, node=None, level=None, name=None, domain=None, text=None, since=None,
                  limit=200):
    """Filter log / event records (dicts): the query language of the web status (``veles/web_status.py:113-186``)."""
    levelno = logging.getLevelName(level.upper()) if isinstance(level, str) and level else level
    if not isinstance(levelno, int):
        levelno = None
    out = []
    for r in records:
        if session is not None and r.get("session") != session:
            continue
        if node is not None and r.get("node") != node:
            continue
        if levelno is not None and int(r.get("levelno", 0)) < levelno:
            continue
        if name is not None and r.get("name") != name:
            continue
        if domain is not None and r.get("domain") != domain:
            continue
        if since is not None and float(r.get("time", 0)) < float(since):
            continue
        if text and text.lower() not in (str(r.get("message", "")) + " " + str(r.get("name", ""))).lower():
            continue
        out.append(r)
    return out[-int(limit):] if limit else out


def read_jsonl_dir(directory, suffix):
    """Records of every ``*<suffix>`` file of a directory (other processes' sinks), oldest first."""
    out = []
    if directory and os.path.isdir(directory):
        for fn in sorted(os.listdir(directory)):
            if fn.endswith(suffix):
                try:
                    with open(os.path.join(directory, fn)) as f:
                        for line in f:
                            line = line.strip()
                            if line:
                                try:
                                    out.append(json.loads(line))
                                except ValueError:
                                    pass
                except OSError:
                    pass
    out.sort(key=lambda r: r.get("time", 0))
    return out


class Logger(object):
    """Mixin providing ``self.logger`` and shortcuts, plus ``event()``."""

    _setup_done = False
    _setup_lock = threading.Lock()

    class LoggerHasBeenAlreadySetUp(Exception):
        pass

    def __init__(self, **kwargs):
        self._logger_ = kwargs.get("logger") or logging.getLogger(self.__class__.__name__)
        try:
            super(Logger, self).__init__()
        except TypeError:
            super(Logger, self).__init__(**kwargs)

    # The logger object is never pickled (trailing underscore contract).
    @property
    def logger(self):
        lg = self.__dict__.get("_logger_")
        if lg is None:
            lg = logging.getLogger(self.__class__.__name__)
            self.__dict__["_logger_"] = lg
        return lg

    @logger.setter
    def logger(self, value):
        self.__dict__["_logger_"] = value

    def init_unpickled(self):
        parent = super(Logger, self)
        if hasattr(parent, "init_unpickled"):
            parent.init_unpickled()
        self.__dict__["_logger_"] = logging.getLogger(self.__class__.__name__)

    @staticmethod
    def setup_logging(level=logging.INFO, stream=None):
        with Logger._setup_lock:
            rootlog = logging.getLogger()
            if Logger._setup_done:
                rootlog.setLevel(level)
                return
            Logger._setup_done = True
            handler = logging.StreamHandler(stream or sys.stderr)
            handler.setFormatter(ColorFormatter())
            rootlog.addHandler(handler)
            rootlog.setLevel(level)

    @staticmethod
    def redirect_all_logging_to_file(file_name, max_bytes=1024 * 1024, backups=1):
        from logging.handlers import RotatingFileHandler
        handler = RotatingFileHandler(file_name, maxBytes=max_bytes, backupCount=backups)
        handler.setFormatter(ColorFormatter(use_color=False))
        rootlog = logging.getLogger()
        for h in list(rootlog.handlers):
            if isinstance(h, logging.StreamHandler) and not isinstance(h, logging.FileHandler):
                rootlog.removeHandler(h)
        rootlog.addHandler(handler)
        return handler

    @staticmethod
    def duplicate_all_logging_to_mongo(addr, docid, nodeid):
        try:
            import pymongo  # noqa: F401
        except ImportError:
            logging.getLogger("Logger").warning(
                "pymongo is not installed: logs/events stay in the JSONL sink (asked for %s)", addr)
            return None
        handler = MongoLogHandler(addr, docid, nodeid)
        logging.getLogger().addHandler(handler)
        return handler

    def change_log_message(self, msg):
        return msg

    def log(self, level, msg, *args, **kwargs):
        self.logger.log(level, self.change_log_message(msg), *args, **kwargs)

    def debug(self, msg, *args, **kwargs):
        self.logger.debug(self.change_log_message(msg), *args, **kwargs)

    def info(self, msg, *args, **kwargs):
        self.logger.info(self.change_log_message(msg), *args, **kwargs)

    def warning(self, msg, *args, **kwargs):
        self.logger.warning(self.change_log_message(msg), *args, **kwargs)

    def error(self, msg, *args, **kwargs):
        self.logger.error(self.change_log_message(msg), *args, **kwargs)

    def critical(self, msg, *args, **kwargs):
        self.logger.critical(self.change_log_message(msg), *args, **kwargs)

    def exception(self, msg="Exception", *args, **kwargs):
        self.logger.exception(self.change_log_message(msg), *args, **kwargs)

    def event(self, name, etype, **info):
        """Record a timeline event; ``etype`` is "begin", "end" or "single"."""
        if etype not in ("begin", "end", "single"):
            raise ValueError('Event type must be any of "begin", "end", "single" (got %r)' % etype)
        rec = {"session": events.session, "instance": str(getattr(self, "id", id(self))),
               "time": time.time(), "domain": self.__class__.__name__, "name": name, "type": etype}
        rec.update(info)
        events.emit(rec)
        return rec


class MongoLogHandler(logging.Handler):
    """Optional duplicate of log records into MongoDB (only if pymongo is importable)."""

    def __init__(self, addr, docid, nodeid=None):
        super().__init__()
        import pymongo
        self._client = pymongo.MongoClient("mongodb://" + addr)
        self._collection = self._client.veles.logs
        self.log_id = docid
        self.node_id = nodeid

    def emit(self, record):
        data = dict(record.__dict__)
        for k in ("args", "msg", "exc_info"):
            data.pop(k, None)
        data["message"] = record.getMessage()
        data["session"] = self.log_id
        data["node"] = self.node_id
        try:
            self._collection.insert_one(data)
        except Exception:  # noqa: BLE001 - logging must never raise
            pass
