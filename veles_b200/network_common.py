"""Shared pieces of the job-control plane: machine id, payload framing/compression, shared-memory
transport for same-host payloads.

Parity targets: ``veles/network_common.py:49-154`` (``mid`` = machine-id + MAC, used to pick the
cheapest endpoint), ``veles/txzmq/connection.py:483-562`` (pickle streamed into frames between
``vpb``/``vpe`` markers with optional gzip/snappy/xz), ``veles/txzmq/sharedio.py:45-105`` (POSIX
shm file-like object so same-host job payloads bypass the socket).
"""
from __future__ import annotations

import bz2
import gzip
import lzma
import os
import pickle
import socket
import uuid
import zlib
from multiprocessing import shared_memory

from .config import root, get
from .pickle2 import best_protocol

CODECS = {
    None: (lambda b: b, lambda b: b),
    "": (lambda b: b, lambda b: b),
    "none": (lambda b: b, lambda b: b),
    "gzip": (lambda b: zlib.compress(b, 1), zlib.decompress),
    "bz2": (lambda b: bz2.compress(b, 1), bz2.decompress),
    "xz": (lambda b: lzma.compress(b, preset=0), lzma.decompress),
}
try:  # optional, like in the reference
    import snappy as _snappy
    CODECS["snappy"] = (_snappy.compress, _snappy.decompress)
except ImportError:
    pass

MARK_BEGIN = b"vpb"
MARK_END = b"vpe"


def machine_id():
    """Stable id of this host: /etc/machine-id (or dbus) + MAC address."""
    mid = ""
    for path in ("/etc/machine-id", "/var/lib/dbus/machine-id"):
        try:
            with open(path) as f:
                mid = f.read().strip()
                break
        except OSError:
            continue
    return "%s-%012x" % (mid or socket.gethostname(), uuid.getnode())


def dumps(obj, codec=None):
    """[b'vpb'+codec, payload, b'vpe'] frames for one pickled object."""
    if codec not in CODECS:
        codec = "gzip"
    comp, _ = CODECS[codec]
    return [MARK_BEGIN + (codec or "none").encode(), comp(pickle.dumps(obj, protocol=best_protocol)), MARK_END]


def loads(frames):
    if len(frames) < 3 or not frames[0].startswith(MARK_BEGIN) or frames[-1] != MARK_END:
        raise ValueError("Malformed payload framing")
    codec = frames[0][len(MARK_BEGIN):].decode() or None
    _, decomp = CODECS[codec if codec in CODECS else None]
    return pickle.loads(decomp(b"".join(frames[1:-1])))





def parse_address(address, default_host="127.0.0.1"):
    if address.startswith(("tcp://", "ipc://", "inproc://")):
        return address
    host, _, port = address.rpartition(":")
    return "tcp://%s:%s" % (host or default_host, port)


def default_codec():
    return get(root.common.engine.network_compression, "gzip")


# SharedIO lives with the ZeroMQ transport (txzmq/sharedio.py); imported last: txzmq.connection needs the codecs above
from .txzmq.sharedio import SharedIO  # noqa: E402,F401 - the class lives with the ZeroMQ transport
