"""Text-line loader over HDFS (reference: ``veles/loader/hdfs_loader.py:41-71``, which wrapped the ``hdfs``
WebHDFS client).  Here the transport is pluggable: ``pyarrow.fs.HadoopFileSystem`` when libhdfs is present,
any object with ``open(path) -> file`` (``filesystem=`` kwarg), or a local path for tests.  Each minibatch
is ``chunk_lines_number`` raw lines handed to the next unit through ``output``.
"""
from __future__ import annotations

from ..units import Unit
from .base import LoaderError


class HDFSTextLoader(Unit):
    """Streams a (remote) text file in chunks of lines; sets ``finished`` when the file ends."""

    def __init__(self, workflow, **kwargs):
        super(HDFSTextLoader, self).__init__(workflow, **kwargs)
        self.file_name = kwargs["file"]
        self.chunk_lines_number = int(kwargs.get("chunk", 1000))
        self.address = kwargs.get("address", None)      # "host:port" of the name node
        self.user = kwargs.get("user", None)
        self.filesystem = kwargs.get("filesystem", None)
        self.output = [""] * self.chunk_lines_number
        from ..mutable import Bool
        self.finished = Bool(False)

    def init_unpickled(self):
        super(HDFSTextLoader, self).init_unpickled()
        self.stream_ = None

    def _open(self):
        if self.filesystem is not None:
            return self.filesystem.open(self.file_name)
        if self.address:
            try:
                from pyarrow import fs
                host, _, port = self.address.partition(":")
                hdfs = fs.HadoopFileSystem(host, int(port or 8020), user=self.user)
                return hdfs.open_input_stream(self.file_name)
            except Exception as e:   # libhdfs / JVM not installed
                raise LoaderError("HDFS is not reachable from this installation: %s" % e)
        return open(self.file_name, "rb")

    def initialize(self, **kwargs):
        self.stream_ = self._open()
        self.finished <<= False
        return None

    def run(self):
        n = 0
        for n in range(self.chunk_lines_number):
            line = self.stream_.readline()
            if not line:
                self.finished <<= True
                self.output[n:] = [""] * (self.chunk_lines_number - n)
                return
            self.output[n] = line.decode("utf-8", "replace").rstrip("\n") if isinstance(line, bytes) else line.rstrip("\n")

    def stop(self):
        if self.stream_ is not None:
            self.stream_.close()
            self.stream_ = None
