"""Audio file loading through libsndfile (ctypes), with a pure-Python WAV fallback.

Reference: ``veles/loader/libsndfile.py`` + ``libsndfile_loader.py`` (309 lines of ctypes bindings and the
``SndFileLoaderBase`` mixin: ``open_file``, ``decode_file`` -> int16/float32 samples + sampling rate +
channels, ``file_format_is_supported``).  The shared library is optional on this image, so ``SndFile``
degrades to the standard-library ``wave`` reader for PCM WAV and raises a clear error otherwise.
"""
from __future__ import annotations

import ctypes
import ctypes.util
import wave

import numpy

from .base import LoaderError
from .file_loader import FileLoaderBase

SFM_READ = 0x10


class SF_INFO(ctypes.Structure):
    _fields_ = [("frames", ctypes.c_int64), ("samplerate", ctypes.c_int), ("channels", ctypes.c_int),
                ("format", ctypes.c_int), ("sections", ctypes.c_int), ("seekable", ctypes.c_int)]


class SndFile(object):
    """Thin binding: ``SndFile.read(path) -> (samples float32 [frames, channels], samplerate)``."""
    _lib = None
    _probed = False

    @classmethod
    def lib(cls):
        if not cls._probed:
            cls._probed = True
            name = ctypes.util.find_library("sndfile")
            if name:
                try:
                    L = ctypes.CDLL(name)
                    L.sf_open.restype = ctypes.c_void_p
                    L.sf_open.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(SF_INFO)]
                    L.sf_readf_float.restype = ctypes.c_int64
                    L.sf_readf_float.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_float), ctypes.c_int64]
                    L.sf_close.argtypes = [ctypes.c_void_p]
                    cls._lib = L
                except OSError:
                    cls._lib = None
        return cls._lib

    @classmethod
    def read(cls, path):
        L = cls.lib()
        if L is not None:
            info = SF_INFO()
            h = L.sf_open(path.encode(), SFM_READ, ctypes.byref(info))
            if not h:
                raise LoaderError("libsndfile could not open %s" % path)
            try:
                buf = numpy.zeros((info.frames, info.channels), dtype=numpy.float32)
                got = L.sf_readf_float(h, buf.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), info.frames)
                return buf[:got], info.samplerate
            finally:
                L.sf_close(h)
        try:
            with wave.open(path, "rb") as w:
                n, ch, width = w.getnframes(), w.getnchannels(), w.getsampwidth()
                raw = w.readframes(n)
                rate = w.getframerate()
        except (wave.Error, EOFError) as e:
            raise LoaderError("%s: libsndfile is not installed and the file is not PCM WAV (%s)" % (path, e))
        if width == 2:
            data = numpy.frombuffer(raw, dtype="<i2").astype(numpy.float32) / 32768.0
        elif width == 1:
            data = (numpy.frombuffer(raw, dtype=numpy.uint8).astype(numpy.float32) - 128.0) / 128.0
        elif width == 4:
            data = numpy.frombuffer(raw, dtype="<i4").astype(numpy.float32) / 2147483648.0
        else:
            raise LoaderError("%s: unsupported sample width %d" % (path, width))
        return data.reshape(-1, ch), rate

    @classmethod
    def is_supported(cls, path):
        try:
            cls.read(path)
            return True
        except LoaderError:
            return False


class SndFileLoaderBase(FileLoaderBase):
    """Mixin for file loaders of audio: ``decode_file(path)`` -> dict(data, sampling_rate, channels, name)."""

    def __init__(self, **kwargs):
        kwargs.setdefault("mime_prefix", "audio")
        super(SndFileLoaderBase, self).__init__(**kwargs)

    @staticmethod
    def file_format_is_supported(path):
        return SndFile.is_supported(path)

    @staticmethod
    def decode_file(path):
        data, rate = SndFile.read(path)
        return {"data": data, "sampling_rate": rate, "channels": data.shape[1], "name": path}
