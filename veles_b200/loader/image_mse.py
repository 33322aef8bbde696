"""Image loaders with regression targets: every label names a TARGET image (``target_paths``), the
minibatch target is that image run through the same pipeline without augmentation.

Reference: ``veles/loader/image_mse.py`` (``ImageLoaderMSEMixin``, ``file_image_mse`` /
``auto_label_file_image_mse`` registrations).
"""
from __future__ import annotations

import os

import numpy

from .base import LoaderError, LoaderMSEMixin
from .file_image import AutoLabelFileImageLoader, FileListImageLoader, ImageLoader
from .file_loader import FileFilter


class ImageLoaderMSEMixin(LoaderMSEMixin):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(ImageLoaderMSEMixin, self).__init__(workflow, **kwargs)
        self.target_paths = list(kwargs.get("target_paths", ()))
        self.target_by_label_ = {}

    def load_data(self):
        super(ImageLoaderMSEMixin, self).load_data()
        if not self.target_paths:
            raise LoaderError("target_paths is required by MSE image loaders")
        flt = FileFilter(mime_prefix="image")
        self.target_by_label_ = {}
        for root in self.target_paths:
            files = [root] if os.path.isfile(root) else sorted(
                os.path.join(d, f) for d, _, fs in os.walk(root) for f in fs)
            for path in files:
                if not flt.is_valid_filename(path):
                    continue
                label = self.get_label_from_filename(path)
                img = self.pipeline.decode(path)
                self.target_by_label_[label] = next(iter(self.pipeline.process(img, training=False)))
        missing = set(l for l in self.key_labels if l not in self.target_by_label_)
        if missing:
            raise LoaderError("No target image for labels %s" % sorted(missing)[:5])
        self._targets_shape = tuple(next(iter(self.target_by_label_.values())).shape)

    def create_minibatch_data(self):
        super(ImageLoaderMSEMixin, self).create_minibatch_data()
        self._minibatch_targets.reset(numpy.zeros((self._max_minibatch_size,) + self._targets_shape,
                                                  dtype=numpy.float32))

    def fill_minibatch(self):
        super(ImageLoaderMSEMixin, self).fill_minibatch()
        infl = self.pipeline.samples_inflation
        for i, idx in enumerate(self.minibatch_indices.mem[:self._minibatch_size_]):
            self._minibatch_targets.mem[i] = self.target_by_label_[self.key_labels[int(idx) // infl]]


class FileListImageLoaderMSE(ImageLoaderMSEMixin, FileListImageLoader):
    """FileListImageLoader with target images."""
    MAPPING = "file_list_image_mse"
    hide_from_registry = False


class AutoLabelFileImageLoaderMSE(ImageLoaderMSEMixin, AutoLabelFileImageLoader):
    """AutoLabelFileImageLoader with per-class target images."""
    MAPPING = "auto_label_file_image_mse"
    hide_from_registry = False
