"""Streaming image loaders reading files on demand: ``file_list_image``, ``auto_label_file_image``
(reference: ``veles/loader/file_image.py``, ``veles/loader/image.py`` ImageLoader)."""
from __future__ import annotations

import numpy

from .base import Loader, LoaderError, TEST, VALID, TRAIN
from .file_loader import FileListLoaderBase, AutoLabelFileLoader
from .image import ImagePipeline


class ImageLoader(Loader):
    """Loader decoding + preprocessing images per minibatch (host path, uploaded each step)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(ImageLoader, self).__init__(workflow, **kwargs)
        self.pipeline = ImagePipeline(prng=self.prng, **kwargs)
        self.keys = []
        self.key_labels = []

    def collect_keys(self):
        raise NotImplementedError

    def get_label_from_filename(self, filename):
        raise NotImplementedError

    def load_data(self):
        class_keys = self.collect_keys()
        infl = self.pipeline.samples_inflation
        self.keys, self.key_labels = [], []
        for ci in (TEST, VALID, TRAIN):
            for k in class_keys[ci]:
                self.keys.append(k)
                self.key_labels.append(self.get_label_from_filename(k))
            self.class_lengths[ci] = len(class_keys[ci]) * infl
        if not self.keys:
            raise LoaderError("No images were found")
        self._has_labels = any(l is not None for l in self.key_labels)
        if self._has_labels:
            diff = [{}, {}, {}]
            pos = 0
            for ci in (TEST, VALID, TRAIN):
                n = len(class_keys[ci])
                for l in self.key_labels[pos:pos + n]:
                    diff[ci][l] = diff[ci].get(l, 0) + 1
                pos += n
            self._setup_labels_mapping(diff)
        self._shape_ = self.pipeline.output_shape(self.pipeline.decode(self.keys[0]))

    def create_minibatch_data(self):
        self.minibatch_data.reset(numpy.zeros((self._max_minibatch_size,) + tuple(self._shape_), dtype=numpy.float32))

    def fill_minibatch(self):
        infl = self.pipeline.samples_inflation
        training = self.minibatch_class == TRAIN
        for i, idx in enumerate(self.minibatch_indices.mem[:self._minibatch_size_]):
            key_index, variant = divmod(int(idx), infl)
            img = self.pipeline.decode(self.keys[key_index])
            for v, out in enumerate(self.pipeline.process(img, training=training)):
                if v == variant:
                    self.minibatch_data.mem[i] = out
                    break
            self._raw_minibatch_labels[i] = self.key_labels[key_index]

    def analyze_dataset(self):
        if self.normalization_type == "none" or self.normalizer.is_initialized:
            return
        super(ImageLoader, self).analyze_dataset()


class FileListImageLoader(FileListLoaderBase, ImageLoader):
    """Image files named by list files (`path label` per line)."""
    MAPPING = "file_list_image"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        FileListLoaderBase.__init__(self, **kwargs)
        ImageLoader.__init__(self, workflow, **kwargs)


class AutoLabelFileImageLoader(AutoLabelFileLoader, ImageLoader):
    """Image files found under directories; the label is derived from the path by `label_regexp`."""
    MAPPING = "auto_label_file_image"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("mime_prefix", "image")
        AutoLabelFileLoader.__init__(self, **kwargs)
        ImageLoader.__init__(self, workflow, **kwargs)
