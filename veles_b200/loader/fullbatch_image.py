"""Full-batch image loaders: decode + preprocess every image once, keep the uint8 dataset resident in HBM
(reference: ``veles/loader/fullbatch_image.py``): ``full_batch_file_list_image``,
``full_batch_auto_label_file_image``, ``full_batch_auto_label_file_image_mse``."""
from __future__ import annotations

import numpy

from .base import TEST, VALID, TRAIN, LoaderError
from .file_loader import FileListLoaderBase, AutoLabelFileLoader
from .fullbatch import FullBatchLoader, FullBatchLoaderMSEMixin
from .image import ImagePipeline


class FullBatchImageLoader(FullBatchLoader):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(FullBatchImageLoader, self).__init__(workflow, **kwargs)
        self.pipeline = ImagePipeline(prng=self.prng, **kwargs)

    def load_data(self):
        if self.original_data and self.restored_from_snapshot:
            return
        class_keys = self.collect_keys()
        samples, labels = [], []
        for ci in (TEST, VALID, TRAIN):
            count = 0
            for key in class_keys[ci]:
                img = self.pipeline.decode(key)
                label = self.get_label_from_filename(key)
                for out in self.pipeline.process(img, training=False):
                    samples.append(out)
                    labels.append(label)
                    count += 1
            self.class_lengths[ci] = count
        if not samples:
            raise LoaderError("No images were found")
        shapes = {s.shape for s in samples}
        if len(shapes) != 1:
            raise LoaderError("Images have different shapes after preprocessing: %s (set scale/crop)" % shapes)
        self.original_data.mem = numpy.stack(samples)
        self.original_labels = labels if any(l is not None for l in labels) else []
        if self.validation_ratio:
            self.resize_validation()


class FullBatchFileListImageLoader(FileListLoaderBase, FullBatchImageLoader):
    """FileListImageLoader that decodes the whole set once and keeps it resident."""
    MAPPING = "full_batch_file_list_image"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        FileListLoaderBase.__init__(self, **kwargs)
        FullBatchImageLoader.__init__(self, workflow, **kwargs)


class FullBatchAutoLabelFileImageLoader(AutoLabelFileLoader, FullBatchImageLoader):
    """AutoLabelFileImageLoader that decodes the whole set once and keeps it resident (device gather per
    minibatch)."""
    MAPPING = "full_batch_auto_label_file_image"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("mime_prefix", "image")
        AutoLabelFileLoader.__init__(self, **kwargs)
        FullBatchImageLoader.__init__(self, workflow, **kwargs)


class FullBatchAutoLabelFileImageLoaderMSE(FullBatchLoaderMSEMixin, FullBatchAutoLabelFileImageLoader):
    """Targets = one image per label from ``target_paths`` (same preprocessing), e.g. denoising / kanji."""
    MAPPING = "full_batch_auto_label_file_image_mse"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(FullBatchAutoLabelFileImageLoaderMSE, self).__init__(workflow, **kwargs)
        self.target_paths = list(kwargs.get("target_paths", []))

    def load_data(self):
        super(FullBatchAutoLabelFileImageLoaderMSE, self).load_data()
        by_label = {}
        for p in self.target_paths:
            for key in self.scan_files(p):
                img = self.pipeline.decode(key)
                by_label[self.get_label_from_filename(key)] = next(self.pipeline.process(img)).astype(numpy.float32)
        missing = set(self.original_labels) - set(by_label)
        if missing:
            raise LoaderError("No target image for labels %s" % sorted(missing, key=str))
        self.original_targets.mem = numpy.stack([by_label[l] for l in self.original_labels])
