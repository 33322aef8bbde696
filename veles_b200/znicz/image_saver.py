"""ImageSaver: dumps misclassified (or all) samples of the current minibatch as image files, grouped by class
(``link_image_saver`` of StandardWorkflow; kwargs ``out_dirs``, ``limit``)."""
from __future__ import annotations

import os

import numpy

from ..config import root, get
from ..distributable import TriviallyDistributable
from ..loader.base import CLASS_NAME
from ..units import Unit


class ImageSaver(Unit, TriviallyDistributable):
    """Writes misclassified (or all) samples of a minibatch as image files grouped by class / outcome."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(ImageSaver, self).__init__(workflow, **kwargs)
        base = get(root.common.dirs.cache, "/tmp")
        self.out_dirs = list(kwargs.get("out_dirs", [os.path.join(base, "images", n) for n in CLASS_NAME]))
        self.limit = kwargs.get("limit", 100)
        self.only_errors = kwargs.get("only_errors", True)
        self.saved = 0
        self.demand("input", "labels", "max_idx", "minibatch_class", "minibatch_size")
        self.indices = None

    def initialize(self, **kwargs):
        for d in self.out_dirs:
            os.makedirs(d, exist_ok=True)

    def run(self):
        if self.saved >= self.limit:
            return
        import cv2
        for a in (self.input, self.labels, self.max_idx):
            a.map_read()
        n = int(self.minibatch_size)
        out_dir = self.out_dirs[int(self.minibatch_class)]
        for i in range(n):
            truth, pred = int(self.labels.mem[i]), int(self.max_idx.mem[i])
            if self.only_errors and truth == pred:
                continue
            img = numpy.asarray(self.input.mem[i], dtype=numpy.float32)
            if img.ndim == 1:
                side = int(round(img.size ** 0.5))
                img = img[:side * side].reshape(side, side)
            lo, hi = img.min(), img.max()
            img = ((img - lo) * (255.0 / (hi - lo)) if hi > lo else img * 0).astype(numpy.uint8)
            idx = int(self.indices.mem[i]) if self.indices is not None and self.indices else self.saved
            cv2.imwrite(os.path.join(out_dir, "%d_as_%d_idx%d.png" % (truth, pred, idx)), img)
            self.saved += 1
            if self.saved >= self.limit:
                return
