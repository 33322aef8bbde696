"""Synthetic full-batch loaders standing in for the dataset loaders of the absent Znicz
package (``mnist_loader``, ``cifar_loader``, ``imagenet_loader``: SURVEY §2.4) — there is no
network on the build or GPU boxes, so datasets are generated with the named shapes.

Samples are class-conditional Gaussians (a fixed random prototype per class plus noise), so
a network can actually learn them: convergence smoke tests are meaningful.
"""
from __future__ import annotations

import numpy

from .base import TEST, VALID, TRAIN
from .fullbatch import FullBatchLoader, FullBatchLoaderMSE


class SyntheticFullBatchLoader(FullBatchLoader):
    """kwargs: sample_shape, n_classes, n_train, n_valid, n_test, dtype ("uint8"|"float32"), noise, seed."""
    MAPPING = "synthetic_fullbatch"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(SyntheticFullBatchLoader, self).__init__(workflow, **kwargs)
        self.sample_shape = tuple(kwargs.get("sample_shape", (784,)))
        self.n_classes = int(kwargs.get("n_classes", 10))
        self.n_train = int(kwargs.get("n_train", 1000))
        self.n_valid = int(kwargs.get("n_valid", 200))
        self.n_test = int(kwargs.get("n_test", 0))
        self.data_dtype = kwargs.get("dtype", "float32")
        self.noise = float(kwargs.get("noise", 0.5))
        self.seed = int(kwargs.get("seed", 1234))

    def load_data(self):
        if self.original_data and self.restored_from_snapshot:
            return
        rng = numpy.random.RandomState(self.seed)
        total = self.n_test + self.n_valid + self.n_train
        protos = self._prototypes(rng)
        labels = rng.randint(0, self.n_classes, size=total)
        data = numpy.empty((total,) + self.sample_shape, dtype=numpy.float32)
        chunk = max(1, (1 << 24) // max(1, int(numpy.prod(self.sample_shape))))
        for i in range(0, total, chunk):
            l = labels[i:i + chunk]
            data[i:i + chunk] = protos[l] + self.noise * rng.standard_normal(
                (len(l),) + self.sample_shape).astype(numpy.float32)
        if self.data_dtype == "uint8":
            data = numpy.clip((data + 2.0) * 64.0, 0, 255).astype(numpy.uint8)
        self.original_data.mem = data
        self.original_labels = [int(v) for v in labels]
        self.class_lengths[TEST] = self.n_test
        self.class_lengths[VALID] = self.n_valid
        self.class_lengths[TRAIN] = self.n_train


    def _prototypes(self, rng):
        """Class prototypes; image-shaped samples get spatially smooth (blocky low-frequency)
        prototypes so that convolution + pooling can separate the classes."""
        shape = self.sample_shape
        if len(shape) == 3 and shape[0] >= 8 and shape[1] >= 8:
            H, W, C = shape
            cell = max(2, min(H, W) // 6)
            gh, gw = -(-H // cell), -(-W // cell)
            coarse = rng.uniform(-1.0, 1.0, size=(self.n_classes, gh, gw, C)).astype(numpy.float32)
            full = numpy.repeat(numpy.repeat(coarse, cell, axis=1), cell, axis=2)
            return numpy.ascontiguousarray(full[:, :H, :W, :])
        return rng.uniform(-1.0, 1.0, size=(self.n_classes,) + shape).astype(numpy.float32)


class MnistLikeLoader(SyntheticFullBatchLoader):
    """MNIST-shaped (784 features / 28x28x1, 10 classes) synthetic data."""
    MAPPING = "mnist_loader"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("sample_shape", (784,))
        kwargs.setdefault("n_train", 6000)
        kwargs.setdefault("n_valid", 1000)
        super(MnistLikeLoader, self).__init__(workflow, **kwargs)


class CifarLikeLoader(SyntheticFullBatchLoader):
    """CIFAR-10-shaped (32x32x3 NHWC uint8, 10 classes) synthetic data."""
    MAPPING = "cifar_loader"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("sample_shape", (32, 32, 3))
        kwargs.setdefault("dtype", "uint8")
        kwargs.setdefault("n_train", 5000)
        kwargs.setdefault("n_valid", 1000)
        super(CifarLikeLoader, self).__init__(workflow, **kwargs)


class ImagenetLikeLoader(SyntheticFullBatchLoader):
    """ImageNet/AlexNet-shaped (227x227x3 NHWC uint8, 1000 classes) synthetic data."""
    MAPPING = "imagenet_loader"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("sample_shape", (227, 227, 3))
        kwargs.setdefault("dtype", "uint8")
        kwargs.setdefault("n_classes", 1000)
        kwargs.setdefault("n_train", 2048)
        kwargs.setdefault("n_valid", 0)
        super(ImagenetLikeLoader, self).__init__(workflow, **kwargs)


class SyntheticMSELoader(FullBatchLoaderMSE):
    """Autoencoder-style regression data: targets = a fixed linear map of the input (or the input)."""
    MAPPING = "synthetic_mse"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(SyntheticMSELoader, self).__init__(workflow, **kwargs)
        self.sample_shape = tuple(kwargs.get("sample_shape", (64,)))
        self.target_size = kwargs.get("target_size", None)
        self.n_train = int(kwargs.get("n_train", 1000))
        self.n_valid = int(kwargs.get("n_valid", 200))
        self.seed = int(kwargs.get("seed", 4321))

    def load_data(self):
        if self.original_data and self.restored_from_snapshot:
            return
        rng = numpy.random.RandomState(self.seed)
        total = self.n_valid + self.n_train
        data = rng.uniform(-1, 1, size=(total,) + self.sample_shape).astype(numpy.float32)
        flat = data.reshape(total, -1)
        if self.target_size is None:
            targets = flat.copy()
        else:
            proj = rng.standard_normal((flat.shape[1], self.target_size)).astype(numpy.float32) / \
                numpy.sqrt(flat.shape[1])
            targets = numpy.tanh(flat.dot(proj))
        self.original_data.mem = data
        self.original_targets.mem = targets
        self.class_lengths[TEST] = 0
        self.class_lengths[VALID] = self.n_valid
        self.class_lengths[TRAIN] = self.n_train
