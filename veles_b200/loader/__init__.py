"""Data loaders (reference: veles/loader/)."""
from .base import (Loader, ILoader, LoaderMSEMixin, LoaderMSE, UserLoaderRegistry, LoaderError,  # noqa: F401
                   TEST, VALID, TRAIN, CLASS_NAME, TARGET)
from .fullbatch import FullBatchLoader, FullBatchLoaderMSE, IFullBatchLoader  # noqa: F401
