"""Data loaders (reference: veles/loader/)."""
from .base import (Loader, ILoader, LoaderMSEMixin, LoaderMSE, UserLoaderRegistry, LoaderError,  # noqa: F401
                   TEST, VALID, TRAIN, CLASS_NAME, TARGET)
from .fullbatch import FullBatchLoader, FullBatchLoaderMSE, IFullBatchLoader  # noqa: F401
# importing the concrete loaders registers their MAPPING names in UserLoaderRegistry.loaders
from . import (synthetic, stream, file_image, fullbatch_image, image_mse, pickles, loader_hdf5, saver, interactive,  # noqa: F401,E402
               restful, ensemble)
