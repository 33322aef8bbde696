"""Report publishing (reference: veles/publishing/)."""
from .publisher import Publisher  # noqa: F401
from .backend import Backend, MarkdownBackend, HTMLBackend, JSONBackend  # noqa: F401
