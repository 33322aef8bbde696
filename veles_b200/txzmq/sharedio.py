"""``SharedIO``: same-host job payloads through POSIX shared memory (reference ``txzmq/sharedio.py:45-105``)."""
from ..network_common import SharedIO  # noqa: F401
