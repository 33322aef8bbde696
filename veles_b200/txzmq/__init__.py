"""ZeroMQ transport helpers (successor of the vendored ``veles/txzmq`` package: streaming pickler frames,
random-port endpoints, shared-memory payloads).  The implementation lives in ``network_common``; this
package keeps the reference's import locations."""
from .connection import ZmqConnection, ZmqEndpoint, ZmqEndpointType  # noqa: F401
from .sharedio import SharedIO  # noqa: F401
