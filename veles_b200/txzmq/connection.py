"""A ZeroMQ socket that speaks the framework's pickled-object framing.

Reference: ``veles/txzmq/connection.py:97-617`` (a Twisted ``IReadDescriptor`` around a zmq socket with a
streaming pickler: ``b'vpb'+codec`` ... ``b'vpe'`` frames, optional gzip/snappy/xz, ``rndtcp://`` random-bind
endpoints).  Here it is a plain blocking/polling wrapper (the agents run their own threads), sharing the codec
table and frame markers with ``network_common`` so master, slaves and external ZeroMQLoader clients agree.
"""
from __future__ import annotations

from ..network_common import dumps, loads, default_codec


class ZmqEndpointType(object):
    bind = "bind"
    connect = "connect"


class ZmqEndpoint(object):
    def __init__(self, type, address):
        self.type = type
        self.address = address

    def __repr__(self):
        return "ZmqEndpoint(%s, %s)" % (self.type, self.address)


class ZmqConnection(object):
    """``socket_type`` is a zmq constant; endpoints may use ``rndtcp://host`` to bind a random port."""

    def __init__(self, socket_type, endpoints=(), identity=None, codec=None, context=None):
        import zmq
        self.context = context or zmq.Context.instance()
        self.socket = self.context.socket(socket_type)
        if identity is not None:
            self.socket.setsockopt(zmq.IDENTITY, identity if isinstance(identity, bytes) else str(identity).encode())
        self.codec = codec or default_codec()
        self.endpoints = []
        for ep in endpoints:
            self.add_endpoint(ep)

    def add_endpoint(self, ep):
        addr = ep.address
        if ep.type == ZmqEndpointType.bind:
            if addr.startswith("rndtcp://"):
                host = addr[len("rndtcp://"):].split(":")[0] or "127.0.0.1"
                port = self.socket.bind_to_random_port("tcp://" + host)
                addr = "tcp://%s:%d" % (host, port)
            else:
                self.socket.bind(addr)
        else:
            self.socket.connect(addr)
        self.endpoints.append(ZmqEndpoint(ep.type, addr))
        return addr

    def send(self, obj, *prefix_frames, **kwargs):
        """Pickle ``obj`` into marker-delimited frames after the given routing frames."""
        frames = list(prefix_frames) + dumps(obj, kwargs.get("codec", self.codec))
        self.socket.send_multipart(frames)

    def receive(self, timeout_ms=None):
        """(routing frames, object) or None on timeout."""
        import zmq
        if timeout_ms is not None and not self.socket.poll(timeout_ms, zmq.POLLIN):
            return None
        frames = self.socket.recv_multipart()
        start = next(i for i, f in enumerate(frames) if f[:3] == b"vpb")
        return frames[:start], loads(frames[start:])

    def shutdown(self):
        self.socket.close(0)
