"""RestfulLoader: collects requests arriving through RESTfulAPI into minibatches — a minibatch is served when
it is full or ``max_response_time`` elapsed since its first request (reference: ``veles/loader/restful.py:52-130``)."""
from __future__ import annotations

import queue
import time

import numpy

from .base import TEST
from .interactive import InteractiveLoader


class RestfulLoader(InteractiveLoader):
    """Collects samples posted to the RESTful API into minibatches (until full or `max_response_time`) and
    hands results back to the requests."""
    MAPPING = "restful"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("minibatch_size", 16)
        super(RestfulLoader, self).__init__(workflow, **kwargs)
        self.max_response_time = float(kwargs.get("max_response_time", 0.1))
        self.requests = []   # per-row opaque request handles of the current minibatch

    def feed(self, obj, request=None):
        self.queue_.put((obj, request))

    def run(self):
        count = 0
        del self.requests[:]          # in place: RESTfulAPI holds a reference to this very list
        self.minibatch_data.map_invalidate()
        deadline = None
        while count < self._max_minibatch_size:
            try:
                timeout = None if deadline is None else max(0.0, deadline - time.time())
                item = self.queue_.get(timeout=timeout if timeout is not None else self.timeout)
            except queue.Empty:
                break
            if item is None or item[0] is None:
                self.complete <<= True
                break
            obj, request = item
            if deadline is None:
                deadline = time.time() + self.max_response_time
            self.minibatch_data.mem[count] = self._decode(obj)
            self.requests.append(request)
            count += 1
        self.minibatch_data.mem[count:] = 0
        self._minibatch_size_ = count
        self.minibatch_class = TEST
        self.minibatch_data.unmap()
        self.samples_served += count
