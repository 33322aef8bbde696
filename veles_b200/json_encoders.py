"""JSON encoders for numpy / Config values (reference: veles/json_encoders.py)."""
import json

import numpy


class NumpyJSONEncoder(json.JSONEncoder):
    def default(self, o):
        if isinstance(o, numpy.ndarray):
            return o.tolist()
        if isinstance(o, numpy.integer):
            return int(o)
        if isinstance(o, numpy.floating):
            return float(o)
        if isinstance(o, numpy.bool_):
            return bool(o)
        if isinstance(o, (set, frozenset)):
            return sorted(o, key=str)
        if isinstance(o, bytes):
            return o.decode("utf-8", "replace")
        from .config import Config
        if isinstance(o, Config):
            return o.to_dict()
        if callable(o):
            return getattr(o, "__name__", repr(o))
        try:
            import torch
            if isinstance(o, torch.Tensor):
                return o.detach().cpu().tolist()
        except ImportError:
            pass
        return json.JSONEncoder.default(self, o)


ConfigJSONEncoder = NumpyJSONEncoder
