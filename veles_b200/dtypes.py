"""Precision / dtype tables (successor of ``veles/opencl_types.py:41-78``: dtype name maps and the
``dtype``/``sizeof_dtype`` defines injected into kernel sources).

There is no source templating here (kernels are compiled ahead of time for sm_100a), so the tables map the
reference's ``root.common.engine.precision_type`` names onto host numpy dtypes, device torch dtypes and the
dtype codes of the C ABI (``include/veles_b200.h``).
"""
from __future__ import annotations

import numpy

#: precision_type -> numpy dtype of host-side parameters
dtypes = {"float": numpy.float32, "double": numpy.float64, "float32": numpy.float32, "bfloat16": numpy.float32,
          "half": numpy.float16}
#: C ABI dtype codes used by the vb_* entry points
abi_codes = {"float32": 0, "bfloat16": 1, "uint8": 2, "int32": 3}
#: element sizes on the device
sizeof = {"float32": 4, "bfloat16": 2, "uint8": 1, "int32": 4, "float16": 2, "float64": 8}


def numpy_dtype_to_abi(dtype):
    dtype = numpy.dtype(dtype)
    if dtype == numpy.float32:
        return abi_codes["float32"]
    if dtype == numpy.uint8:
        return abi_codes["uint8"]
    if dtype == numpy.int32:
        return abi_codes["int32"]
    raise ValueError("dtype %s has no device representation" % dtype)


def torch_dtype(name):
    import torch
    return {"float32": torch.float32, "float": torch.float32, "bfloat16": torch.bfloat16, "uint8": torch.uint8,
            "int32": torch.int32, "half": torch.float16, "float16": torch.float16, "double": torch.float64}[name]


def defines(precision_type="float"):
    """The (name -> value) pairs the reference injected as ``#define``s; kept for tools that print them."""
    dt = dtypes[precision_type]
    return {"dtype": precision_type, "sizeof_dtype": numpy.dtype(dt).itemsize,
            "PRECISION_LEVEL": 0, "compute_dtype": "bfloat16", "accumulate_dtype": "float32"}
