"""Stand-alone activation layers and their backward units:
``activation_tanh, activation_sigmoid, activation_relu, activation_str, activation_log,
activation_tanhlog, activation_sincos, activation_mul``
(``docs/source/manualrst_veles_workflow_parameters.rst:473-503``).
tanh/sigmoid/relu/str differentiate through the output, log/tanhlog/sincos/mul through the input.
"""
from __future__ import annotations

import numpy

from .nn_units import ForwardBase, GDBase, act_forward_numpy, act_deriv_numpy

CODES = {"tanh": 1, "sigmoid": 2, "relu": 3, "str": 4, "log": 5, "tanhlog": 6, "sincos": 7, "mul": 8}


def activation_forward_numpy(code, x, factor=1.0):
    if code <= 4:
        return act_forward_numpy(code, x)
    if code == 5:
        return numpy.log(x + numpy.sqrt(x * x + 1.0))
    if code == 6:
        d, a, b = 3.0, 0.242528761112, 305.459953195
        inner = 1.7159 * numpy.tanh(0.6666 * x)
        big = numpy.log(numpy.maximum(numpy.abs(x), 1e-30) * b) * a
        return numpy.where(x > d, big, numpy.where(x < -d, -big, inner))
    if code == 7:
        idx = numpy.arange(x.size).reshape(x.shape) & 1
        return numpy.where(idx == 1, numpy.sin(x), numpy.cos(x))
    return x * factor


def activation_backward_numpy(code, x, y, factor=1.0):
    if code <= 4:
        return act_deriv_numpy(code, y)
    if code == 5:
        return 1.0 / numpy.sqrt(x * x + 1.0)
    if code == 6:
        d, a = 3.0, 0.242528761112
        return numpy.where(numpy.abs(x) > d, a / numpy.maximum(numpy.abs(x), 1e-30),
                           y * y * (-0.388484177) + 1.14381894)
    if code == 7:
        idx = numpy.arange(x.size).reshape(x.shape) & 1
        return numpy.where(idx == 1, numpy.cos(x), -numpy.sin(x))
    return numpy.full_like(x, factor)


class ActivationForward(ForwardBase):
    hide_from_registry = True
    CODE = 0
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0023"

    def __init__(self, workflow, **kwargs):
        super(ActivationForward, self).__init__(workflow, **kwargs)
        self.factor = float(kwargs.get("factor", 1.0))

    @property
    def ACTIVATION(self):  # output-differentiable kinds can be fused by the downstream GD unit
        return self.CODE if self.CODE <= 4 else 0

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_output(self.input.shape)
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
        self._backend_init_()
        return None

    def package_export(self):
        return {"activation": self.CODE, "factor": self.factor}

    def numpy_run(self):
        self.input.map_read()
        self.output.map_invalidate()
        self.output.mem[...] = activation_forward_numpy(self.CODE, self.input.mem.astype(numpy.float32), self.factor)

    def b200_run(self):
        from ..ops import kernels
        x = self.input.devmem
        y = self.output.devmem
        if x.is_contiguous() and y.is_contiguous():
            kernels.activation_forward(self.CODE, x, y, self.factor)
        else:
            tmp = x.contiguous()
            kernels.activation_forward(self.CODE, tmp, tmp, self.factor)
            y.copy_(tmp.view(y.shape))
        self.output.mark_device_written()


class ActivationBackward(GDBase):
    hide_from_registry = True
    CODE = 0

    def __init__(self, workflow, **kwargs):
        super(ActivationBackward, self).__init__(workflow, **kwargs)
        self.demand("input", "output")

    @property
    def ACTIVATION(self):
        return self.CODE if self.CODE <= 4 else 0

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_err_input()
        if self.on_device:
            self.err_input.dev_dtype = self.device_.compute_dtype
            self.err_input.initialize(self.device_)
        self._backend_init_()
        return None

    @property
    def factor(self):
        return getattr(self.forward, "factor", 1.0) if self.forward is not None else 1.0

    def numpy_run(self):
        self.err_output.map_read()
        self.input.map_read()
        self.output.map_read()
        eo = self.err_output.mem.reshape(self.input.shape).astype(numpy.float32)
        if self.derivative_applied:
            ei = eo
        else:
            ei = eo * activation_backward_numpy(self.CODE, self.input.mem.astype(numpy.float32),
                                                self.output.mem.astype(numpy.float32), self.factor)
        self.err_input.map_invalidate()
        self.err_input.mem[...] = ei

    def b200_run(self):
        from ..ops import kernels
        eo = self.err_output.devmem
        ei = self.err_input.devmem
        src = eo.contiguous().view(ei.shape) if (not eo.is_contiguous() or eo.dtype != ei.dtype) else eo.view(ei.shape)
        if src.dtype != ei.dtype:
            src = src.to(ei.dtype)
        if self.derivative_applied:
            ei.copy_(src)
        else:
            x = self.input.devmem.reshape(ei.shape)
            y = self.output.devmem.reshape(ei.shape)
            if x.dtype != ei.dtype:
                x = x.to(ei.dtype)
            kernels.activation_backward(self.CODE, src, x.contiguous(), y.contiguous(), ei, self.factor)
        self.err_input.mark_device_written()


def _make(name, code):
    fwd = type("Forward" + name, (ActivationForward,), {"MAPPING": {"activation_" + name.lower()}, "CODE": code,
                                                        "__id__": ActivationForward.__id__,
                                                        "hide_from_registry": False, "__module__": __name__})
    bwd = type("Backward" + name, (ActivationBackward,), {"MAPPING": {"activation_" + name.lower()}, "CODE": code,
                                                          "hide_from_registry": False, "__module__": __name__})
    return fwd, bwd


ForwardTanh, BackwardTanh = _make("Tanh", 1)
ForwardSigmoid, BackwardSigmoid = _make("Sigmoid", 2)
ForwardRELU, BackwardRELU = _make("RELU", 3)
ForwardStrictRELU, BackwardStrictRELU = _make("Str", 4)
ForwardLog, BackwardLog = _make("Log", 5)
ForwardTanhLog, BackwardTanhLog = _make("TanhLog", 6)
ForwardSinCos, BackwardSinCos = _make("SinCos", 7)
ForwardMul, BackwardMul = _make("Mul", 8)
