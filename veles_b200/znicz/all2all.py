"""Fully connected layers: ``output = f(input · Wᵀ + bias)``.

Layer types (``docs/source/manualrst_veles_workflow_parameters.rst:469-498``): ``all2all``,
``all2all_tanh`` (scaled tanh 1.7159·tanh(0.6666·s)), ``all2all_relu`` (= log(1+eˢ)),
``all2all_str`` (strict ReLU max(0,s)), ``all2all_sigmoid``, ``softmax``.
Device path = one tcgen05 GEMM with the bias+activation epilogue (csrc/gemm_tcgen05.cu);
the reference did cuBLAS/OpenCL GEMM + a separate bias/activation kernel.
"""
from __future__ import annotations

import numpy

from ..memory import Array
from .nn_units import Forward, act_forward_numpy


class All2All(Forward):
    """Linear fully connected layer."""
    hide_from_registry = False
    MAPPING = {"all2all"}
    ACTIVATION = 0
    TWO_PHASE_SPLIT_K = False   # small-batch layers: split-K GEMM into an fp32 scratch + bias/activation pass
    PEER_WEIGHTS_OK = True   # data parallel "peer_tma" delivery: GEMMs may read weight tiles from the owner rank
    # weights stored (n_in, n_out): the GEMMs read the same buffer with the other operand major, no transposed copy
    TRANSPOSED_OK = True
    __id__ = "58a5eadf-ae1e-498f-bf35-7d93939c4c86"

    def __init__(self, workflow, **kwargs):
        super(All2All, self).__init__(workflow, **kwargs)
        shape = kwargs.get("output_sample_shape", kwargs.get("output_samples_number", kwargs.get("output_shape")))
        if shape is None:
            raise ValueError("output_sample_shape is required")
        self._output_sample_shape = shape
        self.output_dtype = kwargs.get("output_dtype", None)

    @property
    def output_sample_shape(self):
        shape = self._output_sample_shape
        if hasattr(shape, "__iter__"):
            return tuple(int(v) for v in shape)
        return (int(shape),)  # ints and deferred counts (resolved once the loader is initialized)

    @output_sample_shape.setter
    def output_sample_shape(self, value):
        self._output_sample_shape = value

    @property
    def neurons_number(self):
        return int(numpy.prod(self.output_sample_shape))

    def __getstate__(self):
        state = super(All2All, self).__getstate__()
        state["_output_sample_shape"] = self.output_sample_shape if self.is_initialized else \
            (state["_output_sample_shape"] if isinstance(state["_output_sample_shape"], (int, tuple, list))
             else None)
        return state

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True  # retry after the producer allocated its output
        self.device = device
        batch = self.input.shape[0]
        n_in = self.input.size // batch
        n_out = self.neurons_number
        self.create_weights(n_out, n_in)
        if not self.output or self.output.shape != (batch,) + self.output_sample_shape:
            self.output.reset(numpy.zeros((batch,) + self.output_sample_shape, dtype=numpy.float32))
        self._backend_init_()
        return None

    # -- numpy oracle --------------------------------------------------------------------------------------
    def numpy_linear(self):
        self.input.map_read()
        self.weights.map_read()
        x = self.input.mem.reshape(self.input.shape[0], -1).astype(numpy.float32)
        w = self.weights.mem
        s = x.dot(w if self.weights_transposed else w.T)
        if self.include_bias:
            self.bias.map_read()
            s += self.bias.mem
        return s

    def numpy_run(self):
        self.output.map_invalidate()
        self.output.mem[...] = act_forward_numpy(self.ACTIVATION, self.numpy_linear()).reshape(self.output.shape)

    # -- device ---------------------------------------------------------------------------------------------------
    def b200_init(self):
        import torch
        from ..ops import kernels
        dev = self.device_
        self.place_parameters()
        self.output.dev_dtype = dev.compute_dtype
        self.output._pitched = True
        self.output.initialize(dev)
        self._x_stage_ = None
        n_in = self.weights_bf16_.shape[0 if self.weights_transposed else 1]
        xin = self.input.devmem
        # "float" precision (root.common.engine.precision_type): fp32 activations and master weights go straight into
        # the tf32 tensor-core GEMM, no bf16 staging and no weight shadow on this layer's path
        self.tf32_ = (dev.compute_dtype == torch.float32 and xin.dtype == torch.float32 and n_in % 4 == 0
                      and not self.weights_transposed)
        # Small-batch layers stream the whole weight matrix through few M tiles: splitting K over more CTAs (fp32
        # reduce-add into a scratch, then one tiny bias+activation pass) cuts the re-reads of the activations.
        n_out = self.weights_bf16_.shape[1 if self.weights_transposed else 0]
        self.acc_ = None
        # (measured on AlexNet's fc layers: GEMMs 15-30 % faster, but the extra pass eats the gain -> opt-in)
        from ..config import root, get
        if get(root.common.engine.two_phase_split_k, False):
            self.TWO_PHASE_SPLIT_K = True
        if (self.TWO_PHASE_SPLIT_K and not self.tf32_ and n_out % 4 == 0 and not self.weights_transposed
                and kernels.gemm_plan(self.input.shape[0], n_out, n_in)[1] >= 2):
            self.acc_ = torch.zeros(self.input.shape[0], n_out, dtype=torch.float32, device=dev.torch_device)
        need_stage = not self.tf32_ and ((xin.dtype != torch.bfloat16) or (n_in % 8 != 0))
        if need_stage:
            self._x_stage_ = kernels.pitched_empty(self.input.shape[0], n_in, torch.bfloat16, dev.torch_device)

    def device_input_2d(self):
        """[batch, n_in] view of the input with a TMA-legal pitch (bf16, or fp32 in tf32 mode)."""
        from ..ops import kernels
        x = self.input.devmem
        x2 = x if x.dim() == 2 else x.reshape(x.shape[0], -1)
        if getattr(self, "tf32_", False) and x2.stride(0) % 4 != 0:
            x2 = x2.contiguous()
        if self._x_stage_ is not None:
            kernels.copy_2d(x2, self._x_stage_)
            return self._x_stage_
        return x2

    def b200_run(self):
        from ..ops import kernels
        self.input.unmap()
        x = self.device_input_2d()
        out = self.output.devmem
        out2 = out if out.dim() == 2 else out.reshape(out.shape[0], -1)
        if self.tf32_:
            self.weights.unmap()
            kernels.gemm_tf32(x, self.weights.devmem.view(self.weights_bf16_.shape), out2,
                              bias=self.bias.devmem if self.include_bias else None, act=self.ACTIVATION)
        elif self.acc_ is not None and getattr(self, "peer_weights_", None) is None:
            kernels.gemm(x, self.weights_bf16_, self.acc_, accumulate=True, split_k=0)
            kernels.bias_act(self.acc_, out2, bias=self.bias.devmem if self.include_bias else None, act=self.ACTIVATION)
        else:
            kernels.gemm(x, self.weights_bf16_, out2, b_mn=self.weights_transposed,
                         bias=self.bias.devmem if self.include_bias else None,
                         act=self.ACTIVATION, b_peers=getattr(self, "peer_weights_", None))
        self.output.mark_device_written()


class All2AllTanh(All2All):
    """Fully connected layer with y = 1.7159 tanh(0.6666 (x W^T + b))."""
    MAPPING = {"all2all_tanh"}
    ACTIVATION = 1
    __id__ = "b3a2bd5c-3c01-46ef-978a-fef22e008f31"


class All2AllRELU(All2All):
    """Fully connected layer with softplus activation y = log(1 + exp(x W^T + b))."""
    MAPPING = {"all2all_relu"}
    ACTIVATION = 3
    __id__ = "5b7f36d8-f8c8-4eb7-8af3-75eb3cfca3fe"


class All2AllStrictRELU(All2All):
    """Fully connected layer with y = max(0, x W^T + b)."""
    MAPPING = {"all2all_str"}
    ACTIVATION = 4
    __id__ = "fe63baf0-4fe4-4cf3-bafb-ef1215bf27a8"


class All2AllSigmoid(All2All):
    """Fully connected layer with logistic activation."""
    MAPPING = {"all2all_sigmoid"}
    ACTIVATION = 2
    __id__ = "a27974ec-1764-4944-925d-4862de237881"


class All2AllSoftmax(All2All):
    """Linear layer + softmax over the outputs; also emits ``max_idx`` (argmax per sample).

    On the device the softmax is normally *deferred* to EvaluatorSoftmax, which runs one fused
    softmax+argmax+evaluator kernel over ``logits``; a stand-alone run (no
    evaluator linked, e.g. inference) launches the same kernel without labels.
    """
    MAPPING = {"softmax"}
    ACTIVATION = 0
    __id__ = "420219fc-3e1a-45b1-87f8-aaa0c1540de4"

    def __init__(self, workflow, **kwargs):
        super(All2AllSoftmax, self).__init__(workflow, **kwargs)
        self.max_idx = Array(shallow_pickle=True)
        self.defer_softmax = False  # set by EvaluatorSoftmax when it fuses the softmax

    def init_unpickled(self):
        super(All2AllSoftmax, self).init_unpickled()
        self.logits_ = None

    def initialize(self, device=None, **kwargs):
        res = super(All2AllSoftmax, self).initialize(device=device, **kwargs)
        if res:
            return res
        if not self.max_idx or len(self.max_idx) != self.output.shape[0]:
            self.max_idx.reset(numpy.zeros(self.output.shape[0], dtype=numpy.int32))
        if self.on_device:
            self.max_idx.initialize(self.device_)
        return None

    def numpy_run(self):
        s = self.numpy_linear()
        s -= s.max(axis=1, keepdims=True)
        e = numpy.exp(s)
        self.output.map_invalidate()
        self.output.mem[...] = (e / e.sum(axis=1, keepdims=True)).reshape(self.output.shape)
        self.max_idx.map_invalidate()
        self.max_idx.mem[...] = self.output.mem.reshape(self.output.shape[0], -1).argmax(axis=1)

    def b200_init(self):
        import torch
        from ..ops import kernels
        super(All2AllSoftmax, self).b200_init()
        self.logits_ = kernels.pitched_empty(self.output.shape[0], self.neurons_number, torch.float32,
                                             self.device_.torch_device, align=4)

    def b200_run(self):
        from ..ops import kernels
        self.input.unmap()
        x = self.device_input_2d()
        if self.tf32_:
            self.weights.unmap()
            kernels.gemm_tf32(x, self.weights.devmem.view(self.weights_bf16_.shape), self.logits_,
                              bias=self.bias.devmem if self.include_bias else None)
        elif self.acc_ is not None and getattr(self, "peer_weights_", None) is None:
            self.logits_.zero_()    # K-splits reduce-add into the logits; the first split adds the bias
            kernels.gemm(x, self.weights_bf16_, self.logits_, bias=self.bias.devmem if self.include_bias else None,
                         accumulate=True, split_k=0)
        else:
            kernels.gemm(x, self.weights_bf16_, self.logits_, b_mn=self.weights_transposed,
                         bias=self.bias.devmem if self.include_bias else None,
                         b_peers=getattr(self, "peer_weights_", None))
        if not self.defer_softmax:
            out = self.output.devmem
            kernels.softmax_eval(self.logits_, out if out.dim() == 2 else out.reshape(out.shape[0], -1), None, None,
                                 self.max_idx.devmem)
            self.output.mark_device_written()
            self.max_idx.mark_device_written()
