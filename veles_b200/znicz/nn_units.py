"""Base classes of the Znicz NN engine: Forward, GradientDescentBase, parameter arena.

The Znicz package is absent from the reference checkout (``veles/znicz`` is an
empty directory), so these classes are specified by the docs
(``docs/source/manualrst_veles_workflow_parameters.rst:469-578``: layer types, kwargs lists,
GD hyper-parameters), by the native export fixture
(``libVeles/tests/workflow_files/contents.json``: ``weights``, ``bias``, ``include_bias``,
``weights_transposed``, ``activation_mode``), and by our own numpy oracle.

Pinned conventions (numpy oracle == device kernels):
  * ``weights.shape == (n_out, n_in)`` (``(n_in, n_out)`` if ``weights_transposed``),
    ``output = f(input · Wᵀ + bias)``;
  * a GD unit receives ``err_output = dL/d(output)`` and produces ``err_input = dL/d(input)``;
    gradient ``g = err_outputᵀ · input`` (the evaluator already divided by the batch size);
    ``g += weights_decay * ((1 - l1_vs_l2) * W + 0.5 * l1_vs_l2 * sign(W))``;
    ``v = gradient_moment * v - learning_rate * g``;  ``W += v``  (bias likewise with *_bias);
  * derivative fusion protocol: a GD unit whose ``prev_activation`` is set multiplies its
    ``err_input`` by the derivative of the unit that produced its ``input`` (that unit's GD
    then has ``derivative_applied = True`` and skips the multiplication).
"""
from __future__ import annotations

import time

import numpy

from .. import prng
from ..accelerated_units import AcceleratedUnit, AcceleratedWorkflow
from ..config import root, get
from ..memory import Array
from ..units import Unit

ACTIVATIONS = {"ACTIVATION_LINEAR": 0, "ACTIVATION_TANH": 1, "ACTIVATION_SIGMOID": 2, "ACTIVATION_RELU": 3,
               "ACTIVATION_STRICT_RELU": 4}


def act_forward_numpy(act, s):
    if act == 1:
        return 1.7159 * numpy.tanh(0.6666 * s)
    if act == 2:
        return 1.0 / (1.0 + numpy.exp(-s))
    if act == 3:
        return numpy.where(s > 15, s, numpy.log1p(numpy.exp(numpy.minimum(s, 15))))
    if act == 4:
        return numpy.maximum(s, 0)
    return s


def act_deriv_numpy(act, y):
    """Derivative through the OUTPUT y."""
    if act == 1:
        return y * y * (-0.388484177) + 1.14381894
    if act == 2:
        return y * (1.0 - y)
    if act == 3:
        return 1.0 - numpy.exp(-y)
    if act == 4:
        return (y > 0).astype(y.dtype)
    return numpy.ones_like(y)


class ParameterArena(object):
    """Bump allocator over big device slabs so that all trainable state of a workflow is a
    handful of contiguous ranges.  ``weights`` slabs are the master layout; ``mirror()`` hands
    out views of other kinds (gradients, velocity, ...) at the SAME byte offsets, so one
    (slab, offset, numel) triple addresses weight, gradient and optimizer state alike — which is
    what the fused all-reduce+GD kernel (veles_b200.parallel) needs.  In data-parallel mode the
    slabs of the exchanged kinds come from symmetric memory (identical offsets on every rank,
    multicast-mapped): ``allocator(kind, nbytes)`` decides."""

    ALIGN = 256  # bytes
    SMALL_KINDS = {"dp_flags": 1 << 16, "dp_counters": (1 << 18) + 4096}

    def __init__(self, device, slab_bytes=None, allocator=None):
        self.device = device
        self.slab_bytes = slab_bytes or int(get(root.common.engine.arena_slab_bytes, 512 << 20))
        self.allocator = allocator  # callable(kind, nbytes) -> uint8 tensor or None (= default)
        self.slabs = {}             # kind -> list of [tensor(uint8), used_bytes]
        self.records = {}           # kind -> list of (name, slab_index, offset_bytes, nbytes)

    def _new_slab(self, kind, nbytes):
        import torch
        t = self.allocator(kind, nbytes) if self.allocator is not None else None
        if t is None:
            return torch.zeros(nbytes, dtype=torch.uint8, device=self.device.torch_device)
        t.zero_()
        return t

    def alloc(self, kind, name, numel, dtype):
        import torch
        esize = torch.empty((), dtype=dtype).element_size()
        nbytes = (numel * esize + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        slabs = self.slabs.setdefault(kind, [])
        if not slabs or slabs[-1][1] + nbytes > slabs[-1][0].numel():
            slabs.append([self._new_slab(kind, max(self.SMALL_KINDS.get(kind, self.slab_bytes), nbytes)), 0])
        slab = slabs[-1]
        off = slab[1]
        slab[1] += nbytes
        self.records.setdefault(kind, []).append((name, len(slabs) - 1, off, numel * esize))
        return slab[0][off:off + numel * esize].view(dtype)

    def locate(self, tensor, kind="weights"):
        """(slab index, byte offset) of a tensor that lives in a slab of ``kind``."""
        ptr = tensor.data_ptr()
        for i, (t, _) in enumerate(self.slabs.get(kind, [])):
            base = t.data_ptr()
            if base <= ptr < base + t.numel():
                return i, ptr - base
        raise ValueError("tensor does not live in a %r slab of this arena" % kind)

    def mirror(self, kind, ref_tensor, ref_kind="weights"):
        """View of ``kind`` storage at the same (slab, byte offset) as ``ref_tensor`` has in ``ref_kind``."""
        idx, off = self.locate(ref_tensor, ref_kind)
        ref_slabs = self.slabs[ref_kind]
        slabs = self.slabs.setdefault(kind, [])
        while len(slabs) <= idx:
            j = len(slabs)
            slabs.append([self._new_slab(kind, ref_slabs[j][0].numel()), ref_slabs[j][0].numel()])
        nbytes = ref_tensor.numel() * ref_tensor.element_size()
        return slabs[idx][0][off:off + nbytes].view(ref_tensor.dtype)

    def used_ranges(self, kind, dtype):
        out = []
        for t, used in self.slabs.get(kind, []):
            if used:
                out.append(t[:used].view(dtype))
        return out


class NNWorkflow(AcceleratedWorkflow):
    """Workflow of NN units; owns the parameter arena on device backends."""
    hide_from_registry = True

    def init_unpickled(self):
        super(NNWorkflow, self).init_unpickled()
        self.arena_ = None

    def get_arena(self, device):
        if self.arena_ is None or self.arena_.device is not device:
            self.arena_ = ParameterArena(device, allocator=self.__dict__.get("arena_allocator_"))
        return self.arena_

    def run(self):
        decision = getattr(self, "decision", None)
        if (decision is not None and bool(getattr(decision, "complete", False)) and not self.is_slave
                and not self.testing):
            # resumed from the snapshot of a finished training and the limits allow nothing more: the loader's gate
            # stays shut, so nothing would ever reach the end point
            self.warning("training is already complete (epoch %s): nothing to run; raise decision.max_epochs / "
                         "fail_iterations to continue from this snapshot", getattr(decision, "epoch_number", "?"))
            self._run_time_started_ = time.time()
            self._finished_ = False
            self._is_running = True
            self.on_workflow_finished()
            return
        super(NNWorkflow, self).run()


def find_arena(unit, device):
    wf = unit.workflow
    while wf is not None:
        if hasattr(wf, "get_arena"):
            return wf.get_arena(device)
        wf = getattr(wf, "workflow", None) if isinstance(wf, Unit) else None
    holder = unit.workflow
    arena = getattr(holder, "_fallback_arena_", None)
    if arena is None or arena.device is not device:
        arena = ParameterArena(device, slab_bytes=1 << 20)
        try:
            holder._fallback_arena_ = arena
        except AttributeError:
            pass
    return arena


class ForwardBase(AcceleratedUnit):
    """Base of every forward-propagation unit (with or without trainable weights)."""
    hide_from_registry = True
    ACTIVATION = 0
    VIEW_GROUP = "WORKER"
    MAPPING = set()
    HAS_WEIGHTS = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "WORKER")
        super(ForwardBase, self).__init__(workflow, **kwargs)
        self.input = None
        self.output = Array(shallow_pickle=True)
        self.demand("input")
        self.forward_mode = False

    def alloc_output(self, shape, like_input_layout=False):
        if not self.output or self.output.shape != tuple(shape):
            self.output.reset(numpy.zeros(shape, dtype=numpy.float32))

    def device_tensor(self, arr, shape=None):
        """Contiguous NHWC (or given shape) view of an Array's device tensor."""
        t = arr.devmem
        shape = tuple(shape or arr.shape)
        if tuple(t.shape) == shape:
            return t
        return t.reshape(shape)

    def generate_data_for_slave(self, slave):
        return None

    def generate_data_for_master(self):
        return None

    def apply_data_from_master(self, data):
        pass

    def apply_data_from_slave(self, data, slave):
        pass

    def drop_slave(self, slave):
        pass


class Forward(ForwardBase):
    """Base of forward-propagation units with trainable weights."""
    hide_from_registry = True
    HAS_WEIGHTS = True

    def __init__(self, workflow, **kwargs):
        super(Forward, self).__init__(workflow, **kwargs)
        self.weights_stddev = kwargs.get("weights_stddev")
        self.bias_stddev = kwargs.get("bias_stddev", self.weights_stddev)
        self.weights_filling = kwargs.get("weights_filling", "uniform")
        self.bias_filling = kwargs.get("bias_filling", "uniform")
        self.rand = kwargs.get("rand", prng.get())
        self.weights_transposed = kwargs.get("weights_transposed", False)
        self.include_bias = kwargs.get("include_bias", True)
        self.weights = Array()
        self.bias = Array()
        self.exports = ["weights", "bias", "include_bias", "weights_transposed"]

    def init_unpickled(self):
        super(Forward, self).init_unpickled()
        self.weights_bf16_ = None   # bf16 shadow of the weights TMA-loaded by the GEMMs
        self.job_weights_ = None

    # -- weights ---------------------------------------------------------------------------------------
    def fill_array(self, filling, arr, stddev):
        if filling == "uniform":
            self.rand.fill(arr, -stddev, stddev)
        elif filling == "gaussian":
            self.rand.fill_normal_real(arr, 0, stddev)
        elif filling == "constant":
            arr[:] = stddev
        else:
            raise ValueError("Invalid filling type %s" % filling)

    def default_weights_stddev(self, fan_in, fan_out):
        return float(min(numpy.sqrt(6.0 / (fan_in + fan_out)), 0.5))

    def create_weights(self, n_out, n_in):
        if self.weights and self.weights.shape == ((n_in, n_out) if self.weights_transposed else (n_out, n_in)):
            return  # restored from a snapshot or linked from another unit
        if self.weights_stddev is None:
            self.weights_stddev = self.default_weights_stddev(n_in, n_out)
        if self.bias_stddev is None:
            self.bias_stddev = self.weights_stddev
        w = numpy.zeros((n_out, n_in), dtype=numpy.float32)
        self.fill_array(self.weights_filling, w, self.weights_stddev)
        if self.weights_transposed:
            w = numpy.ascontiguousarray(w.T)
        self.weights.reset(w)
        if self.include_bias:
            b = numpy.zeros(n_out, dtype=numpy.float32)
            self.fill_array(self.bias_filling, b, self.bias_stddev)
            self.bias.reset(b)

    def place_parameters(self):
        """Move weights/bias into the device parameter arena and create the bf16 shadow."""
        import torch
        from ..ops import kernels
        if self.weights_transposed and not getattr(self, "TRANSPOSED_OK", False):
            raise NotImplementedError("%s: weights_transposed=True is supported on the numpy backend only"
                                      % type(self).__name__)
        arena = find_arena(self, self.device_)
        if self.weights.devmem is None:
            self.weights.initialize(self.device_, storage=arena.alloc("weights", self.name + ".weights",
                                                                      self.weights.size, torch.float32))
            if self.include_bias and self.bias:
                self.bias.initialize(self.device_, storage=arena.alloc("weights", self.name + ".bias",
                                                                       self.bias.size, torch.float32))
        n_out, n_in = self.weights.shape[0], self.weights.size // self.weights.shape[0]
        pitch = (n_in + 7) // 8 * 8
        if self.weights_bf16_ is None:
            buf = arena.alloc("weights_bf16", self.name + ".weights_bf16", n_out * pitch, torch.bfloat16)
            self.weights_bf16_ = buf.view(n_out, pitch)[:, :n_in]
        self.refresh_shadow()

    def refresh_shadow(self):
        """fp32 master -> bf16 shadow (after loading a snapshot or applying external weights)."""
        from ..ops import kernels
        self.weights.unmap()
        w2d = self.weights.devmem.view(self.weights_bf16_.shape[0], -1)
        kernels.copy_2d(w2d, self.weights_bf16_)

    def package_export(self):
        data = {}
        for attr in self.exports:
            v = getattr(self, attr)
            if isinstance(v, Array):
                v.map_read()
                if v:
                    data[attr] = v.mem
            else:
                data[attr] = v
        data["activation_mode"] = {v: k for k, v in ACTIVATIONS.items()}[self.ACTIVATION]
        return data

    # -- job-level master/slave protocol (the reference's asynchronous parameter-server mode) --------------
    def generate_data_for_slave(self, slave):
        """Master: ship the current weights with every job."""
        self.weights.map_read()
        data = {"weights": self.weights.mem.copy()}
        if self.include_bias and self.bias:
            self.bias.map_read()
            data["bias"] = self.bias.mem.copy()
        return data

    def generate_data_for_master(self):
        return None

    def apply_data_from_master(self, data):
        """Slave: adopt the job's weights (remembered so the GD unit can return a delta)."""
        self.weights.map_invalidate()
        numpy.copyto(self.weights.mem, data["weights"])
        self.weights.unmap()
        if "bias" in data and self.bias:
            self.bias.map_invalidate()
            numpy.copyto(self.bias.mem, data["bias"])
            self.bias.unmap()
        self.job_weights_ = (data["weights"], data.get("bias"))
        if self.on_device and self.weights_bf16_ is not None:
            self.refresh_shadow()

    def apply_data_from_slave(self, data, slave):
        pass

    def drop_slave(self, slave):
        pass


class GDBase(AcceleratedUnit):
    """Base of every backward unit: err_output -> err_input, derivative fusion flags."""
    hide_from_registry = True
    ACTIVATION = 0
    VIEW_GROUP = "TRAINER"
    MAPPING = set()

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "TRAINER")
        super(GDBase, self).__init__(workflow, **kwargs)
        self.err_output = None
        self.err_input = Array(shallow_pickle=True)
        self.input = None
        self.output = None
        self.demand("err_output")
        self.need_err_input = kwargs.get("need_err_input", True)
        self.prev_activation = kwargs.get("prev_activation", 0)
        self.derivative_applied = kwargs.get("derivative_applied", False)
        self.forward = kwargs.get("forward", None)
        self.batch_size = None

    def alloc_err_input(self):
        if self.need_err_input and (not self.err_input or self.err_input.shape != self.input.shape):
            self.err_input.reset(numpy.zeros(self.input.shape, dtype=numpy.float32))

    def generate_data_for_slave(self, slave):
        return None

    def generate_data_for_master(self):
        return None

    def apply_data_from_master(self, data):
        pass

    def apply_data_from_slave(self, data, slave):
        pass

    def drop_slave(self, slave):
        pass


class GradientDescentBase(GDBase):
    """Base of backward units with weights: hyper-parameters, gradient buffers, fused update."""
    hide_from_registry = True
    SOLVERS = {"momentum": 0, "sgd": 0, "adagrad": 1, "adadelta": 2}

    def __init__(self, workflow, **kwargs):
        super(GradientDescentBase, self).__init__(workflow, **kwargs)
        self.weights = None
        self.bias = None
        self.learning_rate = kwargs.get("learning_rate", 0.01)
        self.learning_rate_bias = kwargs.get("learning_rate_bias", self.learning_rate)
        self.weights_decay = kwargs.get("weights_decay", 0.00005)
        self.weights_decay_bias = kwargs.get("weights_decay_bias", 0.0)
        self.l1_vs_l2 = kwargs.get("l1_vs_l2", 0.0)
        self.l1_vs_l2_bias = kwargs.get("l1_vs_l2_bias", self.l1_vs_l2)
        self.gradient_moment = kwargs.get("gradient_moment", 0.0)
        self.gradient_moment_bias = kwargs.get("gradient_moment_bias", self.gradient_moment)
        self.accumulate_gradient = kwargs.get("accumulate_gradient", False)
        self.apply_gradient = kwargs.get("apply_gradient", True)
        self.factor_ortho = kwargs.get("factor_ortho", 0.0)
        self.fast_learning_rate = kwargs.get("fast_learning_rate", None)
        self.adagrad_epsilon = kwargs.get("adagrad_epsilon", 1e-6)
        self.adadelta_momentum = kwargs.get("adadelta_momentum", 0.9)
        self.adadelta_adom = kwargs.get("adadelta_adom", 0.3)
        self.adadelta_epsilon = kwargs.get("adadelta_epsilon", 1e-6)
        solvers = kwargs.get("solvers", set())
        if isinstance(solvers, str):
            solvers = {solvers}
        self.solvers = set(solvers) or {"momentum"}
        self.variant_gradient = kwargs.get("variant_gradient", True)
        self.variant_moment_gradient = kwargs.get("variant_moment_gradient", True)
        self.last_minibatch = kwargs.get("last_minibatch", None)
        # host-side optimizer state (kept as Arrays so snapshots carry momentum buffers)
        self.gradient_weights = Array(shallow_pickle=True)
        self.gradient_bias = Array(shallow_pickle=True)
        self.gradient_weights_with_moment = Array()
        self.gradient_bias_with_moment = Array()
        self.accumulated_gradient_weights = Array()
        self.accumulated_gradient_bias = Array()


    @property
    def solver_code(self):
        for name in ("adadelta", "adagrad"):
            if name in self.solvers:
                return self.SOLVERS[name]
        return 0

    def init_unpickled(self):
        super(GradientDescentBase, self).init_unpickled()
        self.dp_group_ = None          # veles_b200.parallel.DataParallel when world_size > 1
        self.col_scratch_ = None

    @property
    def current_batch_size(self):
        bs = self.batch_size
        if bs is None:
            return self.err_output.shape[0]
        return int(bs)

    # -- numpy oracle of the update rule -----------------------------------------------------------------------
    def eff(self, lr, moment):
        """(learning rate, momentum) after the Znicz variant switches: ``variant_gradient = False`` drops the plain
        gradient step (the weights then only coast on their momentum), ``variant_moment_gradient = False`` drops the
        momentum term."""
        return (lr if self.variant_gradient else 0.0), (moment if self.variant_moment_gradient else 0.0)

    def ortho_term(self, w):
        """``factor_ortho`` regulariser (decorrelates the rows of a weight matrix): every row is pulled towards the
        sum of the OTHER rows' negatives, ``g_i += factor_ortho / n_rows * (sum_k w_k - w_i)``.  Weights only
        (2-D); the transposed layout sums along the other axis."""
        if not self.factor_ortho or w.ndim != 2:
            return None
        transposed = bool(getattr(self.forward, "weights_transposed", False)) if self.forward is not None else False
        if transposed:
            return (w.sum(axis=1, keepdims=True) - w) * (self.factor_ortho / w.shape[1])
        return (w.sum(axis=0, keepdims=True) - w) * (self.factor_ortho / w.shape[0])

    def numpy_update(self, w, g, vel, acc, lr, wd, l12, moment):
        lr, moment = self.eff(lr, moment)
        g = g + wd * ((1.0 - l12) * w + 0.5 * l12 * numpy.sign(w))
        ortho = self.ortho_term(w)
        if ortho is not None:
            g = g + ortho
        solver = self.solver_code
        if solver == 0:
            vel[...] = moment * vel - lr * g
            w += vel
        elif solver == 1:
            acc += g * g
            w -= lr * g / numpy.sqrt(acc + self.adagrad_epsilon)
        else:
            rho = self.adadelta_momentum
            acc[...] = rho * acc + (1 - rho) * g * g
            d = -numpy.sqrt(vel + self.adadelta_epsilon) / numpy.sqrt(acc + self.adadelta_epsilon) * g
            vel[...] = rho * vel + (1 - rho) * d * d
            w += d

    def warn_inert_options(self):
        """``fast_learning_rate`` is the one Znicz hyper-parameter whose meaning the checkout does not document
        (``manualrst_veles_workflow_parameters.rst:556-578`` only lists the name): a run that sets it is refused
        rather than trained with a silently different rule."""
        if self.fast_learning_rate is not None:
            raise ValueError("%s: fast_learning_rate=%r is not supported (its semantics are not documented in the "
                             "reference checkout); remove it from the layer's '<-' parameters"
                             % (self.name, self.fast_learning_rate))

    def _ensure_state(self, like, vel, acc):
        self.warn_inert_options()
        if not vel or vel.shape != like.shape:
            vel.reset(numpy.zeros(like.shape, dtype=numpy.float32))
        if self.solver_code != 0 and (not acc or acc.shape != like.shape):
            acc.reset(numpy.zeros(like.shape, dtype=numpy.float32))

    def place_state(self):
        """Device: gradient / velocity buffers mirror the weights' arena layout."""
        import torch
        arena = find_arena(self, self.device_)
        for w, grad, vel, acc in ((self.weights, self.gradient_weights, self.gradient_weights_with_moment,
                                   self.accumulated_gradient_weights),
                                  (self.bias, self.gradient_bias, self.gradient_bias_with_moment,
                                   self.accumulated_gradient_bias)):
            if w is None or not w:
                continue
            self._ensure_state(w, vel, acc)
            if not grad or grad.shape != w.shape:
                grad.reset(numpy.zeros(w.shape, dtype=numpy.float32))
            try:
                arena.locate(w.devmem)
                mirrored = True
            except (ValueError, AttributeError):
                mirrored = False
            for kind, arr in (("grads", grad), ("velocity", vel), ("accum", acc)):
                if arr and arr.devmem is None:
                    if mirrored:
                        arr.initialize(self.device_, storage=arena.mirror(kind, w.devmem.view(-1)))
                    else:
                        arr.initialize(self.device_, storage=arena.alloc(kind, "%s.%s" % (self.name, kind),
                                                                         arr.size, torch.float32))

    def dp_reduce_host(self, *grads):
        """numpy backend under a host-mode DataParallel group: all-reduce the freshly computed gradients."""
        dp = self.dp_group_
        if dp is None or getattr(dp, "mode", None) != "host":
            return grads
        return tuple(dp.allreduce_host(g) for g in grads)

    def grad_tensors(self):
        """(dW, db) device tensors the backward GEMMs accumulate into this step."""
        if self.dp_group_ is not None:
            return self.dp_group_.grads_of(self)
        return (self.gradient_weights.devmem,
                self.gradient_bias.devmem if (self.bias is not None and self.bias) else None)

    def device_bias_sum(self, eo2, gb_t):
        """Bias gradient = column sums of the 2-D err_output.  On the plain single-GPU momentum path the reduction's
        second phase applies the bias GD step itself (``vb_col_sum_gd``): one latency-bound launch less per layer."""
        from ..ops import kernels
        fuse = (self.apply_gradient and self.dp_group_ is None and self.solver_code == 0
                and not self.accumulate_gradient and self.gradient_bias_with_moment
                and self.gradient_bias_with_moment.devmem is not None
                and get(root.common.engine.fuse_bias_update, True))
        if fuse:
            lr_b, mom_b = self.eff(self.learning_rate_bias, self.gradient_moment_bias)
            kernels.col_sum_gd(eo2, gb_t, self.col_scratch_, self.bias.devmem, self.gradient_bias_with_moment.devmem,
                               lr=lr_b, wd=self.weights_decay_bias, l1_vs_l2=self.l1_vs_l2_bias, moment=mom_b)
            self.bias.mark_device_written()
            self.gradient_bias_with_moment.mark_device_written()
        else:
            kernels.col_sum(eo2, gb_t, self.col_scratch_, accumulate=True)
            self.gradient_bias.mark_device_written()
        self.bias_updated_ = fuse

    def device_bias_from_partials(self, gb_t, partials, slabs):
        """The bias gradient arrives as ``slabs`` partial rows (written by the kernel that produced err_output): the
        second phase of the column-sum kernels folds them, and applies the bias step where :meth:`device_bias_sum`
        would have fused it (leaving the gradient buffer zeroed like it does)."""
        from ..ops import kernels
        fuse = (self.apply_gradient and self.dp_group_ is None and self.solver_code == 0
                and not self.accumulate_gradient and self.gradient_bias_with_moment
                and self.gradient_bias_with_moment.devmem is not None
                and get(root.common.engine.fuse_bias_update, True))
        if fuse:
            lr_b, mom_b = self.eff(self.learning_rate_bias, self.gradient_moment_bias)
            kernels.col_sum_finish(partials, slabs, gb_t.view(-1), self.bias.devmem.view(-1),
                                   self.gradient_bias_with_moment.devmem.view(-1), lr=lr_b, wd=self.weights_decay_bias,
                                   l1_vs_l2=self.l1_vs_l2_bias, moment=mom_b)
            self.bias.mark_device_written()
            self.gradient_bias_with_moment.mark_device_written()
        else:
            kernels.col_sum_finish(partials, slabs, gb_t.view(-1))
            self.gradient_bias.mark_device_written()
        self.bias_updated_ = fuse

    def device_ortho(self):
        """Device form of :meth:`ortho_term`: adds the regulariser to the weight gradient of this step (column sums of
        the fp32 master, then one element-wise pass).  In a data-parallel group the exchanged gradients are summed,
        so each rank contributes its share."""
        if not self.factor_ortho or self.weights is None or not self.weights:
            return
        import torch
        from ..ops import kernels
        gw_t, _ = self.grad_tensors()
        rows = self.weights.shape[0]
        w2 = self.weights.devmem.view(rows, -1)
        g2 = gw_t.view(rows, -1)
        transposed = bool(getattr(self.forward, "weights_transposed", False)) if self.forward is not None else False
        share = 1.0
        dp = self.dp_group_
        if dp is not None and getattr(dp, "mode", "") != "host" and getattr(dp, "grad_scale", 1.0) == 1.0:
            share = 1.0 / dp.world_size
        if getattr(self, "_ortho_sums_", None) is None:
            self._ortho_sums_ = torch.zeros(w2.shape[0] if transposed else w2.shape[1], dtype=torch.float32,
                                            device=w2.device)
            self._ortho_scratch_ = torch.empty(kernels.col_sum_scratch_elems(w2.shape[0], w2.shape[1]),
                                               dtype=torch.float32, device=w2.device)
        if transposed:
            kernels.row_sum(w2, self._ortho_sums_)
        else:
            kernels.col_sum(w2, self._ortho_sums_, self._ortho_scratch_)
        kernels.ortho_grad(w2, g2, self._ortho_sums_, self.factor_ortho * share / (w2.shape[1] if transposed
                                                                                   else w2.shape[0]), transposed)

    def device_update(self, forward=None, skip_weights=False):
        """U1: fused single-GPU GD step on weights and bias (or hand over to the DP group).  ``skip_weights``: the
        weight-gradient GEMM already applied the step in its epilogue (``kernels.gemm_fused_gd``)."""
        from ..ops import kernels
        if not self.apply_gradient:
            return
        self.device_ortho()
        if self.dp_group_ is not None:
            self.dp_group_.update_unit(self)
            return
        solver = self.solver_code
        rho = self.adadelta_momentum
        eps = self.adagrad_epsilon if solver == 1 else self.adadelta_epsilon
        fwd = self.forward
        shadow = fwd.weights_bf16_ if fwd is not None else None
        if self.weights is not None and self.weights and not skip_weights:
            w_shadow = shadow if (shadow is not None and shadow.is_contiguous()) else None
            kernels.gd_update(self.weights.devmem, self.gradient_weights.devmem,
                              self.gradient_weights_with_moment.devmem,
                              self.accumulated_gradient_weights.devmem if self.accumulated_gradient_weights else None,
                              w_shadow, lr=self.eff(self.learning_rate, 0)[0], wd=self.weights_decay,
                              l1_vs_l2=self.l1_vs_l2, moment=self.eff(0, self.gradient_moment)[1], solver=solver,
                              eps=eps, rho=rho,
                              zero_grad=not getattr(self, "dw_store_", False))
            if shadow is not None and w_shadow is None:
                kernels.copy_2d(self.weights.devmem.view(shadow.shape[0], -1), shadow)
            self.weights.mark_device_written()
            self.gradient_weights_with_moment.mark_device_written()
        bias_done, self.bias_updated_ = getattr(self, "bias_updated_", False), False
        if self.bias is not None and self.bias and not bias_done:
            kernels.gd_update(self.bias.devmem, self.gradient_bias.devmem, self.gradient_bias_with_moment.devmem,
                              self.accumulated_gradient_bias.devmem if self.accumulated_gradient_bias else None,
                              None, lr=self.eff(self.learning_rate_bias, 0)[0], wd=self.weights_decay_bias,
                              l1_vs_l2=self.l1_vs_l2_bias, moment=self.eff(0, self.gradient_moment_bias)[1],
                              solver=solver, eps=eps,
                              rho=rho, zero_grad=True)
            self.bias.mark_device_written()
            self.gradient_bias_with_moment.mark_device_written()

    # -- job-level master/slave protocol: the slave trains one minibatch locally and returns the weight DELTA,
    #    the master adds deltas as they arrive (asynchronous SGD, like the reference's parameter server)
    def generate_data_for_slave(self, slave):
        return {"learning_rate": self.learning_rate, "learning_rate_bias": self.learning_rate_bias,
                "weights_decay": self.weights_decay, "weights_decay_bias": self.weights_decay_bias,
                "gradient_moment": self.gradient_moment, "gradient_moment_bias": self.gradient_moment_bias}

    def apply_data_from_master(self, data):
        for k, v in (data or {}).items():
            setattr(self, k, v)

    def generate_data_for_master(self):
        fwd = self.forward
        if fwd is None or fwd.job_weights_ is None or self.weights is None or not self.weights:
            return None
        self.weights.map_read()
        w0, b0 = fwd.job_weights_
        data = {"dw": self.weights.mem - w0}
        if self.bias is not None and self.bias and b0 is not None:
            self.bias.map_read()
            data["db"] = self.bias.mem - b0
        return data

    def apply_data_from_slave(self, data, slave):
        if not data or self.weights is None or not self.weights:
            return
        self.weights.map_write()
        self.weights.mem += data["dw"]
        if "db" in data and self.bias is not None and self.bias:
            self.bias.map_write()
            self.bias.mem += data["db"]

    def drop_slave(self, slave):
        pass
