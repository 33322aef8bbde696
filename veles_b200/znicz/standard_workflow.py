"""StandardWorkflow: builds a trainable workflow from a ``layers`` configuration.

Specified by ``docs/source/manualrst_veles_workflow_creation.rst:115-143`` (the default
``create_workflow``) and ``:274-645`` (link rules), config keys by
``docs/source/manualrst_veles_workflow_parameters.rst:38-79, 104-192, 420-445`` —
``layers=[{"type": ..., "->": {forward kwargs}, "<-": {gd kwargs}}]``, ``loss_function`` or
(``decision_name``, ``evaluator_name``), ``loader_name`` + ``loader_config``,
``decision_config``, ``snapshotter_config``, ``mcdnnic_topology`` strings such as
``"12x256x256-32C4-MP2-64C4-MP3-32N-4N"`` (``:583-599``).

Topology (one iteration = one minibatch):

    Repeater -> Loader -> Forward[0..L-1] -> Evaluator -> Decision -> Snapshotter
        ^                                                               |
        +---- GD[0] <- ... <- GD[L-1] <---------------------------------+
    EndPoint (gate_block = ~decision.complete)
"""
from __future__ import annotations

import numpy

import re

from ..config import root, get, Config
from ..loader.base import UserLoaderRegistry
from ..mutable import Bool
from ..plumbing import Repeater
from ..snapshotter import SnapshotterRegistry, SnapshotterToFile
from ..unit_registry import UnitRegistry
from .nn_units import NNWorkflow, Forward, GradientDescentBase


def _cfg_to_dict(cfg):
    if cfg is None:
        return {}
    if isinstance(cfg, Config):
        return cfg.to_dict()
    return dict(cfg)


class LayerRegistry(object):
    """type string -> (Forward class, GD class) discovered from ``MAPPING`` sets."""

    @staticmethod
    def _scan(base):
        found = {}
        for cls in list(UnitRegistry.units) + list(UnitRegistry.hidden_units):
            mapping = cls.__dict__.get("MAPPING")
            if isinstance(mapping, (set, frozenset)) and issubclass(cls, base):
                for m in mapping:
                    found[m] = cls
        return found

    @staticmethod
    def forwards():
        LayerRegistry._import_all()
        from .nn_units import Forward, ForwardBase
        return LayerRegistry._scan(ForwardBase)

    @staticmethod
    def gds():
        LayerRegistry._import_all()
        from .nn_units import GradientDescentBase, GDBase
        return LayerRegistry._scan(GDBase)

    @staticmethod
    def _import_all():
        from . import all2all, gd  # noqa: F401
        for mod in ("conv", "gd_conv", "pooling", "gd_pooling", "activation", "dropout", "normalization",
                    "cutter", "deconv", "gd_deconv", "depooling", "channel_splitting", "resizable_all2all",
                    "rprop_all2all", "weights_zerofilling", "lr_adjust", "recurrent"):
            try:
                __import__("%s.%s" % (__package__, mod))
            except ImportError:
                pass


def parse_mcdnnic_topology(topology, parameters=None):
    """"12x256x256-32C4-MP2-64C4-MP3-32N-4N" -> layers list (conv_tanh/max_pooling/all2all_tanh/softmax)."""
    parameters = parameters or {}
    fwd_p = dict(parameters.get("->", {}))
    gd_p = dict(parameters.get("<-", {}))
    parts = topology.split("-")
    if not re.match(r"^\d+x\d+x\d+$", parts[0]):
        raise ValueError("mcdnnic_topology must start with CxHxW, got %r" % parts[0])
    layers = []
    body = parts[1:]
    for i, tok in enumerate(body):
        m = re.match(r"^(\d+)C(\d+)$", tok)
        if m:
            k = int(m.group(2))
            layers.append({"type": "conv_tanh", "->": dict(fwd_p, n_kernels=int(m.group(1)), kx=k, ky=k),
                           "<-": dict(gd_p)})
            continue
        m = re.match(r"^MP(\d+)$", tok)
        if m:
            k = int(m.group(1))
            layers.append({"type": "max_pooling", "->": {"kx": k, "ky": k, "sliding": (k, k)}})
            continue
        m = re.match(r"^(\d+)N$", tok)
        if m:
            last = i == len(body) - 1
            layers.append({"type": "softmax" if last else "all2all_tanh",
                           "->": dict(fwd_p, output_sample_shape=int(m.group(1))), "<-": dict(gd_p)})
            continue
        raise ValueError("Unrecognised mcdnnic token %r" % tok)
    return layers


class StandardWorkflowBase(NNWorkflow):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(StandardWorkflowBase, self).__init__(workflow, **kwargs)
        self.layers = kwargs.get("layers")
        topo = kwargs.get("mcdnnic_topology")
        if topo is not None:
            self.layers = parse_mcdnnic_topology(topo, kwargs.get("mcdnnic_parameters"))
        if isinstance(self.layers, Config) or self.layers is None:
            raise ValueError("layers (or mcdnnic_topology) must be defined for %s" % type(self).__name__)
        self.layers = [_cfg_to_dict(l) if not isinstance(l, dict) else l for l in self.layers]
        self.loader_name = kwargs.get("loader_name")
        self.loader_factory = kwargs.get("loader_factory")
        self.loader_config = _cfg_to_dict(kwargs.get("loader_config", kwargs.get("loader")))
        self.preprocessing = kwargs.get("preprocessing", False)
        self.forwards = []
        self.gds = []
        self.loader = None
        self.repeater = None

    def link_repeater(self, *parents):
        self.repeater = Repeater(self)
        self.repeater.link_from(*parents)
        return self.repeater

    def link_loader(self, *parents):
        if self.loader_factory is not None:
            self.loader = self.loader_factory(self, **self.loader_config)
        elif self.loader_name is not None:
            self.loader = UserLoaderRegistry.get_factory(self.loader_name)(self, **self.loader_config)
        else:
            raise ValueError("loader_name (or loader_factory) must be defined")
        self.loader.link_from(*parents)
        return self.loader

    def link_end_point(self, *parents):
        self.end_point.link_from(*parents)
        return self.end_point

    # -- forwards -------------------------------------------------------------------------------------------
    def _layer_kwargs(self, layer, key):
        kw = dict(layer.get(key, {}))
        # flat (legacy) layer dicts: every non-reserved key belongs to the forward unit
        if key == "->":
            for k, v in layer.items():
                if k not in ("type", "name", "->", "<-"):
                    kw.setdefault(k, v)
        return kw

    def link_forwards(self, init_attrs, *parents):
        registry = LayerRegistry.forwards()
        self.forwards = []
        prev = None
        n_layers = len(self.layers)
        for i, layer in enumerate(self.layers):
            tpe = layer["type"]
            if tpe not in registry:
                raise ValueError("Unknown layer type %r; known: %s" % (tpe, sorted(registry)))
            kw = self._layer_kwargs(layer, "->")
            if "name" in layer:
                kw["name"] = layer["name"]
            kw.setdefault("name", "%s_%d" % (tpe, i))
            if i == n_layers - 1 and issubclass(registry[tpe], Forward) and \
                    "output_sample_shape" not in kw and hasattr(registry[tpe], "neurons_number") \
                    and tpe not in ("conv", ):
                kw["output_sample_shape"] = self._last_layer_size
            unit = registry[tpe](self, **kw)
            if prev is None:
                unit.link_from(*parents)
                unit.link_attrs(parents[0], init_attrs)
            else:
                unit.link_from(prev)
                unit.link_attrs(prev, ("input", "output"))
            for attr in getattr(unit, "LOADER_ATTRS", ()):
                unit.link_attrs(self.loader, attr)
            self.forwards.append(unit)
            prev = unit
        return prev

    @property
    def _last_layer_size(self):
        loader = self.loader
        targets = getattr(loader, "targets_shape", None)
        if targets:
            n = 1
            for d in targets:
                n *= d
            return n
        return _DeferredLabelsCount(loader)


class _DeferredLabelsCount(object):
    """The class count is known only after the loader is initialized; All2All resolves it."""

    def __init__(self, loader):
        self.loader = loader

    def __int__(self):
        tgt = getattr(self.loader, "minibatch_targets", None)
        if tgt is not None and tgt:
            return int(tgt.size // tgt.shape[0])
        return int(self.loader.unique_labels_count)

    __index__ = __int__


class _Indexed(object):
    """``value`` = obj.<field>[i][j...] at the moment it is read (plotter inputs that live inside lists)."""

    def __init__(self, obj, field, *index):
        self.obj, self.field, self.index = obj, field, index

    @property
    def value(self):
        v = getattr(self.obj, self.field)
        for i in self.index:
            v = v[i]
        return v


class _FirstRow(object):
    """``value`` = the first sample of an Array, flattened."""

    def __init__(self, arr):
        self.arr = arr

    @property
    def value(self):
        self.arr.map_read()
        return numpy.asarray(self.arr.mem[0]).ravel()


class _KernelShape(object):
    """Sequence-like (ky, kx, channels) of a convolutional unit, resolved when the plotter draws."""

    def __init__(self, unit):
        self.unit = unit

    def _shape(self):
        u = self.unit
        n = u.weights.shape[0] if not getattr(u, "weights_transposed", False) else u.weights.shape[-1]
        per = u.weights.size // max(n, 1)
        c = max(per // (u.ky * u.kx), 1)
        return (u.ky, u.kx, c)

    def __iter__(self):
        return iter(self._shape())

    def __len__(self):
        return 3

    def __getitem__(self, i):
        return self._shape()[i]


class StandardWorkflow(StandardWorkflowBase):
    """Workflow built from configuration; override :meth:`create_workflow` to change topology."""
    hide_from_registry = False
    WORKFLOW_NAME = "StandardWorkflow"

    def __init__(self, workflow, **kwargs):
        super(StandardWorkflow, self).__init__(workflow, **kwargs)
        self.loss_function = kwargs.get("loss_function")
        self.decision_name = kwargs.get("decision_name")
        self.evaluator_name = kwargs.get("evaluator_name")
        if self.loss_function is None and (self.decision_name is None or self.evaluator_name is None):
            raise ValueError("Either loss_function or (decision_name, evaluator_name) must be defined")
        if self.loss_function not in (None, "softmax", "mse"):
            raise ValueError("Unknown loss_function %r (expected \"softmax\" or \"mse\")" % (self.loss_function,))
        self.decision_config = _cfg_to_dict(kwargs.get("decision_config", kwargs.get("decision")))
        self.snapshotter_config = _cfg_to_dict(kwargs.get("snapshotter_config", kwargs.get("snapshotter")))
        self.evaluator_config = _cfg_to_dict(kwargs.get("evaluator_config", kwargs.get("evaluator")))
        self.lr_adjuster_config = _cfg_to_dict(kwargs.get("lr_adjuster_config", kwargs.get("lr_adjuster")))
        self.image_saver_config = _cfg_to_dict(kwargs.get("image_saver_config"))
        self.data_saver_config = _cfg_to_dict(kwargs.get("data_saver_config"))
        self.publisher_config = _cfg_to_dict(kwargs.get("publisher_config", kwargs.get("publisher")))
        self.result_unit_config = _cfg_to_dict(kwargs.get("result_unit_config"))
        self.fuse_derivatives = kwargs.get("fuse_derivatives", True)
        self.evaluator = None
        self.decision = None
        self.snapshotter = None
        self.lr_adjuster = None
        self.meandispnorm = None
        self.avatar = None
        self.publisher = None
        self.plotters = []
        self.create_workflow()

    def create_workflow(self):
        self.link_repeater(self.start_point)
        self.link_loader(self.repeater)
        self.link_forwards(("input", "minibatch_data"), self.loader)
        self.link_evaluator(self.forwards[-1])
        self.link_decision(self.evaluator)
        self.link_snapshotter(self.decision)
        self.link_gds(self.snapshotter)
        last = self.gds[0]
        if self.lr_adjuster_config:
            last = self.link_lr_adjuster(last)
        self.link_loop(last)
        self.link_end_point(self.snapshotter)

    # -- evaluator / decision / snapshotter -----------------------------------------------------------------------
    def _loss(self):
        if self.loss_function is not None:
            return self.loss_function
        return "mse" if "mse" in (self.evaluator_name or "") else "softmax"

    def link_evaluator(self, *parents):
        from .evaluator import EvaluatorSoftmax, EvaluatorMSE
        last = self.forwards[-1]
        if self._loss() == "softmax":
            self.evaluator = EvaluatorSoftmax(self, softmax_unit=last if hasattr(last, "max_idx") else None,
                                              **self.evaluator_config)
            self.evaluator.link_attrs(self.loader, ("labels", "minibatch_labels"), ("batch_size", "minibatch_size"))
            if hasattr(last, "max_idx"):
                self.evaluator.link_attrs(last, "max_idx")
        else:
            self.evaluator = EvaluatorMSE(self, **self.evaluator_config)
            self.evaluator.link_attrs(self.loader, ("target", "minibatch_targets"),
                                      ("batch_size", "minibatch_size"))
            if getattr(self.loader, "shard_world", 1) > 1:
                self.evaluator.link_attrs(self.loader, ("local_batch_size", "local_count"))
        self.evaluator.link_attrs(last, "output")
        self.evaluator.link_from(*parents)
        return self.evaluator

    def link_decision(self, *parents):
        from .decision import DecisionGD, DecisionMSE
        cls = DecisionGD if self._loss() == "softmax" else DecisionMSE
        self.decision = cls(self, **self.decision_config)
        self.decision.link_attrs(self.loader, "minibatch_class", "minibatch_size", "last_minibatch",
                                 "class_lengths", "epoch_ended", "epoch_number")
        if self._loss() == "softmax":
            self.decision.link_attrs(self.evaluator, ("minibatch_n_err", "n_err"),
                                     ("minibatch_confusion_matrix", "confusion_matrix"),
                                     ("minibatch_max_err_y_sum", "max_err_output_sum"),
                                     ("minibatch_loss_sum", "loss_sum"))
        else:
            self.decision.link_attrs(self.evaluator, ("minibatch_metrics", "metrics"), ("minibatch_n_err", "n_err"))
        self.decision.link_from(*parents)
        self.repeater.gate_block = self.decision.complete
        return self.decision

    def link_snapshotter(self, *parents):
        cfg = dict(self.snapshotter_config)
        kind = cfg.pop("type", "file")
        cfg.setdefault("prefix", type(self).__name__.lower())
        self.snapshotter = SnapshotterRegistry.snapshotters.get(kind, SnapshotterToFile)(self, **cfg)
        self.snapshotter.link_from(*parents)
        self.snapshotter.link_attrs(self.decision, ("suffix", "snapshot_suffix"))
        self.snapshotter.gate_skip = ~(self.loader.epoch_ended & self.decision.improved)
        self.snapshotter.skip = get(root.common.disable.snapshotting, False) or self.snapshotter.skip
        return self.snapshotter

    # -- backward ------------------------------------------------------------------------------------------------------
    def link_gds(self, *parents):
        registry = LayerRegistry.gds()
        self.gds = [None] * len(self.forwards)
        first_trainable = 0
        prev_gd = None
        for i in reversed(range(len(self.forwards))):
            fwd = self.forwards[i]
            layer = self.layers[i]
            tpe = layer["type"]
            if tpe not in registry:
                raise ValueError("Layer type %r has no backward unit; known: %s" % (tpe, sorted(registry)))
            kw = dict(layer.get("<-", {}))
            kw.setdefault("name", "gd_" + fwd.name)
            kw["forward"] = fwd
            # nothing upstream needs the error w.r.t. the data
            if i == 0:
                kw.setdefault("need_err_input", False)
            gd = registry[tpe](self, **kw)
            self.gds[i] = gd
            if prev_gd is None:
                gd.link_from(*parents)
                gd.link_attrs(self.evaluator, "err_output")
            else:
                gd.link_from(prev_gd)
                gd.link_attrs(prev_gd, ("err_output", "err_input"))
            gd.link_attrs(fwd, "input", "output")
            if getattr(fwd, "HAS_WEIGHTS", False):
                gd.link_attrs(fwd, "weights", "bias")
            for attr in getattr(gd, "FORWARD_ATTRS", ()):
                gd.link_attrs(fwd, attr)
            gd.link_attrs(self.loader, ("batch_size", "minibatch_size"))
            gd.gate_skip = self.decision.gd_skip
            if prev_gd is None and not self.is_slave:
                # once training is complete the backward chain must not start (EndPoint is stopping the units)
                gd.gate_block = self.decision.complete
            prev_gd = gd
        if self.fuse_derivatives:
            self._plan_derivative_fusion()
        self._plan_lrn_pool_fusion()
        self._plan_bias_sum_fusion()
        return self.gds[0]

    def _plan_lrn_pool_fusion(self):
        """norm -> max pooling pairs: the pooling layer's backward unit hands its work to the LRN backward unit, whose
        device kernel scatters the pooling gradient by the argmax bytes and applies the LRN derivative in one pass
        (the dense gradient of the LRN output never touches memory).  Decided per run on the device; the numpy
        backend and unsupported geometries keep the two units separate."""
        from .normalization import LRNormalizerForward, LRNormalizerBackward
        from .gd_pooling import GDPooling
        for i in range(len(self.forwards) - 1):
            f, nxt = self.forwards[i], self.forwards[i + 1]
            gd, gd_pool = self.gds[i], self.gds[i + 1]
            if (isinstance(f, LRNormalizerForward) and isinstance(gd, LRNormalizerBackward)
                    and isinstance(gd_pool, GDPooling) and getattr(nxt, "KIND", "") in ("max", "maxabs")
                    and gd.need_err_input):
                gd.pool_gd = gd_pool
                gd_pool.lrn_consumer = gd

    def _plan_bias_sum_fusion(self):
        """conv -> LRN: the LRN backward kernel writes the convolution's err_output; it also accumulates its column
        sums (= the convolution's bias gradient), so the convolution's GD unit does not read that tensor again.
        Needs the derivative of the convolution's activation to be applied by the LRN unit (or to be absent)."""
        from .gd_conv import GradientDescentConv
        from .normalization import LRNormalizerBackward
        if not get(root.common.engine.fuse_bias_sum, True):
            return
        for i in range(1, len(self.forwards)):
            prev_f, gd, prev_gd = self.forwards[i - 1], self.gds[i], self.gds[i - 1]
            if not (isinstance(gd, LRNormalizerBackward) and isinstance(prev_gd, GradientDescentConv)):
                continue
            if not gd.need_err_input or not getattr(prev_f, "include_bias", True):
                continue
            act = getattr(prev_f, "ACTIVATION", 0)
            if act and not (gd.prev_activation == act and getattr(prev_gd, "derivative_applied", False)):
                continue
            gd.bias_sum_for = prev_gd

    def _plan_derivative_fusion(self):
        """gd[i] multiplies its err_input by f'_{i-1}(input) when forward[i-1] has an activation that
        differentiates through its output; gd[i-1] then skips its own derivative pass."""
        for i in range(1, len(self.forwards)):
            prev_f, gd, prev_gd = self.forwards[i - 1], self.gds[i], self.gds[i - 1]
            act = getattr(prev_f, "ACTIVATION", 0)
            if act and getattr(gd, "CAN_FUSE_PREV_DERIVATIVE", False) and gd.need_err_input \
                    and getattr(prev_gd, "ACTIVATION", 0) == act and hasattr(prev_gd, "derivative_applied"):
                gd.prev_activation = act
                prev_gd.derivative_applied = True

    def link_loop(self, *parents):
        self.repeater.link_from(*parents)
        return self.repeater

    def link_end_point(self, *parents):
        if self.is_slave and self.gds:
            parents = (self.gds[0],)  # a slave job ends after the LAST backward unit of its single pass
        self.end_point.link_from(*parents)
        self.end_point.gate_block = ~self.decision.complete
        return self.end_point

    def link_lr_adjuster(self, *parents):
        from .lr_adjust import LearningRateAdjust
        self.lr_adjuster = LearningRateAdjust(self, **self.lr_adjuster_config)
        for gd in self.gds:
            if hasattr(gd, "learning_rate"):
                self.lr_adjuster.add_gd_unit(gd)
        self.lr_adjuster.link_from(*parents)
        self.lr_adjuster.gate_skip = self.decision.gd_skip
        return self.lr_adjuster

    def link_meandispnorm(self, *parents):
        from ..mean_disp_normalizer import MeanDispNormalizer
        self.meandispnorm = MeanDispNormalizer(self)
        self.meandispnorm.link_attrs(self.loader, ("input", "minibatch_data"), "mean", "rdisp")
        self.meandispnorm.link_from(*parents)
        return self.meandispnorm

    def link_avatar(self, *extra_attrs):
        from ..avatar import Avatar
        self.avatar = Avatar(self)
        self.avatar.reals[self.loader] = list(
            ("minibatch_data", "minibatch_labels", "minibatch_indices") + tuple(extra_attrs))
        self.avatar.clone()
        self.avatar.link_from(self.loader)
        return self.avatar

    def link_ipython(self, *parents):
        from ..interaction import Shell
        self.ipython = Shell(self)
        self.ipython.link_from(*parents)
        self.ipython.gate_skip = ~self.decision.epoch_ended if self.decision.epoch_ended is not None else Bool(True)
        return self.ipython

    def link_publisher(self, *parents):
        from ..publishing import Publisher
        self.publisher = Publisher(self, **self.publisher_config)
        self.publisher.link_from(*parents)
        self.publisher.gate_skip = ~self.decision.complete
        return self.publisher

    def link_data_saver(self, *parents):
        from ..loader.saver import MinibatchesSaver
        self.data_saver = MinibatchesSaver(self, **self.data_saver_config)
        self.data_saver.link_attrs(self.loader, "shuffle_limit", "minibatch_class", "minibatch_data",
                                   "minibatch_labels", "class_lengths", "max_minibatch_size", "minibatch_size")
        self.data_saver.link_from(*parents)
        return self.data_saver

    def link_error_plotter(self, *parents):
        from ..plotting_units import AccumulatingPlotter
        prev = parents
        self.error_plotters = []
        for klass, style in ((1, "g-"), (2, "b-")):
            p = AccumulatingPlotter(self, name="Errors", plot_style=style, label=("", "validation", "train")[klass])
            p.input = self.decision
            p.input_field = "epoch_n_err_pt"
            p.input_offset = klass
            p.link_from(*prev)
            p.gate_skip = ~self.loader.epoch_ended
            self.error_plotters.append(p)
            self.plotters.append(p)
            prev = (p,)
        return prev[0]

    def link_result_unit(self, *parents):
        return parents[0]

    # -- service units documented in manualrst_veles_workflow_creation.rst:454-472, 550-644 ---------------------------
    def _chain(self, units, parents, gate=None):
        """Link ``units`` one after another behind ``parents``; returns the last one."""
        prev = parents
        for u in units:
            u.link_from(*prev)
            if gate is not None:
                u.gate_skip = gate
            self.plotters.append(u)
            prev = (u,)
        return prev[0]

    def link_image_saver(self, *parents):
        """ImageSaver: dumps (misclassified) samples of every minibatch, by class (``image_saver_config``)."""
        from .image_saver import ImageSaver
        self.image_saver = ImageSaver(self, **self.image_saver_config)
        self.image_saver.link_from(*parents)
        self.image_saver.link_attrs(self.forwards[0], "input")
        self.image_saver.link_attrs(self.loader, ("indices", "minibatch_indices"), ("labels", "minibatch_labels"),
                                    "minibatch_class", "minibatch_size")
        self.image_saver.link_attrs(self.evaluator if self.evaluator is not None else self.forwards[-1], "max_idx")
        if self.snapshotter is not None:
            self.image_saver.gate_skip = ~self.decision.improved
        return self.image_saver

    def link_conf_matrix_plotter(self, *parents):
        from ..plotting_units import MatrixPlotter
        self.conf_matrix_plotters = []
        for klass in (1, 2):
            p = MatrixPlotter(self, name="Confusion matrix (%s)" % ("", "validation", "train")[klass],
                              reversed_labels_mapping=getattr(self.loader, "reversed_labels_mapping", None))
            p.input = _Indexed(self.decision, "confusion_matrixes", klass)
            p.input_field = "value"
            self.conf_matrix_plotters.append(p)
        return self._chain(self.conf_matrix_plotters, parents, ~self.loader.epoch_ended)

    def link_err_y_plotter(self, *parents):
        from ..plotting_units import AccumulatingPlotter
        self.err_y_plotters = []
        for klass, style in ((1, "g-"), (2, "b-")):
            p = AccumulatingPlotter(self, name="Last layer max gradient sum", plot_style=style,
                                    label=("", "validation", "train")[klass], input_field="max_err_y_sums",
                                    input_offset=klass)
            p.input = self.decision
            self.err_y_plotters.append(p)
        return self._chain(self.err_y_plotters, parents, ~self.loader.epoch_ended)

    def link_mse_plotter(self, *parents):
        from ..plotting_units import AccumulatingPlotter
        self.mse_plotters = []
        for klass, style in ((1, "g-"), (2, "b-")):
            p = AccumulatingPlotter(self, name="rmse", plot_style=style, label=("", "validation", "train")[klass],
                                    input_field="epoch_rmse" if hasattr(self.decision, "epoch_rmse") else "epoch_loss",
                                    input_offset=klass)
            p.input = self.decision
            self.mse_plotters.append(p)
        return self._chain(self.mse_plotters, parents, ~self.loader.epoch_ended)

    def link_min_max_plotter(self, is_min, *parents):
        """Per-epoch minimum (``is_min``) or maximum of the per-sample MSE (``DecisionMSE.epoch_metrics``)."""
        from ..plotting_units import AccumulatingPlotter
        plotters = []
        for klass, style in ((1, "g-"), (2, "b-")):
            p = AccumulatingPlotter(self, name="mse %s" % ("min" if is_min else "max"), plot_style=style,
                                    label=("", "validation", "train")[klass], input_field="value")
            p.input = _Indexed(self.decision, "epoch_metrics", klass, 2 if is_min else 1)
            plotters.append(p)
        setattr(self, "min_plotters" if is_min else "max_plotters", plotters)
        return self._chain(plotters, parents, ~self.loader.epoch_ended)

    def link_multi_hist_plotter(self, layers=None, limit=None, *parents):
        """One MultiHistogram (per-neuron weight histograms) per trainable layer."""
        from ..plotting_units import MultiHistogram
        self.multi_hist_plotters = []
        for i, f in enumerate(self.forwards):
            if not getattr(f, "HAS_WEIGHTS", False) or (layers is not None and i not in layers
                                                        and self.layers[i].get("type") not in layers):
                continue
            p = MultiHistogram(self, name="Histogram %s %d" % (self.layers[i].get("type", "layer"), i + 1),
                               hist_number=limit or 8)
            p.input = f.weights
            p.input_field = None
            self.multi_hist_plotters.append(p)
        return self._chain(self.multi_hist_plotters, parents, ~self.loader.epoch_ended)

    def link_weights_plotter(self, layers=None, limit=64, weights_input="weights", *parents):
        """One Weights2D per trainable layer (kernels / neurons as image tiles)."""
        from .nn_plotting_units import Weights2D
        self.weights_plotters = []
        for i, f in enumerate(self.forwards):
            if not getattr(f, "HAS_WEIGHTS", False) or (layers is not None and i not in layers
                                                        and self.layers[i].get("type") not in layers):
                continue
            p = Weights2D(self, name="%s %d %s" % (self.layers[i].get("type", "layer"), i + 1, weights_input),
                          limit=limit, transposed=bool(getattr(f, "weights_transposed", False)))
            p.input = getattr(f, weights_input)
            p.input_field = None
            if hasattr(f, "kx") and hasattr(f, "ky"):
                p.get_shape_from = _KernelShape(f)
            self.weights_plotters.append(p)
        return self._chain(self.weights_plotters, parents, ~self.loader.epoch_ended)

    def link_similar_weights_plotter(self, layers=None, limit=64, magnitude_threshold=0.65, form_threshold=1.1,
                                     peak_threshold=0.5, *parents):
        from .nn_plotting_units import SimilarWeights2D
        self.similar_weights_plotters = []
        for i, f in enumerate(self.forwards):
            if not getattr(f, "HAS_WEIGHTS", False) or (layers is not None and i not in layers
                                                        and self.layers[i].get("type") not in layers):
                continue
            p = SimilarWeights2D(self, name="%s %d similar weights" % (self.layers[i].get("type", "layer"), i + 1),
                                 limit=limit, magnitude_threshold=magnitude_threshold, form_threshold=form_threshold,
                                 peak_threshold=peak_threshold,
                                 transposed=bool(getattr(f, "weights_transposed", False)))
            p.input = f.weights
            p.input_field = None
            if hasattr(f, "kx") and hasattr(f, "ky"):
                p.get_shape_from = _KernelShape(f)
            self.similar_weights_plotters.append(p)
        return self._chain(self.similar_weights_plotters, parents, ~self.loader.epoch_ended)

    def link_image_plotter(self, *parents):
        """First samples of the current minibatch next to the last layer's outputs."""
        from ..plotting_units import ImagePlotter
        self.image_plotter = ImagePlotter(self, name="output sample")
        self.image_plotter.inputs = [self.loader.minibatch_data, self.forwards[-1].output]
        self.image_plotter.input_fields = [None, None]
        return self._chain([self.image_plotter], parents, ~self.loader.epoch_ended)

    def link_immediate_plotter(self, *parents):
        """Target, output and input of the first sample of the minibatch as curves (MSE workflows)."""
        from ..plotting_units import ImmediatePlotter
        self.immediate_plotter = ImmediatePlotter(self, name="ImmediatePlotter")
        p = self.immediate_plotter
        for obj, style in ((getattr(self.loader, "minibatch_targets", None), "g-"),
                           (self.forwards[-1].output, "r-"), (self.loader.minibatch_data, "b-")):
            if obj is None:
                continue
            p.inputs.append(_FirstRow(obj))
            p.input_fields.append("value")
            p.input_styles.append(style)
        return self._chain([p], parents, ~self.loader.epoch_ended)

    def link_table_plotter(self, *parents):
        """Max / min of every layer's weights, biases, outputs and gradients."""
        from ..plotting_units import TableMaxMin
        self.table_plotter = TableMaxMin(self, name="Max, Min")
        t = self.table_plotter
        for i, f in enumerate(self.forwards):
            tag = "%s %d" % (self.layers[i].get("type", "layer"), i + 1)
            if getattr(f, "HAS_WEIGHTS", False):
                t.y += [f.weights, f.bias]
                t.row_labels += [tag + " weights", tag + " bias"]
            t.y.append(f.output)
            t.row_labels.append(tag + " output")
        for i, g in enumerate(self.gds or []):
            if getattr(g, "gradient_weights", None) is not None and g.gradient_weights is not None:
                t.y.append(g.gradient_weights)
                t.row_labels.append("gd %d gradient" % (i + 1))
        return self._chain([t], parents, ~self.loader.epoch_ended)


    # -- helpers --------------------------------------------------------------------------------------------------------
    def extract_forward_workflow(self, loader_name=None, loader_config=None, result_unit_factory=None, cyclic=False):
        """A new workflow with only Loader -> forwards (weights shared by reference) for inference
        (``docs/source/manualrst_veles_example_advanced.rst:352-366``)."""
        from ..units import Unit
        wf = ForwardWorkflow(self.workflow, layers=self.layers, loader_name=loader_name or self.loader_name,
                             loader_config=loader_config if loader_config is not None else self.loader_config,
                             source=self, cyclic=cyclic)
        return wf

    def initialize(self, device=None, **kwargs):
        res = super(StandardWorkflow, self).initialize(device=device, **kwargs)
        if not res:
            # steady-state unit sequences replay as CUDA graphs under the ordinary scheduler (wf.run(), the CLI)
            from ..parallel.step import install_graph_replay
            install_graph_replay(self)
        return res


class ForwardWorkflow(StandardWorkflowBase):
    """Loader -> forward chain only (inference / feature extraction)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        source = kwargs.pop("source", None)
        cyclic = kwargs.pop("cyclic", False)
        super(ForwardWorkflow, self).__init__(workflow, **kwargs)
        self.link_repeater(self.start_point)
        self.link_loader(self.repeater)
        last = self.link_forwards(("input", "minibatch_data"), self.loader)
        for f in self.forwards:
            f.forward_mode = True
        if source is not None:
            for mine, theirs in zip(self.forwards, source.forwards):
                if getattr(theirs, "HAS_WEIGHTS", False):
                    mine.weights = theirs.weights
                    mine.bias = theirs.bias
        if cyclic:
            self.repeater.link_from(last)
            self.end_point.link_from(last)
            self.end_point.gate_block = ~self.loader.epoch_ended
            self.repeater.gate_block = self.loader.epoch_ended
        else:
            self.end_point.link_from(last)
