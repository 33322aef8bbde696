"""Learning-rate schedules applied to every GD unit each training iteration
(``docs/source/manualrst_veles_workflow_parameters.rst:655-680``): policies ``exp``, ``fixed``,
``step_exp``, ``inv``, ``arbitrary_step``."""
from __future__ import annotations

from ..distributable import TriviallyDistributable
from ..units import Unit


class LRPolicy(object):
    def __call__(self, it):
        raise NotImplementedError


class FixedAdjustPolicy(LRPolicy):
    def __init__(self, base_lr, **kwargs):
        self.base_lr = base_lr

    def __call__(self, it):
        return self.base_lr


class ExpPolicy(LRPolicy):
    """lr = base_lr * gamma^it"""

    def __init__(self, base_lr, gamma=0.999, a_ratio=1.0, **kwargs):
        self.base_lr, self.gamma, self.a_ratio = base_lr, gamma, a_ratio

    def __call__(self, it):
        return self.base_lr * (self.gamma ** (self.a_ratio * it))


class StepExpPolicy(LRPolicy):
    """lr = base_lr * gamma^floor(it / step)"""

    def __init__(self, base_lr, gamma=0.1, step=1000, **kwargs):
        self.base_lr, self.gamma, self.step = base_lr, gamma, step

    def __call__(self, it):
        return self.base_lr * (self.gamma ** (it // self.step))


class InvAdjustPolicy(LRPolicy):
    """lr = base_lr * (1 + gamma * it)^(-pow)"""

    def __init__(self, base_lr, gamma=0.0001, pow_ratio=0.75, **kwargs):
        self.base_lr, self.gamma, self.pow_ratio = base_lr, gamma, pow_ratio

    def __call__(self, it):
        return self.base_lr * (1.0 + self.gamma * it) ** (-self.pow_ratio)


class ArbitraryStepPolicy(LRPolicy):
    """lrs_with_lengths = [(lr, n_iterations), ...]; the last lr persists."""

    def __init__(self, base_lr=None, lrs_with_lengths=((1.0, 1),), **kwargs):
        self.lrs_with_lengths = list(lrs_with_lengths)

    def __call__(self, it):
        acc = 0
        for lr, length in self.lrs_with_lengths:
            acc += length
            if it < acc:
                return lr
        return self.lrs_with_lengths[-1][0]


POLICIES = {"exp": ExpPolicy, "fixed": FixedAdjustPolicy, "step_exp": StepExpPolicy, "inv": InvAdjustPolicy,
            "arbitrary_step": ArbitraryStepPolicy}


class LearningRateAdjust(Unit, TriviallyDistributable):
    """Applies learning-rate schedules (exp, inv, step, arbitrary-step policies) to GD units as training
    proceeds."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "TRAINER")
        super(LearningRateAdjust, self).__init__(workflow, **kwargs)
        self.lr_policy_name = kwargs.get("lr_policy_name", kwargs.get("lr_policy"))
        self.bias_lr_policy_name = kwargs.get("bias_lr_policy_name", self.lr_policy_name)
        self.lr_parameters = dict(kwargs.get("lr_parameters", {}))
        self.bias_lr_parameters = dict(kwargs.get("bias_lr_parameters", self.lr_parameters))
        self._gd_units = []
        self._minibatches_count = 0

    def add_gd_unit(self, gd):
        wp = bp = None
        if self.lr_policy_name is not None:
            wp = POLICIES[self.lr_policy_name](gd.learning_rate, **self.lr_parameters)
        if self.bias_lr_policy_name is not None:
            bp = POLICIES[self.bias_lr_policy_name](gd.learning_rate_bias, **self.bias_lr_parameters)
        self._gd_units.append((gd, wp, bp))

    def initialize(self, **kwargs):
        pass

    def run(self):
        it = self._minibatches_count
        for gd, wp, bp in self._gd_units:
            if wp is not None:
                gd.learning_rate = float(wp(it))
            if bp is not None:
                gd.learning_rate_bias = float(bp(it))
        self._minibatches_count += 1
