# -*- coding: utf-8 -*-
"""Drop-in for rmsprop_applier.py: per-tensor clip_by_norm + RMSProp with the "rms" slot at 1.0 and
the "momentum" slot at 0 (rmsprop_applier.py:34-39,79-106), executed by pn_rmsprop_apply."""
import numpy as np


class ApplyOp(object):
    """One per-variable update op; the caller runs a subset (a3c_training_thread.py:205-209)."""

    def __init__(self, applier, var):
        self.applier, self.var = applier, var


class RMSPropApplier(object):
    def __init__(self, learning_rate, decay=0.9, momentum=0.0, epsilon=1e-10, clip_norm=40.0, device="/cpu:0",
                 name="RMSPropApplier"):
        self._name = name
        self._learning_rate = learning_rate
        self._decay = decay
        self._momentum = momentum
        self._epsilon = epsilon
        self._clip_norm = clip_norm
        self._device = device
        self._net = None

    def get_slot(self, var, name):
        """"rms" or "momentum" slot of `var` as a NumPy array (rmsprop_applier.py:63-67)."""
        pop = var.net.pop
        if name == "rms":
            return pop.view(pop.ms, var.index).detach().cpu().numpy().copy()
        if name == "momentum":
            if pop.mom is None:
                return np.zeros(var.get_shape(), dtype=np.float32)
            return pop.view(pop.mom, var.index).detach().cpu().numpy().copy()
        return None

    def apply_gradients(self, var_list, accum_grad_list, name=None):
        """Returns a LIST of per-variable ops, index-aligned with var_list (rmsprop_applier.py:92-106).
        The gradients live in the network's gradient slab (accum_grad_list is accepted for signature
        compatibility)."""
        if var_list:
            self._net = var_list[0].net
            if self._momentum != 0.0 and self._net.pop.mom is None:
                raise ValueError("momentum != 0 needs PathNetPopulation(momentum_slot=True)")
        return [ApplyOp(self, v) for v in var_list]

    def run(self, ops, learning_rate):
        """Run a subset of the ops: every tensor whose op is absent is frozen for this update."""
        if not ops:
            return
        net = ops[0].var.net
        run = np.zeros(len(net.pop.tensors), dtype=np.uint8)
        for op in ops:
            run[op.var.index] = 1
        net.pop.set_frozen(1 - run)
        net.pop.apply_gradients(float(learning_rate), self._decay, self._momentum, self._epsilon, self._clip_norm)
