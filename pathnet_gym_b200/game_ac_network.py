# -*- coding: utf-8 -*-
"""Drop-in for game_ac_network.py: the actor-critic PathNet super-net behind the reference's
class / method names, evaluated by libpathnet_b200.so.

The reference builds one TF graph per worker process and feeds it through a Session; here a
network object wraps a PathNetPopulation (P paths at once; P = 1 reproduces one reference worker).
`sess` arguments are accepted and ignored.  Variables are views into the fp32 parameter slab.
"""
import numpy as np
import torch

from . import pathnet
from .constants import ACTION_SIZEZ
from .population import PathNetConfig, PathNetPopulation


class SlabVar(object):
    """A trainable tensor = a view into the parameter slab (stands in for tf.Variable)."""

    def __init__(self, net, index, name):
        self.net, self.index, self.name = net, index, name

    @property
    def value(self):
        return self.net.pop.view(self.net.pop.params, self.index)

    def eval(self, sess=None):
        return self.value.detach().cpu().numpy().copy()

    def assign(self, array):
        self.value.copy_(torch.as_tensor(np.asarray(array, dtype=np.float32)).to(self.value.device).view_as(self.value))
        return self

    def get_shape(self):
        return tuple(self.value.shape)

    def __repr__(self):
        return "<SlabVar %s %s>" % (self.name, self.get_shape())


class GenotypeSlot(object):
    """Stands in for one worker's L*M gate variables + their assign ops (game_ac_network.py:158-170)."""

    def __init__(self, net, path):
        self.net, self.path = net, path

    def assign(self, geopath):
        self.net._set_path_geopath(self.path, geopath)


class GameACNetwork(object):
    """Base class (game_ac_network.py:10-110)."""

    def __init__(self, training_stage, thread_index, device="/cpu:0"):
        self.training_stage = training_stage
        self._action_size = ACTION_SIZEZ[self.training_stage]
        self._thread_index = thread_index
        self._device = device
        self.entropy_beta = 0.01

    def prepare_loss(self, entropy_beta):
        """game_ac_network.py:20-59: the loss itself is fused into the head kernel; only beta is state."""
        self.entropy_beta = float(entropy_beta)

    def run_policy_and_value(self, sess, s_t):
        raise NotImplementedError()

    def run_policy(self, sess, s_t):
        raise NotImplementedError()

    def run_value(self, sess, s_t):
        raise NotImplementedError()

    def get_vars(self):
        raise NotImplementedError()

    def sync_from(self, src_netowrk, name=None):
        """game_ac_network.py:73-85 (unused by the reference's driver): copy every variable."""
        for s, d in zip(src_netowrk.get_vars(), self.get_vars()):
            d.assign(s.eval())
        return None


class GameACPathNetNetwork(GameACNetwork):
    """Feed-forward PathNet actor-critic (game_ac_network.py:113-271).

    FLAGS fields read: task_index, L, M, kernel_num, stride_size, worker_hosts_num (as the reference),
    plus optional N, H, W (frame size; the reference fork hard-codes 160x120, game_ac_network.py:180),
    population (paths evaluated by this object, default 1), batch_max, gpu.
    """

    LSTM = False

    def __init__(self, training_stage, thread_index, device="/cpu:0", FLAGS=""):
        GameACNetwork.__init__(self, training_stage, thread_index, device)
        self.task_index = FLAGS.task_index
        kernel_num = tuple(int(x) for x in str(FLAGS.kernel_num).split(","))
        stride_size = tuple(int(x) for x in str(FLAGS.stride_size).split(","))
        self.cfg = PathNetConfig(L=FLAGS.L, M=FLAGS.M, N=getattr(FLAGS, "N", FLAGS.L), H=getattr(FLAGS, "H", 160),
                                 W=getattr(FLAGS, "W", 120), C=4, kernels=kernel_num, strides=stride_size, F=8,
                                 hidden=256, A=max(ACTION_SIZEZ), lstm=self.LSTM)
        self.P = int(getattr(FLAGS, "population", 1))
        self.pop = PathNetPopulation(self.cfg, self.P, int(getattr(FLAGS, "batch_max", 32)),
                                     device=int(getattr(FLAGS, "gpu", 0)))
        self.pop.init_params(int(getattr(FLAGS, "seed", 1)))
        L, M = self.cfg.L, self.cfg.M
        names = {}
        for i, t in enumerate(self.pop.tensors):
            names[(t["kind"], t["layer"], t["module"])] = i
        self.W_conv = np.zeros((L - 1, M), dtype=object)
        self.b_conv = np.zeros((L - 1, M), dtype=object)
        for l in range(L - 1):
            for j in range(M):
                self.W_conv[l, j] = SlabVar(self, names[(0, l, j)], "W_conv[%d,%d]" % (l, j))
                self.b_conv[l, j] = SlabVar(self, names[(1, l, j)], "b_conv[%d,%d]" % (l, j))
        self.W_lin = np.zeros(M, dtype=object)
        self.b_lin = np.zeros(M, dtype=object)
        for j in range(M):
            self.W_lin[j] = SlabVar(self, names[(2, L - 1, j)], "W_lin[%d]" % j)
            self.b_lin[j] = SlabVar(self, names[(3, L - 1, j)], "b_lin[%d]" % j)
        self.W_fc2, self.b_fc2 = SlabVar(self, names[(4, L, -1)], "W_fc2"), SlabVar(self, names[(5, L, -1)], "b_fc2")
        self.W_fc3, self.b_fc3 = SlabVar(self, names[(6, L, -1)], "W_fc3"), SlabVar(self, names[(7, L, -1)], "b_fc3")
        # genotype of every worker (all gates 1.0 at construction, pathnet.py:25-30)
        W = max(int(FLAGS.worker_hosts_num), self.P)
        self.geopath_set = np.zeros(W, dtype=object)
        for i in range(W):
            self.geopath_set[i] = pathnet.geopath_initializer(L, M)
        self.geopath_update_ops_set = np.zeros(W, dtype=object)
        self.geopath_update_placeholders_set = np.zeros(W, dtype=object)
        for i in range(W):
            self.geopath_update_ops_set[i] = GenotypeSlot(self, i)
        self.fixed_path = np.zeros((L, M), dtype=float)
        self._push_genotypes()
        self.s = self.a = self.td = self.r = None       # feeds are plain arguments here
        self.pi = self.v = self.total_loss = None

    # -- genotype plumbing ---------------------------------------------------
    def _path_ids(self):
        """Worker indices evaluated by this object: [task_index] for P = 1, else 0..P-1."""
        return [self.task_index] if self.P == 1 else list(range(self.P))

    def _push_genotypes(self):
        masks = np.stack([pathnet.masks_from_geopath(self.geopath_set[i]) for i in self._path_ids()])
        self.pop.set_genotypes(masks)

    def _set_path_geopath(self, path, geopath):
        self.geopath_set[path] = np.asarray(geopath, dtype=float).copy()
        if path in self._path_ids():
            self._push_genotypes()

    def set_training_stage(self, training_stage):
        self.training_stage = training_stage

    # -- acting forward (game_ac_network.py:209-227) ---------------------------
    def _forward_one(self, s_t):
        x = torch.as_tensor(np.ascontiguousarray(s_t, dtype=np.float32)).to(self.pop.device)
        x = x.view(1, 1, self.cfg.H, self.cfg.W, self.cfg.C).expand(self.P, 1, -1, -1, -1).contiguous()
        pi, v = self.pop.forward(x)
        k = 0 if self.P == 1 else self.task_index
        return pi[k, 0].cpu().numpy(), float(v[k, 0])

    def run_policy_and_value(self, sess, s_t):
        return self._forward_one(s_t)

    def run_policy(self, sess, s_t):
        return self._forward_one(s_t)[0]

    def run_value(self, sess, s_t):
        return self._forward_one(s_t)[1]

    def get_geopath(self, sess):
        return np.array(self.geopath_set[self.task_index], dtype=float)

    def set_fixed_path(self, fixed_path):
        self.fixed_path = np.asarray(fixed_path, dtype=float)

    def all_vars(self):
        res = []
        for l in range(self.W_conv.shape[0]):
            for j in range(self.W_conv.shape[1]):
                res += [self.W_conv[l, j], self.b_conv[l, j]]
        for j in range(len(self.W_lin)):
            res += [self.W_lin[j], self.b_lin[j]]
        return res + [self.W_fc2, self.b_fc2, self.W_fc3, self.b_fc3]

    def get_vars(self):
        """game_ac_network.py:239-253: variables not on the fixed path (+ heads)."""
        return [v for v, on in zip(self.all_vars(), self._idx_full()) if on]

    def _idx_full(self):
        res = []
        for l in range(self.W_conv.shape[0]):
            for j in range(self.W_conv.shape[1]):
                res += [0, 0] if self.fixed_path[l, j] != 0.0 else [1, 1]
        for j in range(len(self.W_lin)):
            res += [0, 0] if self.fixed_path[-1, j] != 0.0 else [1, 1]
        return res + [1, 1, 1, 1]

    def get_vars_idx(self):
        """game_ac_network.py:255-271: 1 = update op is run, 0 = frozen; index-aligned with the FULL list."""
        return self._idx_full()


class GameACPathNetLSTMNetwork(GameACPathNetNetwork):
    """PathNet-LSTM actor-critic (game_ac_network.py:303-521), the reference's default (constants.py:31).

    Same super-net without the /M (:394); the module sum feeds BasicLSTMCell(256) unrolled over the fed frames
    (batch 1 per path, :396-416) whose outputs feed the heads.  Extra variables W_lstm / b_lstm are appended to
    get_vars() and are never frozen (:437-438,502-520).  The cell state is carried on the host object exactly as the
    reference does it: run_policy_and_value / run_policy adopt the new state, run_value rolls it back (:468-481),
    reset_state zeroes it (:441-443)."""

    LSTM = True

    def __init__(self, training_stage, thread_index, device="/cpu:0", FLAGS=""):
        GameACPathNetNetwork.__init__(self, training_stage, thread_index, device, FLAGS)
        names = {(t["kind"]): i for i, t in enumerate(self.pop.tensors) if t["kind"] in (8, 9)}
        self.W_lstm, self.b_lstm = SlabVar(self, names[8], "W_lstm"), SlabVar(self, names[9], "b_lstm")
        self._state_next = torch.zeros(self.P, 2, self.cfg.hidden, dtype=torch.float32, device=self.pop.device)
        self.initial_lstm_state = self.step_size = None     # feeds are plain arguments here
        self.reset_state()

    def reset_state(self):
        """game_ac_network.py:441-443: zero (c, h)."""
        self.lstm_state_out = torch.zeros(self.P, 2, self.cfg.hidden, dtype=torch.float32, device=self.pop.device)

    def _forward_one(self, s_t, adopt_state=True):
        x = torch.as_tensor(np.ascontiguousarray(s_t, dtype=np.float32)).to(self.pop.device)
        x = x.view(1, 1, self.cfg.H, self.cfg.W, self.cfg.C).expand(self.P, 1, -1, -1, -1).contiguous()
        nxt = torch.empty_like(self.lstm_state_out)
        self.pop.bind_lstm_state(self.lstm_state_out, nxt)          # step_size = 1 from the carried state
        pi, v = self.pop.forward(x)
        if adopt_state:
            self.lstm_state_out = nxt
        k = 0 if self.P == 1 else self.task_index
        return pi[k, 0].cpu().numpy(), float(v[k, 0])

    def run_value(self, sess, s_t):
        """game_ac_network.py:468-481: the bootstrap value does not advance the state."""
        return self._forward_one(s_t, adopt_state=False)[1]

    def all_vars(self):
        return GameACPathNetNetwork.all_vars(self) + [self.W_lstm, self.b_lstm]

    def _idx_full(self):
        return GameACPathNetNetwork._idx_full(self) + [1, 1]
