"""CPU restatement of the PathNet-gym hot path.  TEST INFRASTRUCTURE ONLY.

This file is the parity *checker* for the CUDA library, and the `cpu_baseline`
arm of bench.py.  It must never sit on the product path: only `tests/`,
`__graft_entry__.smoke()` and `bench.py` (cpu_baseline / --impl reference) may
import it.  The product (pathnet_gym_b200) fails loudly without its CUDA
library and has no CPU fallback.

PARITY STATUS
  * genotype ops (get_geopath / mutation): PINNED - tests/golden/make_golden.py
    drives the *real* reference functions (/root/reference/pathnet.py:50-63,
    78-87, TensorFlow import stubbed) with this file's counter-based uniform
    stream and stores the results in tests/golden/genotype_golden.npz.
  * floating-point path (forward / loss / backward / clip / RMSProp):
    PARITY UNPINNED.  The arithmetic lives in TensorFlow 1.x (unpinned version,
    not vendored, not installable here) and the reference ships no golden
    vectors or numeric tests.  This file restates the TF-1 op semantics from the
    reference call sites; two independent restatements (torch autograd on the
    literal all-M gated graph, and a NumPy closed-form gathered evaluation) are
    cross-checked against each other in tests/test_oracle.py.

Every function cites the reference file:line it follows (paths relative to
/root/reference/).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

import numpy as np

# ----------------------------------------------------------------------------
# constants.py:4-32 (values are part of the numerical contract)
# ----------------------------------------------------------------------------
LOCAL_T_MAX = 20
RMSP_ALPHA = 0.99
RMSP_EPSILON = 0.1
INITIAL_ALPHA_LOW = 1e-4
INITIAL_ALPHA_HIGH = 1e-2
INITIAL_ALPHA_LOG_RATE = 0.4226
GAMMA = 0.99
ENTROPY_BETA = 0.01
MAX_TIME_STEP = 4 * 10 ** 6
GRAD_NORM_CLIP = 40.0
SCORE_UNSET = -1000.0  # doom_pathnet.py:131,267


@dataclass(frozen=True)
class OracleConfig:
    """Shape parameters of the super-net (game_ac_network.py:124-155,180)."""
    L: int = 3
    M: int = 10
    N: int = 3
    H: int = 84
    W: int = 84
    C: int = 4
    kernels: Tuple[int, ...] = (8, 4, 3)      # doom_pathnet.py:356-357
    strides: Tuple[int, ...] = (4, 2, 1)      # doom_pathnet.py:358-359
    F: int = 8                                 # feature_num, game_ac_network.py:132
    hidden: int = 256                          # game_ac_network.py:146
    A: int = 18                                # max(ACTION_SIZEZ), game_ac_network.py:152
    lstm: bool = False                         # GameACPathNetLSTMNetwork (game_ac_network.py:303-521): no /M, BasicLSTMCell(256)

    def conv_dims(self):
        """[(IH, IW, Cin, k, s, OH, OW)] for conv layers 0..L-2 (VALID padding)."""
        out = []
        ih, iw, cin = self.H, self.W, self.C
        for l in range(self.L - 1):
            k, s = self.kernels[l], self.strides[l]
            oh, ow = (ih - k) // s + 1, (iw - k) // s + 1
            out.append((ih, iw, cin, k, s, oh, ow))
            ih, iw, cin = oh, ow, self.F
        return out

    @property
    def last_lin(self) -> int:
        if self.L == 1:
            return self.H * self.W * self.C
        d = self.conv_dims()[-1]
        return d[5] * d[6] * self.F


def tensor_table(cfg: OracleConfig):
    """Variable order of get_vars() with fixed_path == 0 (game_ac_network.py:239-253).

    Returns a list of dicts: name, shape, layer, module (-1 for heads), kind
    (0 conv W, 1 conv b, 2 lin W, 3 lin b, 4 fc2 W, 5 fc2 b, 6 fc3 W, 7 fc3 b).
    """
    t = []
    dims = cfg.conv_dims()
    for l in range(cfg.L - 1):
        _, _, cin, k, _, _, _ = dims[l]
        for j in range(cfg.M):
            t.append(dict(name=f"W_conv[{l},{j}]", shape=(k, k, cin, cfg.F), layer=l, module=j, kind=0))
            t.append(dict(name=f"b_conv[{l},{j}]", shape=(cfg.F,), layer=l, module=j, kind=1))
    for j in range(cfg.M):
        t.append(dict(name=f"W_lin[{j}]", shape=(cfg.last_lin, cfg.hidden), layer=cfg.L - 1, module=j, kind=2))
        t.append(dict(name=f"b_lin[{j}]", shape=(cfg.hidden,), layer=cfg.L - 1, module=j, kind=3))
    t.append(dict(name="W_fc2", shape=(cfg.hidden, cfg.A), layer=cfg.L, module=-1, kind=4))
    t.append(dict(name="b_fc2", shape=(cfg.A,), layer=cfg.L, module=-1, kind=5))
    t.append(dict(name="W_fc3", shape=(cfg.hidden, 1), layer=cfg.L, module=-1, kind=6))
    t.append(dict(name="b_fc3", shape=(1,), layer=cfg.L, module=-1, kind=7))
    if cfg.lstm:
        # appended by the LSTM class's get_vars (game_ac_network.py:502-520); tf.contrib.rnn.BasicLSTMCell variables
        # ".../basic_lstm_cell/kernel" [input + units, 4 * units] and ".../bias" [4 * units] (:437-438)
        t.append(dict(name="W_lstm", shape=(2 * cfg.hidden, 4 * cfg.hidden), layer=cfg.L + 1, module=-1, kind=8))
        t.append(dict(name="b_lstm", shape=(4 * cfg.hidden,), layer=cfg.L + 1, module=-1, kind=9))
    return t


def init_params(cfg: OracleConfig, seed: int = 1) -> List[np.ndarray]:
    """U(-d, d), d = 1/sqrt(fan_in); the bias uses the weight's d
    (game_ac_network.py:89-107).  PCG64(seed) replaces tf.random_uniform."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = []
    d = 1.0
    for t in tensor_table(cfg):
        shp = t["shape"]
        if t["kind"] == 0:
            d = 1.0 / math.sqrt(shp[2] * shp[0] * shp[1])
        elif t["kind"] in (2, 4, 6):
            d = 1.0 / math.sqrt(shp[0])
        elif t["kind"] == 8:
            # BasicLSTMCell creates its kernel with the variable scope's default initializer, glorot_uniform:
            # U(-l, l), l = sqrt(6 / (fan_in + fan_out)); the bias is zeros_initializer (TF 1.x rnn_cell_impl.py)
            d = math.sqrt(6.0 / (shp[0] + shp[1]))
        if t["kind"] == 9:
            out.append(np.zeros(shp, dtype=np.float32))
            continue
        out.append(rng.uniform(-d, d, size=shp).astype(np.float32))
    return out


def frames_to_float(frames_u8: np.ndarray) -> np.ndarray:
    """game_state.py:49-50: x_t.astype(float32) * (1.0/255.0) in fp32."""
    return frames_u8.astype(np.float32) * np.float32(1.0 / 255.0)


def frozen_flags(cfg: OracleConfig, fixed_path: np.ndarray) -> np.ndarray:
    """1 = frozen.  Complement of get_vars_idx() (game_ac_network.py:255-271)."""
    fl = []
    for t in tensor_table(cfg):
        if t["module"] < 0:
            fl.append(0)
        else:
            row = t["layer"] if t["layer"] < cfg.L - 1 else -1
            fl.append(1 if fixed_path[row, t["module"]] == 1.0 else 0)
    return np.asarray(fl, dtype=np.uint8)


# ----------------------------------------------------------------------------
# Restatement #1: torch, literal graph (all M modules evaluated and gated)
# ----------------------------------------------------------------------------
def lstm_torch(x_seq, W, b, state):
    """tf.contrib.rnn.BasicLSTMCell(256, state_is_tuple=True) under tf.nn.dynamic_rnn, batch 1 (game_ac_network.py:396-416):
    [i | j | f | o] = [x_t, h] @ kernel + bias;  c' = c * sigmoid(f + 1.0) + sigmoid(i) * tanh(j);  h' = tanh(c') * sigmoid(o)
    (TF 1.x rnn_cell_impl.py BasicLSTMCell.call, forget_bias = 1.0).  x_seq [T, units]; state = (c, h), each [units]."""
    import torch
    c, h = state
    n = x_seq.shape[1]
    outs = []
    for t in range(x_seq.shape[0]):
        a = torch.cat([x_seq[t], h]) @ W + b
        i, j, f, o = a[:n], a[n:2 * n], a[2 * n:3 * n], a[3 * n:]
        c = c * torch.sigmoid(f + 1.0) + torch.sigmoid(i) * torch.tanh(j)
        h = torch.tanh(c) * torch.sigmoid(o)
        outs.append(h)
    return torch.stack(outs), (c, h)


def forward_torch(cfg: OracleConfig, params, frames, gates, lstm_state=None, return_state=False):
    """game_ac_network.py:182-201.  `frames` [n,H,W,C] float, `gates` [L,M] 0/1.

    Every module of every layer is evaluated and multiplied by its scalar gate,
    summed in ascending j (np.sum over the object array), then /M, then heads.
    cfg.lstm: the LSTM class (:381-426): no /M (:394), the n frames are ONE sequence in time order through the cell,
    whose outputs feed the heads; lstm_state = (c, h) at the start (zeros if None, reset_state :441-443).
    """
    import torch
    import torch.nn.functional as Fn
    tt = tensor_table(cfg)
    idx = {t["name"]: i for i, t in enumerate(tt)}
    net = frames.permute(0, 3, 1, 2)  # NHWC -> NCHW
    for l in range(cfg.L):
        if l == cfg.L - 1:
            # tf.reshape of NHWC => (h, w, c) flatten order (game_ac_network.py:185)
            net = net.permute(0, 2, 3, 1).reshape(net.shape[0], cfg.last_lin)
        acc = None
        for j in range(cfg.M):
            g = gates[l, j]
            if l < cfg.L - 1:
                Wt = params[idx[f"W_conv[{l},{j}]"]].permute(3, 2, 0, 1)  # HWIO -> OIHW
                b = params[idx[f"b_conv[{l},{j}]"]]
                y = torch.relu(Fn.conv2d(net, Wt, stride=cfg.strides[l]) + b.view(1, -1, 1, 1)) * g
            else:
                y = torch.relu(net @ params[idx[f"W_lin[{j}]"]] + params[idx[f"b_lin[{j}]"]]) * g
            acc = y if acc is None else acc + y
        net = acc
    state = None
    if cfg.lstm:
        if lstm_state is None:
            z0 = torch.zeros(cfg.hidden, dtype=net.dtype)
            lstm_state = (z0, z0)
        net, state = lstm_torch(net, params[idx["W_lstm"]], params[idx["b_lstm"]], lstm_state)
    else:
        net = net / cfg.M                                       # :194
    z = net @ params[idx["W_fc2"]] + params[idx["b_fc2"]]
    pi = torch.softmax(z, dim=1)                                # :198
    v = (net @ params[idx["W_fc3"]] + params[idx["b_fc3"]]).reshape(-1)  # :200-201
    if return_state:
        return pi, v, state
    return pi, v


def loss_torch(pi, v, a_onehot, td, r, entropy_beta=ENTROPY_BETA):
    """prepare_loss, game_ac_network.py:38-59 (sums, not means)."""
    import torch
    log_pi = torch.log(torch.clamp(pi, 1e-20, 1.0))
    entropy = -torch.sum(pi * log_pi, dim=1)
    policy_loss = -torch.sum(torch.sum(log_pi * a_onehot, dim=1) * td + entropy * entropy_beta)
    value_loss = 0.5 * (0.5 * torch.sum((r - v) ** 2))          # 0.5 * tf.nn.l2_loss
    return policy_loss + value_loss


def path_grads_torch(cfg, params_np, frames_f, gates, actions, td, r, beta=ENTROPY_BETA, dtype="float64", lstm_state=None):
    """tf.gradients(total_loss, get_vars()) for ONE path (a3c_training_thread.py:55-60).

    Returns (pi, v, total_loss, [grad per tensor]) as NumPy arrays of `dtype`."""
    import torch
    dt = getattr(torch, dtype)
    ps = [torch.tensor(p, dtype=dt, requires_grad=True) for p in params_np]
    x = torch.tensor(np.asarray(frames_f), dtype=dt)
    g = torch.tensor(np.asarray(gates), dtype=dt)
    st = None
    if lstm_state is not None:
        st = (torch.tensor(np.asarray(lstm_state[0]), dtype=dt), torch.tensor(np.asarray(lstm_state[1]), dtype=dt))
    pi, v = forward_torch(cfg, ps, x, g, lstm_state=st)
    a1h = torch.zeros(len(actions), cfg.A, dtype=dt)
    a1h[torch.arange(len(actions)), torch.tensor(np.asarray(actions), dtype=torch.long)] = 1.0
    loss = loss_torch(pi, v, a1h, torch.tensor(np.asarray(td), dtype=dt), torch.tensor(np.asarray(r), dtype=dt), beta)
    grads = torch.autograd.grad(loss, ps, allow_unused=True)
    gn = [(gr.detach().numpy() if gr is not None else np.zeros(p.shape, dtype=dtype)) for gr, p in zip(grads, params_np)]
    return pi.detach().numpy(), v.detach().numpy(), float(loss.detach()), gn


# ----------------------------------------------------------------------------
# Restatement #2: NumPy closed form, gathered active modules only
# ----------------------------------------------------------------------------
def _im2col(x, k, s):
    n, H, W, C = x.shape
    oh, ow = (H - k) // s + 1, (W - k) // s + 1
    sn, sh, sw, sc = x.strides
    cols = np.lib.stride_tricks.as_strided(
        x, shape=(n, oh, ow, k, k, C), strides=(sn, sh * s, sw * s, sh, sw, sc), writeable=False)
    return cols.reshape(n * oh * ow, k * k * C), oh, ow


def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def forward_backward_numpy(cfg, params, frames_f, gates, actions, td, r, beta=ENTROPY_BETA, dtype=np.float64,
                           want_grads=True, relu_override=None, pre_out=None, lstm_state=None, state_out=None):
    """Closed-form forward + backward for ONE path.

    Only gated-on modules are evaluated: adding exact +0.0 terms does not change
    the sum (game_ac_network.py:186-193).  Returns pi, v, loss, grads.

    Test aids: `pre_out` (a dict) receives every pre-activation, keyed (layer, module): conv layers
    [B*OH*OW, F], the linear layer [B, hidden].  `relu_override` (same keys, bool arrays of the same
    shapes) replaces the decision `pre > 0` of that module: the parity tests use it to evaluate the graph
    with the ReLU decisions another implementation took, which separates arithmetic differences from
    decisions on pre-activations that lie below fp32 resolution."""
    relu_override = relu_override or {}
    tt = tensor_table(cfg)
    idx = {t["name"]: i for i, t in enumerate(tt)}
    P = [np.asarray(p, dtype=dtype) for p in params]
    x = np.ascontiguousarray(np.asarray(frames_f, dtype=dtype))
    n = x.shape[0]
    saved = []
    dims = cfg.conv_dims()
    for l in range(cfg.L - 1):
        _, _, cin, k, s, oh, ow = dims[l]
        cols, _, _ = _im2col(x, k, s)
        acc = np.zeros((n * oh * ow, cfg.F), dtype=dtype)
        masks = {}
        for j in range(cfg.M):
            if gates[l, j] != 1.0:
                continue
            pre = cols @ P[idx[f"W_conv[{l},{j}]"]].reshape(-1, cfg.F) + P[idx[f"b_conv[{l},{j}]"]]
            if pre_out is not None:
                pre_out[(l, j)] = pre
            masks[j] = np.asarray(relu_override[(l, j)]).reshape(pre.shape) if (l, j) in relu_override else pre > 0
            acc = acc + np.where(masks[j], pre, 0)
        saved.append((x.shape, cols, masks))
        x = np.ascontiguousarray(acc.reshape(n, oh, ow, cfg.F))
    flat = x.reshape(n, cfg.last_lin)
    h = np.zeros((n, cfg.hidden), dtype=dtype)
    lmask = {}
    for j in range(cfg.M):
        if gates[cfg.L - 1, j] != 1.0:
            continue
        pre = flat @ P[idx[f"W_lin[{j}]"]] + P[idx[f"b_lin[{j}]"]]
        if pre_out is not None:
            pre_out[(cfg.L - 1, j)] = pre
        key = (cfg.L - 1, j)
        lmask[j] = np.asarray(relu_override[key]).reshape(pre.shape) if key in relu_override else pre > 0
        h = h + np.where(lmask[j], pre, 0)
    lstm_saved = None
    if cfg.lstm:
        # no /M (game_ac_network.py:394); the n transitions are one sequence through BasicLSTMCell (see lstm_torch)
        n_u = cfg.hidden
        Wl, bl = P[idx["W_lstm"]], P[idx["b_lstm"]]
        c = np.zeros(n_u, dtype=dtype) if lstm_state is None else np.asarray(lstm_state[0], dtype=dtype)
        hp = np.zeros(n_u, dtype=dtype) if lstm_state is None else np.asarray(lstm_state[1], dtype=dtype)
        x_seq, outs, lstm_saved = h, [], []
        for t in range(n):
            a = np.concatenate([x_seq[t], hp]) @ Wl + bl
            ig, gg, fg, og = _sigmoid(a[:n_u]), np.tanh(a[n_u:2 * n_u]), _sigmoid(a[2 * n_u:3 * n_u] + 1.0), _sigmoid(a[3 * n_u:])
            c_new = c * fg + ig * gg
            tc = np.tanh(c_new)
            lstm_saved.append((ig, gg, fg, og, c, tc, hp))
            c, hp = c_new, tc * og
            outs.append(hp)
        if state_out is not None:
            state_out[0], state_out[1] = c, hp
        h = np.stack(outs)
    else:
        h = h / cfg.M
    z = h @ P[idx["W_fc2"]] + P[idx["b_fc2"]]
    zs = z - z.max(axis=1, keepdims=True)
    e = np.exp(zs)
    pi = e / e.sum(axis=1, keepdims=True)
    v = (h @ P[idx["W_fc3"]] + P[idx["b_fc3"]]).reshape(-1)
    td = np.asarray(td, dtype=dtype)
    r = np.asarray(r, dtype=dtype)
    a1h = np.zeros((n, cfg.A), dtype=dtype)
    a1h[np.arange(n), np.asarray(actions)] = 1.0
    inside = (pi >= 1e-20) & (pi <= 1.0)
    logp = np.log(np.clip(pi, 1e-20, 1.0))
    ent = -(pi * logp).sum(axis=1)
    loss = -np.sum((logp * a1h).sum(axis=1) * td + ent * beta) + 0.25 * np.sum((r - v) ** 2)
    if not want_grads:
        return pi, v, loss, None
    # d loss / d pi (literal expression; clip gradient passes only inside [1e-20, 1])
    dpi = -(td[:, None] * a1h) * np.where(inside, 1.0 / np.clip(pi, 1e-20, 1.0), 0.0) \
        + beta * (logp + np.where(inside, 1.0, 0.0))
    dz = pi * (dpi - (dpi * pi).sum(axis=1, keepdims=True))       # softmax backward
    dv = -0.5 * (r - v)
    G = [np.zeros(p.shape, dtype=dtype) for p in P]
    G[idx["W_fc2"]] = h.T @ dz
    G[idx["b_fc2"]] = dz.sum(axis=0)
    G[idx["W_fc3"]] = h.T @ dv[:, None]
    G[idx["b_fc3"]] = dv.sum(keepdims=True)
    dh = dz @ P[idx["W_fc2"]].T + dv[:, None] * P[idx["W_fc3"]].T
    if cfg.lstm:
        # back-propagation through time; dh becomes the gradient at the cell INPUT (the module sum)
        n_u = cfg.hidden
        Wl = P[idx["W_lstm"]]
        dW, db = np.zeros_like(Wl), np.zeros(4 * n_u, dtype=dtype)
        dx = np.zeros_like(dh)
        dh_next, dc_next = np.zeros(n_u, dtype=dtype), np.zeros(n_u, dtype=dtype)
        for t in range(n - 1, -1, -1):
            ig, gg, fg, og, c_prev, tc, h_prev = lstm_saved[t]
            dht = dh[t] + dh_next
            dc = dht * og * (1.0 - tc * tc) + dc_next
            da = np.concatenate([dc * gg * ig * (1.0 - ig), dc * ig * (1.0 - gg * gg), dc * c_prev * fg * (1.0 - fg),
                                 dht * tc * og * (1.0 - og)])
            dW += np.outer(np.concatenate([x_seq[t], h_prev]), da)
            db += da
            dxh = Wl @ da
            dx[t], dh_next, dc_next = dxh[:n_u], dxh[n_u:], dc * fg
        G[idx["W_lstm"]], G[idx["b_lstm"]] = dW, db
        dh = dx
    else:
        dh = dh / cfg.M
    dflat = np.zeros_like(flat)
    for j, mk in lmask.items():
        dpre = dh * mk
        G[idx[f"W_lin[{j}]"]] = flat.T @ dpre
        G[idx[f"b_lin[{j}]"]] = dpre.sum(axis=0)
        dflat = dflat + dpre @ P[idx[f"W_lin[{j}]"]].T
    dy = dflat
    for l in range(cfg.L - 2, -1, -1):
        _, _, cin, k, s, oh, ow = dims[l]
        xshape, cols, masks = saved[l]
        dy2 = dy.reshape(n * oh * ow, cfg.F)
        dcols = np.zeros_like(cols) if l > 0 else None
        for j, mk in masks.items():
            dpre = dy2 * mk
            G[idx[f"W_conv[{l},{j}]"]] = (cols.T @ dpre).reshape(k, k, cin, cfg.F)
            G[idx[f"b_conv[{l},{j}]"]] = dpre.sum(axis=0)
            if l > 0:
                dcols = dcols + dpre @ P[idx[f"W_conv[{l},{j}]"]].reshape(-1, cfg.F).T
        if l > 0:
            dx = np.zeros(xshape, dtype=dtype)
            dc = dcols.reshape(n, oh, ow, k, k, cin)
            for kh in range(k):
                for kw in range(k):
                    dx[:, kh:kh + s * oh:s, kw:kw + s * ow:s, :] += dc[:, :, :, kh, kw, :]
            dy = dx
    return pi, v, loss, G


# ----------------------------------------------------------------------------
# clip_by_norm + ApplyRMSProp   (rmsprop_applier.py:34-39,79-106)
# ----------------------------------------------------------------------------
def clip_by_norm(g: np.ndarray, clip_norm: float) -> np.ndarray:
    """tf.clip_by_norm (TF 1.x clip_ops.py): t*clip_norm * min(rsqrt(sum(t*t)), 1/clip_norm),
    evaluated in g's dtype.  A zero tensor gives rsqrt = inf -> min picks 1/clip."""
    dt = g.dtype.type
    ss = dt(np.sum(g * g, dtype=g.dtype))
    with np.errstate(divide="ignore"):
        l2inv = dt(1.0) / np.sqrt(ss) if ss > 0 else dt(np.inf)
    scale = min(l2inv, dt(1.0) / dt(clip_norm))
    return (g * dt(clip_norm)) * dt(scale)


def apply_rms_prop(var, ms, mom, lr, rho, momentum, eps, grad):
    """TF core/kernels/training_ops.cc ApplyRMSProp (CPU functor), same dtype as var:
        ms  += (grad^2 - ms) * (1 - rho)
        mom  = mom * momentum + (grad * lr) / sqrt(eps + ms)
        var -= mom
    Returns new (var, ms, mom).  Slots start at ms=1, mom=0 (rmsprop_applier.py:34-39)."""
    dt = var.dtype.type
    ms2 = ms + (grad * grad - ms) * (dt(1.0) - dt(rho))
    mom2 = mom * dt(momentum) + (grad * dt(lr)) / np.sqrt(dt(eps) + ms2)
    return var - mom2, ms2, mom2


def apply_gradients(params, ms, mom, grads, frozen, lr, rho=RMSP_ALPHA, momentum=0.0, eps=RMSP_EPSILON,
                    clip=GRAD_NORM_CLIP):
    """RMSPropApplier.apply_gradients restricted to the ops the caller runs:
    tensors with get_vars_idx()==0 are skipped entirely (a3c_training_thread.py:205-209)."""
    P2, S2, Q2 = [], [], []
    for p, s, q, g, fz in zip(params, ms, mom, grads, frozen):
        if fz:
            P2.append(p.copy()); S2.append(s.copy()); Q2.append(q.copy())
            continue
        gc = clip_by_norm(np.asarray(g, dtype=p.dtype), clip)
        a, b, c = apply_rms_prop(p, s, q, lr, rho, momentum, eps, gc)
        P2.append(a); S2.append(b); Q2.append(c)
    return P2, S2, Q2


def log_uniform(lo, hi, rate):
    """doom_pathnet.py:46-50."""
    return math.exp(math.log(lo) * (1 - rate) + math.log(hi) * rate)


def anneal_learning_rate(initial_lr, global_t, max_t=MAX_TIME_STEP):
    """a3c_training_thread.py:83-87."""
    lr = initial_lr * (max_t - global_t) / max_t
    return lr if lr > 0.0 else 0.0


def nstep_returns(rewards, values, bootstrap, gamma=GAMMA):
    """a3c_training_thread.py:155-180.  Inputs in rollout order; outputs (R, td)
    in rollout order too (the reference builds them reversed; the loss is a sum)."""
    R = float(bootstrap)
    n = len(rewards)
    Rs = np.zeros(n, dtype=np.float64)
    tds = np.zeros(n, dtype=np.float64)
    for i in range(n - 1, -1, -1):
        R = float(np.clip(rewards[i], -1, 1)) + gamma * R
        Rs[i] = R
        tds[i] = R - float(values[i])
    return Rs, tds


def nstep_returns_batch(rewards, values, bootstrap, lengths=None, gamma=GAMMA):
    """nstep_returns for [P, B] arrays with padding (rows past lengths[p] -> 0), rounded to fp32 like the feed."""
    rewards, values = np.asarray(rewards), np.asarray(values)
    P, B = rewards.shape
    R = np.zeros((P, B), np.float32)
    td = np.zeros((P, B), np.float32)
    for p in range(P):
        n = B if lengths is None else int(lengths[p])
        Rs, tds = nstep_returns(rewards[p, :n].astype(np.float64), values[p, :n].astype(np.float64), float(bootstrap[p]), gamma)
        R[p, :n], td[p, :n] = Rs.astype(np.float32), tds.astype(np.float32)
    return R, td


# ----------------------------------------------------------------------------
# Frame preprocessing (game_state.py:37-51,62-66,75-81): OpenCV's 8-bit paths restated
# ----------------------------------------------------------------------------
def bgr2gray_u8(img) -> np.ndarray:
    """cv2.cvtColor(img, cv2.COLOR_BGR2GRAY) for u8: 15-bit fixed point, round half up (pinned against cv2)."""
    img = np.asarray(img)
    b, g, r = (img[..., k].astype(np.int32) for k in range(3))
    return ((3735 * b + 19235 * g + 9798 * r + (1 << 14)) >> 15).astype(np.uint8)


def _resize_taps(dn, sn):
    idx = np.zeros((dn, 2), np.int32)
    a = np.zeros((dn, 2), np.int32)
    scale = float(sn) / float(dn)
    for d in range(dn):
        f = np.float32((d + 0.5) * scale - 0.5)
        s = int(np.floor(f))
        f = np.float32(f - np.float32(s))
        if s < 0:
            s, f = 0, np.float32(0)
        if s >= sn - 1:
            s, f = sn - 1, np.float32(0)
        idx[d] = (s, min(s + 1, sn - 1))
        a[d, 0] = int(np.rint(np.float32((np.float32(1) - f) * np.float32(2048))))
        a[d, 1] = int(np.rint(np.float32(f * np.float32(2048))))
    return idx, a


def resize_linear_u8(src, dw, dh) -> np.ndarray:
    """cv2.resize(src, (dw, dh)) with the default INTER_LINEAR on a u8 image: 11-bit coefficients, the fixed-point
    horizontal pass then `(((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2` vertically (pinned against cv2)."""
    S = np.asarray(src).astype(np.int32)
    sh, sw = S.shape
    xi, xa = _resize_taps(dw, sw)
    yi, ya = _resize_taps(dh, sh)
    rows = S[:, xi[:, 0]] * xa[:, 0] + S[:, xi[:, 1]] * xa[:, 1]
    r0, r1 = rows[yi[:, 0]], rows[yi[:, 1]]
    out = (((ya[:, 0:1] * (r0 >> 4)) >> 16) + ((ya[:, 1:2] * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def preprocess_frames(obs, frames, reset=None) -> np.ndarray:
    """GameState._process_frame + stack update for n environments: obs [n,Hin,Win,3] u8, frames [n,H,W,C] u8 (copy
    returned): new frame appended as the last channel, or all channels on reset."""
    out = np.array(frames, dtype=np.uint8).copy()
    n, H, W, C = out.shape
    for e in range(n):
        x = resize_linear_u8(bgr2gray_u8(obs[e]), W, H)
        if reset is not None and reset[e]:
            out[e] = x[..., None]
        else:
            out[e, :, :, :-1] = out[e, :, :, 1:]
            out[e, :, :, -1] = x
    return out


DOMAIN_ACTION = 3


def choice_from_uniform(p, u) -> int:
    """np.random.choice(range(A), p=p) as numpy's legacy RandomState implements it, with the uniform draw u made
    explicit: cdf = cumsum(float64(p)); cdf /= cdf[-1]; index = searchsorted(cdf, u, side='right')
    (a3c_training_thread.py:89-90).  Pinned against RandomState.choice itself in tests/test_oracle.py."""
    cdf = np.cumsum(np.asarray(p, dtype=np.float64))
    cdf = cdf / cdf[-1]
    return int(min(np.searchsorted(cdf, u, side="right"), len(cdf) - 1))


def sample_actions(pi, seed, epoch, id0=0) -> np.ndarray:
    """choose_action for every row of pi [n, A] with u = Philox(seed; epoch, id0 + row, 0, DOMAIN_ACTION)."""
    pi = np.asarray(pi, dtype=np.float32)
    return np.asarray([choice_from_uniform(pi[t], philox_uniform(seed, epoch, id0 + t, 0, DOMAIN_ACTION, 0))
                       for t in range(pi.shape[0])], dtype=np.int32)


# ----------------------------------------------------------------------------
# Population step = the unit the CUDA library implements
# ----------------------------------------------------------------------------
def gates_from_masks(cfg, masks_row) -> np.ndarray:
    g = np.zeros((cfg.L, cfg.M), dtype=np.float64)
    for l in range(cfg.L):
        for j in range(cfg.M):
            g[l, j] = float((int(masks_row[l]) >> j) & 1)
    return g


def masks_from_gates(g) -> np.ndarray:
    g = np.asarray(g)
    return np.asarray([sum((1 << j) for j in range(g.shape[1]) if g[l, j] == 1.0) for l in range(g.shape[0])],
                      dtype=np.uint32)


def population_step(cfg, params, frames_u8, masks, actions, td, r, beta=ENTROPY_BETA, dtype="float64",
                    impl="torch", relu_override=None, pre_out=None, lstm_states=None):
    """Synchronous population gradient: per-path tf.gradients, summed over paths
    (the shared-module gradient the B200 design all-reduces).  frames_u8 [P,B,H,W,C].
    relu_override / pre_out (impl="numpy" only): lists with one dict per path, see forward_backward_numpy."""
    P = frames_u8.shape[0]
    pis, vs, losses = [], [], []
    G = [np.zeros(p.shape, dtype=dtype) for p in params]
    for p in range(P):
        xf = frames_to_float(frames_u8[p]) if frames_u8.dtype == np.uint8 else frames_u8[p]
        gates = gates_from_masks(cfg, masks[p])
        if impl == "torch":
            pi, v, ls, g = path_grads_torch(cfg, params, xf, gates, actions[p], td[p], r[p], beta, dtype,
                                            lstm_state=lstm_states[p] if lstm_states is not None else None)
        else:
            pi, v, ls, g = forward_backward_numpy(cfg, params, xf, gates, actions[p], td[p], r[p], beta,
                                                  np.dtype(dtype).type,
                                                  relu_override=relu_override[p] if relu_override is not None else None,
                                                  pre_out=pre_out[p] if pre_out is not None else None,
                                                  lstm_state=lstm_states[p] if lstm_states is not None else None)
        pis.append(pi); vs.append(v); losses.append(ls)
        for a, b in zip(G, g):
            a += b
    return np.stack(pis), np.stack(vs), np.asarray(losses), G


# ----------------------------------------------------------------------------
# Counter-based RNG + genotype ops
# ----------------------------------------------------------------------------
DOMAIN_MUTATE = 0
DOMAIN_INIT = 1
DOMAIN_PERM = 2
DOMAIN_ACTION = 3

_M0, _M1 = 0xD2511F53, 0xCD9E8D57
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK32 = 0xFFFFFFFF


def philox4x32_10(ctr: Sequence[int], key: Sequence[int]) -> Tuple[int, int, int, int]:
    """Philox4x32-10 (Salmon et al., SC'11; Random123).  Pure-Python integers."""
    c0, c1, c2, c3 = [int(c) & _MASK32 for c in ctr]
    k0, k1 = int(key[0]) & _MASK32, int(key[1]) & _MASK32
    for rnd in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> 32, p0 & _MASK32
        hi1, lo1 = p1 >> 32, p1 & _MASK32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK32, lo1, (hi0 ^ c3 ^ k1) & _MASK32, lo0
        k0 = (k0 + _W0) & _MASK32
        k1 = (k1 + _W1) & _MASK32
    return c0, c1, c2, c3


def philox_u64(seed: int, epoch: int, path_id: int, pos: int, domain: int, slot: int) -> int:
    """Counter layout: (epoch, path_id, pos, domain<<24 | slot); key = (seed lo, seed hi)."""
    o = philox4x32_10((epoch, path_id, pos, ((domain & 0xFF) << 24) | (slot & 0xFFFFFF)),
                      (seed & _MASK32, (seed >> 32) & _MASK32))
    return (o[0] << 32) | o[1]


def philox_uniform(seed, epoch, path_id, pos, domain, slot) -> float:
    """u = (x >> 11) * 2^-53 in fp64, x = out0*2^32 + out1."""
    return float(philox_u64(seed, epoch, path_id, pos, domain, slot) >> 11) * (2.0 ** -53)


def get_geopath_masks(L, M, N, seed, epoch, path_id) -> np.ndarray:
    """pathnet.get_geopath (pathnet.py:78-87) on bitmasks.  Layer l draws
    slot = 0,1,2,... (pos = l) of `int(rand()*M)` until N distinct bits are set."""
    out = np.zeros(L, dtype=np.uint32)
    for l in range(L):
        m, cnt, slot = 0, 0, 0
        while cnt < N:
            rv = int(philox_uniform(seed, epoch, path_id, l, DOMAIN_INIT, slot) * M)
            slot += 1
            if not (m >> rv) & 1:
                m |= 1 << rv
                cnt += 1
        out[l] = m
    return out


def mutate_masks(masks: np.ndarray, L, M, N, seed, epoch, path_id) -> np.ndarray:
    """pathnet.mutation (pathnet.py:50-63) on bitmasks; row-major scan of the LIVE row.
    At (l,j): slot 0 is the `np.random.rand()` draw, slot 1 the `np.random.randint(0,M)`
    draw (restated as floor(u*M))."""
    out = np.array(masks, dtype=np.uint32).copy()
    for l in range(L):
        row = int(out[l])
        for j in range(M):
            u0 = philox_uniform(seed, epoch, path_id, l * M + j, DOMAIN_MUTATE, 0)
            if (row >> j) & 1:
                if int(u0 * L * N) <= 1:
                    row &= ~(1 << j)
                    c = int(philox_uniform(seed, epoch, path_id, l * M + j, DOMAIN_MUTATE, 1) * M)
                    row |= 1 << c
            else:
                if int(u0 * L * (M - N) * M) <= 1:
                    c = int(philox_uniform(seed, epoch, path_id, l * M + j, DOMAIN_MUTATE, 1) * M)
                    row |= 1 << c
        out[l] = row
    return out


def tournament_permutation(P, seed, epoch) -> np.ndarray:
    """Stand-in for np.random.shuffle(arange(W)) (doom_pathnet.py:227-229,268-270):
    paths ordered by their 64-bit Philox key (ties by path id)."""
    keys = [(philox_u64(seed, epoch, i, 0, DOMAIN_PERM, 0), i) for i in range(P)]
    keys.sort()
    return np.asarray([i for _, i in keys], dtype=np.int32)


def tournament(genotypes, scores, fixed, L, M, N, Bcand, n_groups, seed, epoch):
    """Controller tournament (doom_pathnet.py:241-270), batched over disjoint groups.

    genotypes [P,L] u32 (raw, without the frozen path), scores [P] f32, fixed [L] u32.
    Group g = perm[g*B:(g+1)*B].  A group with any score == -1000 is left untouched
    (the reference sleeps and retries).  Otherwise winner = first arg-max in group order,
    every other member <- mutation(copy(winner)), ALL members' scores <- -1000.
    n_groups = 1 is exactly one reference tournament.
    Returns (genotypes', effective masks = genotypes' | fixed, scores', winners[n_groups])."""
    g = np.array(genotypes, dtype=np.uint32).copy()
    sc = np.array(scores, dtype=np.float32).copy()
    P = g.shape[0]
    perm = tournament_permutation(P, seed, epoch)
    winners = np.full(n_groups, -1, dtype=np.int32)
    for grp in range(n_groups):
        mem = perm[grp * Bcand:(grp + 1) * Bcand]
        if len(mem) < Bcand or np.any(sc[mem] == np.float32(SCORE_UNSET)):
            continue
        w = int(mem[int(np.argmax(sc[mem]))])
        winners[grp] = w
        for m in mem:
            if m != w:
                g[m] = mutate_masks(g[w], L, M, N, seed, epoch, int(m))
            sc[m] = np.float32(SCORE_UNSET)
    eff = g | np.asarray(fixed, dtype=np.uint32)[None, :]
    return g, eff, sc, winners


# ----------------------------------------------------------------------------
# Float-array wrappers with the reference signatures (used by the golden script)
# ----------------------------------------------------------------------------
def get_geopath(L, M, N, seed, epoch, path_id=0) -> np.ndarray:
    return gates_from_masks(OracleConfig(L=L, M=M, N=N), get_geopath_masks(L, M, N, seed, epoch, path_id))


def mutation(geopath, L, M, N, seed, epoch, path_id=0) -> np.ndarray:
    m = mutate_masks(masks_from_gates(geopath), L, M, N, seed, epoch, path_id)
    return gates_from_masks(OracleConfig(L=L, M=M, N=N), m)
