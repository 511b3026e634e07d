"""Checkpoint / interchange.

The reference persists its state through `tf.train.Saver` inside the Supervisor's logdir (doom_pathnet.py:157-166):
every trainable tensor of the chief's `net_0` scope, the RMSProp slots, the genotype scalars and the `fixed_path`
scalars.  Here a checkpoint is one `.npz` with the tensors in `get_vars()` order (game_ac_network.py:239-253) under
readable names, the `rms` (and optional `momentum`) slots, the genotype bitmasks and the frozen-path mask, plus the
TF-Saver variable name of every tensor so that reference-trained weights can be mapped onto the slab if a TF
checkpoint ever appears: variables are created in `get_vars()` order inside `tf.variable_scope("net_<i>")`
(game_ac_network.py:125-148), so tensor k is `net_<i>/Variable` (k = 0) or `net_<i>/Variable_<k>`.

Everything here works on NumPy arrays (CPU); `save_population` / `load_population` move a device population through it.
"""
import json

import numpy as np

FORMAT = "pathnet_b200.checkpoint.v1"


def tensor_names(cfg):
    """Readable tensor names in get_vars() order with nothing frozen."""
    names = []
    for l in range(cfg.L - 1):
        for j in range(cfg.M):
            names += ["W_conv[%d,%d]" % (l, j), "b_conv[%d,%d]" % (l, j)]
    for j in range(cfg.M):
        names += ["W_lin[%d]" % j, "b_lin[%d]" % j]
    names += ["W_fc2", "b_fc2", "W_fc3", "b_fc3"]
    if getattr(cfg, "lstm", False):                 # GameACPathNetLSTMNetwork.get_vars appends them (game_ac_network.py:502-506)
        names += ["W_lstm", "b_lstm"]
    return names


def tf_saver_names(cfg, thread_index=0):
    """Name each tensor has in a reference tf.train.Saver checkpoint: the conv / linear / head tensors are anonymous
    tf.Variables created in get_vars() order (`Variable`, `Variable_1`, ...), the LSTM cell's two are created by
    dynamic_rnn under their own names (game_ac_network.py:411-416,436-438)."""
    lstm = bool(getattr(cfg, "lstm", False))
    n = len(tensor_names(cfg)) - (2 if lstm else 0)
    names = ["net_%d/Variable" % thread_index if k == 0 else "net_%d/Variable_%d" % (thread_index, k) for k in range(n)]
    if lstm:
        names += ["net_%d/basic_lstm_cell/kernel" % thread_index, "net_%d/basic_lstm_cell/bias" % thread_index]
    return names


def save_checkpoint(path, cfg, params, rms, momentum=None, genotypes=None, fixed_path=None, global_t=0, extra=None):
    """params / rms / momentum: lists of arrays in get_vars() order; genotypes [P, L] u32 bitmasks; fixed_path [L, M]."""
    names = tensor_names(cfg)
    if len(params) != len(names) or len(rms) != len(names):
        raise ValueError("expected %d tensors, got %d params / %d rms" % (len(names), len(params), len(rms)))
    out = {"__format__": np.asarray(FORMAT), "__names__": np.asarray(names), "__tf_names__": np.asarray(tf_saver_names(cfg)),
           "__config__": np.asarray(json.dumps({k: getattr(cfg, k) for k in ("L", "M", "H", "W", "C", "F", "hidden", "A", "lstm")
                                                if hasattr(cfg, k)})),
           "__global_t__": np.asarray(int(global_t), dtype=np.int64)}
    for k, nm in enumerate(names):
        out["param/" + nm] = np.asarray(params[k], dtype=np.float32)
        out["rms/" + nm] = np.asarray(rms[k], dtype=np.float32)
        if momentum is not None:
            out["momentum/" + nm] = np.asarray(momentum[k], dtype=np.float32)
    if genotypes is not None:
        out["genotypes"] = np.asarray(genotypes, dtype=np.uint32)
    if fixed_path is not None:
        out["fixed_path"] = np.asarray(fixed_path, dtype=np.float32)
    for k, v in (extra or {}).items():
        out["extra/" + k] = np.asarray(v)
    np.savez(path, **out)


def load_checkpoint(path, cfg):
    """-> dict(params, rms, momentum | None, genotypes | None, fixed_path | None, global_t, extra); shapes are checked
    against cfg's tensor table by the caller (load_population does)."""
    z = np.load(path, allow_pickle=False)
    if str(z["__format__"]) != FORMAT:
        raise ValueError("not a %s file: %s" % (FORMAT, path))
    names = tensor_names(cfg)
    if [str(x) for x in z["__names__"]] != names:
        raise ValueError("checkpoint tensor table does not match this configuration (L, M or the LSTM flag differ)")
    has_mom = ("momentum/" + names[0]) in z.files
    return dict(params=[z["param/" + nm] for nm in names], rms=[z["rms/" + nm] for nm in names],
                momentum=[z["momentum/" + nm] for nm in names] if has_mom else None,
                genotypes=z["genotypes"] if "genotypes" in z.files else None,
                fixed_path=z["fixed_path"] if "fixed_path" in z.files else None, global_t=int(z["__global_t__"]),
                extra={k[6:]: z[k] for k in z.files if k.startswith("extra/")})


def from_tf_variables(cfg, variables, thread_index=0):
    """Map a {tf variable name: array} dict (e.g. read from a reference Saver checkpoint) onto the get_vars() list."""
    return [np.asarray(variables[nm], dtype=np.float32) for nm in tf_saver_names(cfg, thread_index)]


def save_population(path, pop, genotypes=None, fixed_path=None, global_t=0):
    mom = pop.dump(pop.mom) if pop.mom is not None else None
    if genotypes is None:
        genotypes = pop.masks
    save_checkpoint(path, pop.cfg, pop.dump(pop.params), pop.dump(pop.ms), mom, genotypes, fixed_path, global_t)


def load_population(path, pop):
    ck = load_checkpoint(path, pop.cfg)
    for k, (a, shp) in enumerate(zip(ck["params"], pop.shapes)):
        if tuple(a.shape) != tuple(shp):
            raise ValueError("tensor %d has shape %s in the checkpoint, %s here" % (k, a.shape, shp))
    pop.load(pop.params, ck["params"])
    pop.load(pop.ms, ck["rms"])
    if ck["momentum"] is not None and pop.mom is not None:
        pop.load(pop.mom, ck["momentum"])
    if ck["genotypes"] is not None and ck["genotypes"].shape[0] == pop.P:
        pop.set_genotypes(ck["genotypes"])
    if ck["fixed_path"] is not None:
        pop.set_fixed_path(ck["fixed_path"])
    return ck
