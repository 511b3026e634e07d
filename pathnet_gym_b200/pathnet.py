"""Drop-in for the reference's pathnet.py: same function names and argument meaning.

Live functions (pathnet.py:20-30,50-63,78-87) run on the GPU through libpathnet_b200.so with a
counter-based RNG (Philox4x32-10) instead of NumPy's global unseeded MT19937: each call consumes one
`epoch` of the module-level stream, so runs are reproducible after `seed()`.  The MNIST-era builders
the reference never calls (pathnet.py:90-196) are kept as thin torch helpers for API completeness.
"""
import numpy as np
import torch

from . import population as _pop

_state = {"seed": 0x5EED5EED, "epoch": 0, "path_id": 0, "device": 0}


def seed(seed, epoch=0, device=0):
    """Fix the Philox key / starting epoch of get_geopath() and mutation()."""
    _state.update(seed=int(seed), epoch=int(epoch), device=int(device))


def _next_epoch():
    e = _state["epoch"]
    _state["epoch"] = e + 1
    return e


def masks_from_geopath(geopath):
    g = np.asarray(geopath)
    return np.asarray([sum(1 << j for j in range(g.shape[1]) if g[l, j] == 1.0) for l in range(g.shape[0])], dtype=np.uint32)


def geopath_from_masks(masks, M):
    masks = np.asarray(masks, dtype=np.uint32)
    return ((masks[:, None] >> np.arange(M, dtype=np.uint32)[None, :]) & 1).astype(float)


def get_geopath(L, M, N):
    """pathnet.py:78-87 -> ndarray [L, M] of 0.0/1.0 with N distinct modules per layer."""
    m = _pop.geopath_init(1, L, M, N, _state["seed"], _next_epoch(), path_id0=_state["path_id"], device=_state["device"])
    return geopath_from_masks(m.cpu().numpy().view(np.uint32)[0], M)


def mutation(geopath, L, M, N):
    """pathnet.py:50-63 -- mutates `geopath` in place and returns it."""
    dev = torch.device("cuda", _state["device"])
    m = torch.from_numpy(masks_from_geopath(geopath).view(np.int32).reshape(1, L).copy()).to(dev)
    _pop.mutate_(m, M, N, _state["seed"], _next_epoch(), path_id0=_state["path_id"])
    geopath[...] = geopath_from_masks(m.cpu().numpy().view(np.uint32)[0], M)
    return geopath


def geopath_initializer(L, M):
    """pathnet.py:25-30: every gate starts at 1.0."""
    return np.ones((L, M), dtype=float)


def geopath_insert(sess, geopath_update_placeholders, geopath_update_ops, candi, L, M):
    """pathnet.py:20-23: write genotype `candi` [L, M] into a worker's gate variables.  Here
    `geopath_update_ops` is the worker's GenotypeSlot and the L*M scalar assigns are one call."""
    geopath_update_ops.assign(np.asarray(candi, dtype=float).reshape(L, M))


def parameters_backup(var_list_to_learn):
    """pathnet.py:10-14."""
    out = np.zeros(len(var_list_to_learn), dtype=object)
    for i, v in enumerate(var_list_to_learn):
        out[i] = v.eval() if hasattr(v, "eval") else np.array(v)
    return out


def parameters_update(sess, var_update_placeholders, var_update_ops, var_list_backup):
    """pathnet.py:16-18."""
    for i in range(len(var_update_placeholders)):
        var_update_ops[i].assign(var_list_backup[i])


# -- dead in the reference (never called by doom_pathnet.py); kept as small torch helpers -----------
def weight_variable(shape):
    return torch.nn.init.trunc_normal_(torch.empty(*shape), std=0.1, a=-0.2, b=0.2)


def bias_variable(shape):
    return torch.full(tuple(shape), 0.1)


def module_weight_variable(shape):
    return [weight_variable(shape)]


def module_bias_variable(shape):
    return [bias_variable(shape)]


def variable_summaries(var):
    return dict(mean=float(var.mean()), stddev=float(var.std(unbiased=False)), max=float(var.max()), min=float(var.min()))


def module(input_tensor, weights, biases, layer_name, act=torch.relu):
    return act(input_tensor @ weights[0] + biases[0])


def module2(i, input_tensor, weights, biases, layer_name, act=torch.relu):
    if i % 3 == 0:
        return input_tensor
    y = act(input_tensor @ weights[0] + biases[0])
    return y if i % 3 == 1 else y + input_tensor


def conv_module(input_tensor, weights, biases, stride, layer_name, act=torch.relu):
    x = input_tensor.permute(0, 3, 1, 2)
    y = torch.nn.functional.conv2d(x, weights[0].permute(3, 2, 0, 1), stride=stride) + biases[0].view(1, -1, 1, 1)
    return act(y).permute(0, 2, 3, 1)


def nn_layer(input_tensor, weights, biases, layer_name, act=torch.relu):
    return input_tensor @ weights[0] + biases[0]
