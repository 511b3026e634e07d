"""Shared helpers for the parity tests (test infrastructure; may import oracle/)."""
import os

import numpy as np

from oracle import pathnet_oracle as po

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TINY = po.OracleConfig(L=3, M=4, N=2, H=20, W=20, C=4, kernels=(4, 3, 3), strides=(2, 1, 1), F=8, hidden=32, A=5)
TINY4 = po.OracleConfig(L=4, M=5, N=2, H=28, W=24, C=4, kernels=(4, 4, 3), strides=(2, 2, 1), F=8, hidden=64, A=3)
# smallest geometry the tensor-core layer-0 path accepts (k = 2*stride, stride*C = 16)
TC20 = po.OracleConfig(L=3, M=10, N=3, H=20, W=24, C=4, kernels=(8, 2, 3), strides=(4, 1, 1), F=8, hidden=32, A=6)
C1 = po.OracleConfig(L=3, M=10, N=3, H=84, W=84, C=4, A=18)
C3 = po.OracleConfig(L=4, M=10, N=4, H=84, W=84, C=4, A=18)


def synth_inputs(cfg, P, B, seed):
    """Must stay identical to tests/golden/make_golden.py:synth_inputs."""
    rng = np.random.Generator(np.random.PCG64(seed))
    frames = rng.integers(0, 256, size=(P, B, cfg.H, cfg.W, cfg.C), dtype=np.uint8)
    actions = rng.integers(0, cfg.A, size=(P, B)).astype(np.int32)
    td = rng.normal(size=(P, B)).astype(np.float32)
    r = (rng.uniform(-1, 1, size=(P, B)) * 5.0).astype(np.float32)
    masks = np.stack([po.get_geopath_masks(cfg.L, cfg.M, cfg.N, 7, 0, p) for p in range(P)])
    return frames, actions, td, r, masks


def cfg_from_golden(arr):
    L, M, N, H, W, C, F, hidden, A = [int(x) for x in arr[:9]]
    ks = tuple(int(x) for x in arr[9:12])
    st = tuple(int(x) for x in arr[12:15])
    return po.OracleConfig(L=L, M=M, N=N, H=H, W=W, C=C, kernels=ks, strides=st, F=F, hidden=hidden, A=A)


def to_engine_cfg(cfg):
    from pathnet_gym_b200.population import PathNetConfig
    return PathNetConfig(L=cfg.L, M=cfg.M, N=cfg.N, H=cfg.H, W=cfg.W, C=cfg.C, kernels=cfg.kernels,
                         strides=cfg.strides, F=cfg.F, hidden=cfg.hidden, A=cfg.A, lstm=getattr(cfg, "lstm", False))


def rel_err(x, ref):
    """max |x - ref| / max |ref|  -- the tolerance north_star states (1e-3 relative) is applied
    per tensor against the tensor's largest magnitude."""
    x = np.asarray(x, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    den = np.max(np.abs(ref))
    if den == 0:
        return float(np.max(np.abs(x)))
    return float(np.max(np.abs(x - ref)) / den)


def split_flat(cfg, flat):
    out, o = [], 0
    for t in po.tensor_table(cfg):
        n = int(np.prod(t["shape"]))
        out.append(np.asarray(flat[o:o + n]).reshape(t["shape"]))
        o += n
    assert o == len(flat)
    return out

# BASELINE config C4: second-task transfer, L=4 M=10 N=4, 84x84x4 frames, 3-action head
C4 = po.OracleConfig(L=4, M=10, N=4, H=84, W=84, C=4, A=3)

# the LSTM variant (the reference's default network, constants.py:31) on the C2 / C3 nets and on a small geometry
C1_LSTM = po.OracleConfig(L=3, M=10, N=3, H=84, W=84, C=4, A=18, lstm=True)
C3_LSTM = po.OracleConfig(L=4, M=10, N=4, H=84, W=84, C=4, A=18, lstm=True)
TC20_LSTM = po.OracleConfig(L=3, M=10, N=3, H=20, W=24, C=4, kernels=(8, 2, 3), strides=(4, 1, 1), F=8, hidden=256, A=6, lstm=True)
