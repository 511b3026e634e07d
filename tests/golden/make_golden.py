#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/.

Run in the BUILD container only (needs /root/reference):   python tests/golden/make_golden.py

1. genotype_golden.npz -- produced by the REAL reference functions
   pathnet.mutation / pathnet.get_geopath (/root/reference/pathnet.py:50-63,78-87).
   `tensorflow` is stubbed (those two functions only use NumPy) and
   np.random.rand / np.random.randint are replaced by the counter-based uniform
   stream of SURVEY App. A.6, so the reference code itself decides every bit.
   The oracle restatement (oracle/pathnet_oracle.py) and the CUDA kernel are both
   tested against these arrays.
2. float_golden_tiny.npz / float_golden_c1.npz -- fp64 "truth" of the oracle for the
   floating-point path (parity UNPINNED against TensorFlow: TF is not installable
   here and the reference holds no numeric fixtures).  They pin the oracle against
   drift and give the GPU tests precomputed expectations.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import pathnet_oracle as po  # noqa: E402

REF = "/root/reference"


def load_reference_pathnet():
    class _Stub(types.ModuleType):
        """`import tensorflow as tf` must succeed and `tf.nn.relu` (a default argument of the
        dead MNIST-era builders, pathnet.py:122) must resolve; nothing in it is ever called."""
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _Stub(name)
    sys.modules["tensorflow"] = _Stub("tensorflow")
    try:
        spec = importlib.util.spec_from_file_location("ref_pathnet", os.path.join(REF, "pathnet.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        del sys.modules["tensorflow"]   # do not leave the stub around for other importers (torch inspects modules)
    return mod


class SpyArray(np.ndarray):
    """ndarray that remembers the last (i, j) it was read at: the reference reads
    geopath[i,j] immediately before each np.random.rand() of the scan."""
    last = (0, 0)

    def __getitem__(self, key):
        if isinstance(key, tuple) and len(key) == 2:
            SpyArray.last = (int(key[0]), int(key[1]))
        return super().__getitem__(key)


def run_reference_mutation(ref, gates, L, M, N, seed, epoch, path_id):
    arr = np.array(gates, dtype=float).view(SpyArray)
    state = {}

    def rand():
        i, j = SpyArray.last
        state["pos"] = i * M + j
        return po.philox_uniform(seed, epoch, path_id, i * M + j, po.DOMAIN_MUTATE, 0)

    def randint(lo, hi):
        assert lo == 0 and hi == M
        return int(po.philox_uniform(seed, epoch, path_id, state["pos"], po.DOMAIN_MUTATE, 1) * M)

    o_rand, o_randint = np.random.rand, np.random.randint
    np.random.rand, np.random.randint = rand, randint
    try:
        out = ref.mutation(arr, L, M, N)
    finally:
        np.random.rand, np.random.randint = o_rand, o_randint
    return np.asarray(out, dtype=float)


def run_reference_get_geopath(ref, L, M, N, seed, epoch, path_id):
    st = dict(layer=0, slot=0, seen=set())

    def rand():
        u = po.philox_uniform(seed, epoch, path_id, st["layer"], po.DOMAIN_INIT, st["slot"])
        st["slot"] += 1
        st["seen"].add(int(u * M))
        if len(st["seen"]) == N:
            st["layer"] += 1
            st["slot"] = 0
            st["seen"] = set()
        return u

    o_rand = np.random.rand
    np.random.rand = rand
    try:
        out = ref.get_geopath(L, M, N)
    finally:
        np.random.rand = o_rand
    return np.asarray(out, dtype=float)


def make_genotype_golden():
    ref = load_reference_pathnet()
    cases = []
    cfgs = [(3, 10, 3), (4, 10, 4), (3, 4, 2), (2, 5, 1), (4, 16, 6), (5, 32, 4), (3, 8, 8)]
    for (L, M, N) in cfgs:
        for seed in (7, 0xDEADBEEFCAFE, 123456789012345):
            for epoch in range(0, 24):
                pid = (epoch * 37 + 11) % 512
                g0 = run_reference_get_geopath(ref, L, M, N, seed, epoch, pid)
                m0 = po.masks_from_gates(g0)
                g1 = run_reference_mutation(ref, g0, L, M, N, seed, epoch + 1000, pid)
                m1 = po.masks_from_gates(g1)
                # second mutation on an already-mutated (possibly != N per row) genotype
                g2 = run_reference_mutation(ref, g1, L, M, N, seed, epoch + 2000, pid)
                m2 = po.masks_from_gates(g2)
                row = np.zeros(5 + 3 * 8, dtype=np.uint64)
                row[:5] = (L, M, N, epoch, pid)
                row[5:5 + L] = m0
                row[13:13 + L] = m1
                row[21:21 + L] = m2
                cases.append((seed, row))
    seeds = np.asarray([c[0] for c in cases], dtype=np.uint64)
    rows = np.stack([c[1] for c in cases])
    np.savez_compressed(os.path.join(HERE, "genotype_golden.npz"), seeds=seeds, rows=rows)
    print("genotype_golden: %d cases" % len(cases))


TINY = po.OracleConfig(L=3, M=4, N=2, H=20, W=20, C=4, kernels=(4, 3, 3), strides=(2, 1, 1), F=8, hidden=32, A=5)
TINY4 = po.OracleConfig(L=4, M=5, N=2, H=28, W=24, C=4, kernels=(4, 4, 3), strides=(2, 2, 1), F=8, hidden=64, A=3)
C1 = po.OracleConfig(L=3, M=10, N=3, H=84, W=84, C=4, A=18)


def synth_inputs(cfg, P, B, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    frames = rng.integers(0, 256, size=(P, B, cfg.H, cfg.W, cfg.C), dtype=np.uint8)
    actions = rng.integers(0, cfg.A, size=(P, B)).astype(np.int32)
    td = rng.normal(size=(P, B)).astype(np.float32)
    r = (rng.uniform(-1, 1, size=(P, B)) * 5.0).astype(np.float32)
    masks = np.stack([po.get_geopath_masks(cfg.L, cfg.M, cfg.N, 7, 0, p) for p in range(P)])
    return frames, actions, td, r, masks


def make_float_golden(cfg, name, P, B, full_grads, grad_scale=1.0):
    params = po.init_params(cfg, 1)
    frames, actions, td, r, masks = synth_inputs(cfg, P, B, 1234)
    td = (td * grad_scale).astype(np.float32)
    pi, v, loss, G = po.population_step(cfg, params, frames, masks, actions, td, r, dtype="float64")
    # freeze module 0 of every layer in the update check
    fixed = np.zeros((cfg.L, cfg.M))
    fixed[:, 0] = 1.0
    frozen = po.frozen_flags(cfg, fixed)
    lr = 7.0016e-4
    p64 = [p.astype(np.float64) for p in params]
    ms0 = [np.ones_like(p) for p in p64]
    mom0 = [np.zeros_like(p) for p in p64]
    P2, S2, _ = po.apply_gradients(p64, ms0, mom0, G, frozen, lr)
    out = dict(frames=frames, actions=actions, td=td, r=r, masks=masks, pi=pi, v=v, loss=loss, frozen=frozen,
               lr=np.float64(lr), cfg=np.asarray([cfg.L, cfg.M, cfg.N, cfg.H, cfg.W, cfg.C, cfg.F, cfg.hidden, cfg.A,
                                                  *cfg.kernels, *cfg.strides], dtype=np.int64))
    norms = np.asarray([np.sqrt(np.sum(g * g)) for g in G])
    out["grad_norms"] = norms
    out["grad_sums"] = np.asarray([g.sum() for g in G])
    if not full_grads:
        # big config: frames are regenerated in the tests from synth_inputs(seed 1234); keep a checksum only
        import zlib
        out["frames_crc"] = np.uint32(zlib.crc32(frames.tobytes()))
        del out["frames"]
    if full_grads:
        out["grads"] = np.concatenate([g.ravel() for g in G])
        out["params_after"] = np.concatenate([p.ravel() for p in P2])
        out["ms_after"] = np.concatenate([s.ravel() for s in S2])
    else:
        # strided sample of each tensor (every 97th element) keeps the fixture small
        out["grads_sample"] = np.concatenate([g.ravel()[::97] for g in G])
        out["params_after_sample"] = np.concatenate([p.ravel()[::97] for p in P2])
        out["ms_after_sample"] = np.concatenate([s.ravel()[::97] for s in S2])
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "loss", loss, "max grad norm", norms.max())


if __name__ == "__main__":
    make_genotype_golden()
    make_float_golden(TINY, "float_golden_tiny.npz", P=3, B=5, full_grads=True)
    make_float_golden(TINY4, "float_golden_tiny4.npz", P=4, B=3, full_grads=True, grad_scale=40.0)
    make_float_golden(C1, "float_golden_c1.npz", P=1, B=20, full_grads=False)
