"""Timed CPU baseline: the oracle restatement run in the reference's own execution shape.
TEST / BENCH INFRASTRUCTURE ONLY (bench.py `cpu_baseline` and `--impl reference`).

The reference's arithmetic lives in TensorFlow 1.x, which cannot be installed here
(BASELINE.md), so the baseline is kind "port": torch-CPU fp32, every one of the M modules
of every layer evaluated and gated (game_ac_network.py:186-193), ONE path at a time (the
reference runs one path per process), forward + loss + autograd backward + per-tensor
clip_by_norm + RMSProp (rmsprop_applier.py:79-106), frozen tensors skipped
(a3c_training_thread.py:205-209).
"""
import time

import numpy as np
import torch

from . import pathnet_oracle as po


class CpuPathWorker:
    def __init__(self, cfg: po.OracleConfig, seed=1, threads=None):
        if threads:
            torch.set_num_threads(int(threads))
        self.cfg = cfg
        self.params = [torch.tensor(p, dtype=torch.float32, requires_grad=True) for p in po.init_params(cfg, seed)]
        self.ms = [torch.ones_like(p) for p in self.params]
        self.frozen = [0] * len(self.params)

    def step(self, frames_u8, gates, actions, td, r, lr=7.0016e-4, beta=po.ENTROPY_BETA):
        """One a3c update of one path (a3c_training_thread.py:155-216 without the env rollout)."""
        cfg = self.cfg
        x = torch.from_numpy(frames_u8).to(torch.float32) * (1.0 / 255.0)
        g = torch.tensor(np.asarray(gates), dtype=torch.float32)
        pi, v = po.forward_torch(cfg, self.params, x, g)
        a1h = torch.zeros(len(actions), cfg.A)
        a1h[torch.arange(len(actions)), torch.as_tensor(np.asarray(actions), dtype=torch.long)] = 1.0
        loss = po.loss_torch(pi, v, a1h, torch.as_tensor(td), torch.as_tensor(r), beta)
        grads = torch.autograd.grad(loss, self.params, allow_unused=True)
        with torch.no_grad():
            for p, m, gr, fz in zip(self.params, self.ms, grads, self.frozen):
                if fz or gr is None:
                    continue
                nrm = torch.sqrt(torch.sum(gr * gr))
                gc = gr * (po.GRAD_NORM_CLIP / torch.clamp(nrm, min=po.GRAD_NORM_CLIP))
                m.add_((gc * gc - m) * (1.0 - po.RMSP_ALPHA))
                p.sub_(lr * gc / torch.sqrt(m + po.RMSP_EPSILON))
        return float(loss)


def time_cpu_baseline(cfg, B, n_steps, warmup=1, threads=None, seed=1234):
    """Returns (env_steps_per_s, seconds, threads used) for n_steps single-path updates of B transitions."""
    w = CpuPathWorker(cfg, threads=threads)
    rng = np.random.Generator(np.random.PCG64(seed))
    frames = rng.integers(0, 256, size=(B, cfg.H, cfg.W, cfg.C), dtype=np.uint8)
    actions = rng.integers(0, cfg.A, size=B)
    td = rng.normal(size=B).astype(np.float32)
    r = rng.normal(size=B).astype(np.float32)
    gates = [po.gates_from_masks(cfg, po.get_geopath_masks(cfg.L, cfg.M, cfg.N, 7, 0, p)) for p in range(n_steps + warmup)]
    for i in range(warmup):
        w.step(frames, gates[i], actions, td, r)
    t0 = time.perf_counter()
    for i in range(n_steps):
        w.step(frames, gates[warmup + i], actions, td, r)
    dt = time.perf_counter() - t0
    return B * n_steps / dt, dt, torch.get_num_threads()
