#!/usr/bin/env python
"""bench.py -- env-steps/s of the PathNet population step on B200 (BASELINE.json metric).

One step = forward + A3C loss + backward of every local path (P paths x B transitions),
all-reduce of the shared gradient slab when N > 1, per-tensor clip + masked RMSProp.
Workload = BASELINE config C2 per GPU (L=3, M=10, N=3, 64 paths x batch 32, 84x84x4 u8
synthetic frames); N GPUs hold N x 64 paths (weak scaling; C5 = 512 paths on 8 GPUs).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
  python bench.py --impl reference [...]                          # CPU restatement of the reference

For N > 1 launch with torchrun (RANK / LOCAL_RANK / WORLD_SIZE from the env); rank 0 prints
ONE JSON line.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "env-steps/sec (population fwd+bwd+RMSProp)"
UNIT = "env-steps/s"
P_PER_GPU, BATCH, NSETS = 64, 32, 4
LR = 7.0016e-4
DTYPE = "u8 x s8 -> s32 (layer 0), bf16 hi/lo x3 -> f32 (linear), f32 elsewhere"
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def workload_config(n_gpus):
    return {"workload": "C2 per GPU: PathNet A3C L=3 M=10 N=3, %d paths x batch %d, 84x84x4 u8 frames; "
                        "fwd+loss+bwd(+grad all-reduce)+per-tensor clip+masked RMSProp" % (P_PER_GPU, BATCH),
            "paths_per_gpu": P_PER_GPU, "paths_total": P_PER_GPU * n_gpus, "batch": BATCH,
            "parallelism": "population sharded over %d GPU(s)" % n_gpus,
            "l2": "inputs rotate over %d frame sets (%.0f MB > 126 MB L2) so each step reads its frames from HBM"
                  % (NSETS, NSETS * P_PER_GPU * BATCH * 84 * 84 * 4 / 1e6)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return {k: float(d[k]) for k in FALLBACK_PEAKS}, "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return dict(FALLBACK_PEAKS), "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------
# clocks / throttle reasons sampled DURING the timed regions
# ---------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self.active = threading.Event()
        self.stop_flag = False
        self.err = None

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            while not self.stop_flag:
                if self.active.is_set():
                    self.samples.append(int(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                    try:
                        bits = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                    except Exception:
                        bits = int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    for b, name in self.REASONS.items():
                        if bits & b:
                            self.reasons.add(name)
                time.sleep(0.005)
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def summary(self):
        out = {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
               "reasons": sorted(self.reasons), "samples": len(self.samples)}
        if self.err:
            out["error"] = self.err
        return out


# ---------------------------------------------------------------------------
# algorithmic work per kernel launch for the roofline object
# ---------------------------------------------------------------------------
def kernel_work(name, pop, B):
    """(bound, algorithmic flops or bytes per launch) of a kernel, counting ACTIVE modules only."""
    import numpy as np
    cfg = pop.cfg
    pc = lambda l: float(sum(bin(int(m)).count("1") for m in pop.masks[:, l]))   # sum over paths of n_l
    dims, ih, iw, cin = [], cfg.H, cfg.W, cfg.C
    for l in range(cfg.L - 1):
        k, s = cfg.kernels[l], cfg.strides[l]
        oh, ow = (ih - k) // s + 1, (iw - k) // s + 1
        dims.append((oh * ow, k * k * cin))
        ih, iw, cin = oh, ow, cfg.F
    T = pop.P * B
    if name.startswith(("conv_fwd_l", "conv_wgrad_l", "conv_dgrad_l")) and "reduce" not in name:
        l = int(name.rsplit("l", 1)[1])
        return "tensor", 2.0 * dims[l][0] * dims[l][1] * cfg.F * pc(l) * B
    if name in ("lin_fwd", "lin_dgrad", "lin_wgrad"):
        return "tensor", 2.0 * pop.last_lin * cfg.hidden * pc(cfg.L - 1) * B
    if name == "head_loss":          # logits + dh + chunk partials of the head weight gradient in one kernel
        return "tensor", 2.0 * cfg.hidden * (cfg.A + 1) * 3 * T
    nonfrozen = float(sum(t["length"] for t, f in zip(pop.tensors, pop.frozen) if not f))
    if name == "rms_update":
        return "hbm", 20.0 * nonfrozen          # read g, var, ms; write var, ms (momentum slot elided)
    if name == "rms_sumsq":
        return "hbm", 4.0 * nonfrozen
    if name.startswith("conv_wgrad_reduce_l"):
        l = int(name.rsplit("l", 1)[1])
        return "hbm", 4.0 * (dims[l][1] * cfg.F + cfg.F) * (pc(l) + cfg.M)
    return "hbm", 0.0


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
    from pathnet_gym_b200.trainer import PipelinedHostTrainer, PopulationTrainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    cfg = PathNetConfig()
    P, B, K, W = P_PER_GPU, BATCH, args.steps, max(args.warmup, 3)
    pop = PathNetPopulation(cfg, P, B, device=local)
    pop.init_params(1)                                      # same seed on every rank: replicated parameters
    masks = geopath_init(P, cfg.L, cfg.M, cfg.N, seed=7, epoch=0, path_id0=rank * P, device=local)
    pop.set_genotypes(masks.cpu().numpy().view(np.uint32))
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    frames = [torch.randint(0, 256, (P, B, cfg.H, cfg.W, cfg.C), dtype=torch.uint8, device=dev, generator=gen)
              for _ in range(NSETS)]
    actions = torch.randint(0, cfg.A, (P, B), dtype=torch.int32, device=dev, generator=gen)
    td = torch.randn(P, B, device=dev, generator=gen)
    r = torch.randn(P, B, device=dev, generator=gen)

    trainer = PopulationTrainer(pop)          # public API: train_step -> all-reduce(grads) -> clipped RMSProp

    def step(i):
        trainer.step(frames[i % NSETS], actions, td, r, LR)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()
    for i in range(W):
        step(i)
    barrier()
    # timed region: exactly K steps, nothing but the product's own launches on the stream
    launches0 = pop.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active.set()
    barrier()
    ev0.record()
    for i in range(K):
        step(W + i)
    ev1.record()
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = pop.launch_count - launches0
    # the same K steps again with a CUDA-event pair around every kernel launch (on the launching stream): the
    # per-kernel durations behind the roofline objects.  The event records cost a few percent, so this pass is
    # reported beside the headline (ms_per_step_instrumented), not as the headline.
    pop.profile_enable(True)
    barrier()
    ev0.record()
    for i in range(K):
        step(W + K + i)
    ev1.record()
    barrier()
    sampler.active.clear()
    ms_instr = float(ev0.elapsed_time(ev1))
    prof = pop.profile_read()
    pop.profile_enable(False)

    # ---- end to end: host buffers through the C-ABI host entry point ----
    hf = [f.cpu().pin_memory() for f in frames]
    ha, htd, hr = actions.cpu().pin_memory(), td.cpu().pin_memory(), r.cpu().pin_memory()
    hloss = torch.empty(P, dtype=torch.float32).pin_memory()

    # The repo's public host-buffer API: batch i+1 is staged (pinned host -> device, own copy stream) while step i
    # runs; every step's inputs cross PCIe inside the timed region and every step's losses are read back.
    pipe = PipelinedHostTrainer(pop)
    batch = lambda i: (hf[i % NSETS], ha, htd, hr)
    pipe.prime(batch(0))
    for i in range(3):
        pipe.step(batch(i + 1), LR, loss_out=hloss)
    torch.cuda.synchronize()
    sampler.active.set()
    barrier()
    ev0.record()
    for i in range(K):
        pipe.step(batch(3 + i + 1) if i + 1 < K else None, LR, loss_out=hloss)
    ev1.record()
    barrier()
    sampler.active.clear()
    ms2 = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_ms = float(ms2.item())
    sampler.stop_flag = True
    sampler.join(timeout=2)

    if rank == 0:
        peaks, peak_src = load_peaks()
        units = P * B * world
        kern = {n: {"avg_ms": round(a, 5), "launches": c} for n, (a, c) in prof.items()}
        tot = sum(a * c for a, c in prof.values()) or 1.0
        for n, (a, c) in prof.items():
            kern[n]["share"] = round(a * c / tot, 4)
        dom = max(prof, key=lambda n: prof[n][0] * prof[n][1])
        bound, work = kernel_work(dom, pop, B)
        dur = prof[dom][0] * 1e-3
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(dom)
            except Exception:
                traffic = None
        if bound == "tensor":
            ach, peak, runit = work / dur / 1e12, peaks["bf16_tflops_sustained"], "TFLOP/s"
        else:
            ach, peak, runit = work / dur / 1e9, peaks["hbm_gbs"], "GB/s"
        roof = {"kernel": dom, "bound": bound, "achieved": round(ach, 3), "peak": peak, "unit": runit,
                "frac": round(ach / peak, 5), "traffic": traffic, "peak_source": peak_src,
                "work_per_launch": work, "avg_launch_ms": round(prof[dom][0], 5)}
        tensor_rows = []
        for kn in ("conv_fwd_l0", "conv_wgrad_l0", "conv_fwd_l1", "conv_dgrad_l1", "conv_wgrad_l1", "lin_fwd", "lin_dgrad", "lin_wgrad"):
            if kn in prof:
                _, wk = kernel_work(kn, pop, B)
                a = wk / (prof[kn][0] * 1e-3) / 1e12
                tensor_rows.append({"kernel": kn, "bound": "tensor", "achieved": round(a, 2), "peak": peaks["bf16_tflops_sustained"],
                                    "unit": "TFLOP/s", "frac": round(a / peaks["bf16_tflops_sustained"], 4),
                                    "avg_launch_ms": round(prof[kn][0], 5)})
        rb, rwork = kernel_work("rms_update", pop, B)
        rms = prof.get("rms_update")
        roof_rms = None
        if rms:
            a = rwork / (rms[0] * 1e-3) / 1e9
            roof_rms = {"kernel": "rms_update", "bound": "hbm", "achieved": round(a, 2), "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "frac": round(a / peaks["hbm_gbs"], 4), "traffic": None,
                        "note": "slabs (20 MB) are L2-resident between steps; see DESIGN.md"}
        fbytes = P * B * cfg.H * cfg.W * cfg.C
        line = {"metric": METRIC, "value": round(units * K / (ms_total * 1e-3), 1), "unit": UNIT, "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": round(ms_total / K, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
                "config": workload_config(world), "clocks": sampler.summary(),
                "e2e": {"value": round(units * K / (e2e_ms * 1e-3), 1), "unit": UNIT,
                        "h2d_bytes_per_step": fbytes + 3 * P * B * 4, "d2h_bytes_per_step": P * 4,
                        "ms_per_step": round(e2e_ms / K, 4),
                        "api": "PipelinedHostTrainer: pn_stage_inputs_host (pinned host buffers, copy stream) + "
                               "pn_train_step_staged + pn_rmsprop_apply; losses copied back every step"},
                "gpu_launches": int(launches), "roofline": roof, "roofline_rmsprop": roof_rms,
                "roofline_tensor_kernels": tensor_rows, "kernels": kern,
                "ms_per_step_instrumented": round(ms_instr / K, 4),
                "workspace_mb": round(pop.workspace_bytes / 1e6, 1)}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(target_seconds=12.0)
        else:
            line["cpu_baseline"] = None
        emit(line)
    pop.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(target_seconds=12.0, steps=None, warmup=1):
    """Oracle port timed on this box's host cores: one path at a time, all M modules gated."""
    from oracle import cpu_baseline as cb, pathnet_oracle as po
    cores = os.cpu_count() or 1
    cfg = po.OracleConfig()
    if steps is None:
        v, dt, _ = cb.time_cpu_baseline(cfg, BATCH, 4, warmup=warmup, threads=cores)
        steps = int(max(8, min(2000, target_seconds / (dt / 4))))
    v, dt, thr = cb.time_cpu_baseline(cfg, BATCH, steps, warmup=warmup, threads=cores)
    return {"value": round(v, 1), "unit": UNIT, "cores": thr, "kind": "port",
            "sample": "%d single-path updates of batch %d (C2 net, all %d modules/layer evaluated and gated, torch-CPU "
                      "fp32, fwd+loss+bwd+clip+RMSProp), %.1f s; TensorFlow-1 reference itself is not installable"
                      % (steps, BATCH, cfg.M, dt)}


def run_reference(args):
    """Reference arm: the reference's own execution shape on the host cores (oracle port; the
    TF-1 reference cannot be installed here -- see DESIGN.md).  Each step = a bounded sample of
    the C2 workload: 4 of the paths, one after another, batch 32."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline as cb, pathnet_oracle as po
    cores = os.cpu_count() or 1
    K, W, per_step = args.steps, max(args.warmup, 1), 4
    cfg = po.OracleConfig()
    v, dt, thr = cb.time_cpu_baseline(cfg, BATCH, K * per_step, warmup=W * per_step, threads=cores)
    n = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    line = {"impl": "reference", "metric": METRIC, "value": round(v, 1), "unit": UNIT, "n_gpus": n, "steps": K,
            "warmup": W, "ms_per_step": round(dt / K * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (torch CPU)", "data": "synthetic", "config": workload_config(n),
            "cpu_baseline": {"value": round(v, 1), "unit": UNIT, "cores": thr, "kind": "port",
                             "sample": "each step = %d single-path updates of batch %d of the C2 workload on the host "
                                       "cores (all M modules evaluated and gated, torch-CPU fp32)" % (per_step, BATCH)},
            "e2e": {"value": round(v, 1), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the process's original stdout; everything else (NCCL banners, warnings) to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, data)
    else:
        sys.stdout.write(data.decode())
        sys.stdout.flush()


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
