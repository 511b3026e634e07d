#!/usr/bin/env python
"""bench.py -- env-steps/s of the PathNet population step on B200 (BASELINE.json metric).

One step = forward + A3C loss + backward of every local path (P paths x B transitions),
sum of the shared gradient slab over the GPUs when N > 1, per-tensor clip + masked RMSProp,
and -- for the configurations that name it -- the genotype tournament every 1000 steps.

  python bench.py [--config c3] [--gpus N] [--steps K] [--warmup W]   # this repo's CUDA path
  python bench.py --impl reference [...]                              # CPU restatement of the reference

Configurations (BASELINE.json `configs`):
  c2  L=3 M=10 N=3,  64 paths x 32, A=18
  c3  L=4 M=10 N=4, 256 paths x 32, A=18, tournament selection + mutation every 1000 steps   (default:
      the largest single-GPU configuration)
  c4  second-task transfer: the c3 net with a 3-action head, task-1 winner frozen, fresh population OR'ed
      with the frozen path, 64 paths x 32, tournament every 1000 steps
  c5  = `--gpus N`: the main line is weak scaling (the chosen config on every GPU); the `c5_strong` object in
      the same line holds 512 paths of the L=4 net IN TOTAL split over the N GPUs.
At one GPU the line also carries `lstm_variant`: the same configuration with the reference's default network
(GameACPathNetLSTMNetwork, constants.py:31 USE_LSTM = True), device-timed like the main value.

For N > 1 launch with torchrun (RANK / LOCAL_RANK / WORLD_SIZE from the env); rank 0 prints ONE JSON line.
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
BATCH = 32
LR = 7.0016e-4
TOURNAMENT_EVERY = 1000
DTYPE = ("u8 x s8 -> s32 (layer 0), fp16 hi/lo x3 -> f32 (forward of the middle conv layers and linear modules), "
         "bf16 hi/lo x3 -> f32 (their backward), f32 elsewhere")
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}

CONFIGS = {
    "c2": dict(L=3, N=3, A=18, P=64, tournament=False, transfer=False,
               text="C2: PathNet A3C L=3 M=10 N=3, 64 paths x batch 32"),
    "c3": dict(L=4, N=4, A=18, P=256, tournament=True, transfer=False,
               text="C3: PathNet A3C L=4 M=10 N=4, 256 paths x batch 32, tournament selection + mutation every "
                    "1000 steps (population already mutated by 20 tournament rounds)"),
    "c4": dict(L=4, N=4, A=3, P=64, tournament=True, transfer=True,
               text="C4: second-task transfer, L=4 M=10 N=4, 3-action head, task-1 winner's path frozen, fresh "
                    "population OR'ed with it, 64 paths x batch 32, tournament every 1000 steps"),
}


def frame_sets(P):
    """Enough distinct frame sets that every step reads its frames from HBM (> 126 MB L2 in rotation)."""
    per_set = P * BATCH * 84 * 84 * 4
    return max(2, -(-260_000_000 // per_set))


def workload_config(name, n_gpus, P):
    c = CONFIGS[name]
    ns = frame_sets(P)
    return {"workload": c["text"] + ", 84x84x4 u8 synthetic frames; fwd+loss+bwd(+grad sum over GPUs)+per-tensor "
                                    "clip+masked RMSProp per GPU",
            "name": name, "paths_per_gpu": P, "paths_total": P * n_gpus, "batch": BATCH,
            "tournament_every_steps": TOURNAMENT_EVERY if c["tournament"] else None,
            "parallelism": "population sharded over %d GPU(s), weak scaling" % n_gpus,
            "l2": "inputs rotate over %d frame sets (%.0f MB > 126 MB L2) so each step reads its frames from HBM"
                  % (ns, ns * P * BATCH * 84 * 84 * 4 / 1e6)}


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
# algorithmic work per kernel launch for the roofline objects
# ---------------------------------------------------------------------------
def layer_dims(cfg):
    dims, ih, iw, cin = [], cfg.H, cfg.W, cfg.C
    for l in range(cfg.L - 1):
        k, s = cfg.kernels[l], cfg.strides[l]
        oh, ow = (ih - k) // s + 1, (iw - k) // s + 1
        dims.append((oh * ow, k * k * cin))
        ih, iw, cin = oh, ow, cfg.F
    return dims


def kernel_work(name, pop, B):
    """(bound, algorithmic flops or bytes per launch) of a kernel, counting ACTIVE modules only."""
    cfg = pop.cfg
    pc = lambda l: float(sum(bin(int(m)).count("1") for m in pop.masks[:, l]))   # sum over paths of n_l
    dims = layer_dims(cfg)
    T = pop.P * B
    if name.startswith(("conv_fwd_l", "conv_wgrad_l", "conv_dgrad_l")) and "reduce" not in name:
        l = int(name.rsplit("l", 1)[1])
        return "tensor", 2.0 * dims[l][0] * dims[l][1] * cfg.F * pc(l) * B
    if name in ("lin_fwd", "lin_dgrad", "lin_wgrad"):
        return "tensor", 2.0 * pop.last_lin * cfg.hidden * pc(cfg.L - 1) * B
    if name == "head_loss":          # logits + dh + chunk partials of the head weight gradient in one kernel
        return "tensor", 2.0 * cfg.hidden * (cfg.A + 1) * 3 * T
    return "hbm", 0.0


def layer0_hbm_bytes(name, pop, B):
    """Algorithmic HBM bytes of the two layer-0 kernels (they are HBM-side: every frame byte once per kernel)."""
    cfg = pop.cfg
    T = pop.P * B
    npix = layer_dims(cfg)[0][0]
    n0 = float(sum(bin(int(m)).count("1") for m in pop.masks[:, 0])) * B        # (transition, active module) pairs
    frames = float(T) * cfg.H * cfg.W * cfg.C
    if name == "conv_fwd_l0":        # read frames; write the layer output (fp32) and one ReLU byte per (module, pixel)
        return frames + T * npix * cfg.F * 4.0 + n0 * npix
    if name == "conv_wgrad_l0":      # read frames, the output gradient (fp32) and the ReLU bytes
        return frames + T * npix * cfg.F * 4.0 + n0 * npix
    return 0.0


def rms_bytes(pop, first, count, complement):
    """20 B per non-frozen parameter of the selected tensor range (read g, var, ms; write var, ms)."""
    tot = 0.0
    for i, (t, f) in enumerate(zip(pop.tensors, pop.frozen)):
        inside = first <= i < first + count
        if inside != complement and not f:
            tot += t["length"]
    return 20.0 * tot


def synthetic_scores(P_total, rnd, device):
    """Deterministic episode scores, identical on every rank (the bench has no environment)."""
    import torch
    g = torch.Generator(device="cpu")
    g.manual_seed(9000 + rnd)
    return torch.rand(P_total, generator=g).to(device) * 100.0


def build_population(name, P, rank, world, local, pre_rounds=20, lstm=False):
    """Population + controller in the state the configuration describes."""
    import numpy as np
    import torch
    from pathnet_gym_b200.controller import PopulationController
    from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
    c = CONFIGS[name]
    cfg = PathNetConfig(L=c["L"], N=c["N"], A=c["A"], lstm=lstm)
    pop = PathNetPopulation(cfg, P, BATCH, device=local)
    pop.init_params(1)                                      # same seed on every rank: replicated parameters
    ctl = None
    if c["tournament"]:
        ctl = PopulationController(pop, P * world, Bcand=3, seed=7, rank=rank, world_size=world)
        ctl.start_task()
        rounds = 0
        if c["transfer"]:                                   # task 1 -> freeze the winner -> task-2 population
            for rnd in range(pre_rounds):
                ctl.scores.copy_(synthetic_scores(P * world, rnd, pop.device))
                ctl.tournament_step()
            ctl.end_task()
            ctl.start_task()
            rounds = pre_rounds
        for rnd in range(rounds, rounds + pre_rounds):      # a population in mid-task: mutation has added / dropped modules
            ctl.scores.copy_(synthetic_scores(P * world, rnd, pop.device))
            ctl.tournament_step()
    else:
        masks = geopath_init(P, cfg.L, cfg.M, cfg.N, seed=7, epoch=0, path_id0=rank * P, device=local)
        pop.set_genotypes(masks.cpu().numpy().view(np.uint32))
    return cfg, pop, ctl


class Workload(object):
    """Device-resident inputs + the step closure (train step -> gradient sum -> update [-> tournament])."""

    def __init__(self, name, P, rank, world, local, seed_off=0, lstm=False):
        import torch
        from pathnet_gym_b200.trainer import PopulationTrainer
        self.name, self.P, self.rank, self.world = name, P, rank, world
        self.cfg, self.pop, self.ctl = build_population(name, P, rank, world, local, lstm=lstm)
        dev = self.pop.device
        gen = torch.Generator(device=dev)
        gen.manual_seed(1234 + rank + seed_off)
        cfg = self.cfg
        self.nsets = frame_sets(P)
        self.frames = [torch.randint(0, 256, (P, BATCH, cfg.H, cfg.W, cfg.C), dtype=torch.uint8, device=dev, generator=gen)
                       for _ in range(self.nsets)]
        self.actions = torch.randint(0, cfg.A, (P, BATCH), dtype=torch.int32, device=dev, generator=gen)
        self.td = torch.randn(P, BATCH, device=dev, generator=gen)
        self.r = torch.randn(P, BATCH, device=dev, generator=gen)
        self.trainer = PopulationTrainer(self.pop)
        self.gstep = 0
        self.tournaments = 0

    def maybe_tournament(self):
        """doom_pathnet.py:241-270 every TOURNAMENT_EVERY steps: scores in, winners copied + mutated, new masks loaded."""
        if self.ctl is not None and self.gstep % TOURNAMENT_EVERY == 0:
            self.ctl.scores.copy_(synthetic_scores(self.P * self.world, 100 + self.gstep // TOURNAMENT_EVERY, self.pop.device))
            self.ctl.tournament_step()
            self.tournaments += 1

    def step(self):
        self.gstep += 1
        self.trainer.step(self.frames[self.gstep % self.nsets], self.actions, self.td, self.r, LR)
        self.maybe_tournament()

    def close(self):
        self.pop.close()


def timed(wl, K, barrier, torch):
    """Exactly K steps between two events on the launch stream; one tournament falls into the region (the step counter
    is placed so that step K/2 is a multiple of TOURNAMENT_EVERY: with K < 1000 that over-represents its cost)."""
    wl.gstep = TOURNAMENT_EVERY - max(1, K // 2)
    t0 = wl.tournaments
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(K):
        wl.step()
    ev1.record()
    barrier()
    return ev0.elapsed_time(ev1), wl.tournaments - t0


def param_checksum(pop, torch):
    """Order-independent 64-bit checksum of the parameter slab's bit patterns."""
    v = pop.params.view(torch.int32).to(torch.int64)
    idx = torch.arange(v.numel(), device=v.device, dtype=torch.int64)
    return int(((v * (2 * idx + 1)).sum() & 0x7FFFFFFFFFFFFFFF).item())


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from pathnet_gym_b200.trainer import PipelinedHostTrainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    numa_cpus = None
    if world > 1:                                           # one process per GPU: keep its pinned buffers next to its GPU
        from pathnet_gym_b200.parallel import bind_to_gpu_numa_node
        numa_cpus = bind_to_gpu_numa_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    name = args.config
    P, B, K, W = CONFIGS[name]["P"], BATCH, args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxms(ms):
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    wl = Workload(name, P, rank, world, local)
    pop, cfg = wl.pop, wl.cfg
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(W):
        wl.step()
    barrier()
    # ---- timed region: exactly K steps, nothing but the product's own launches on the stream ----
    launches0 = pop.launch_count
    sampler.active.set()
    ms, ntour = timed(wl, K, barrier, torch)
    ms_total = maxms(ms)
    launches = pop.launch_count - launches0
    checksum = param_checksum(pop, torch)
    same = True
    if world > 1:
        cs = torch.tensor([checksum], device=dev, dtype=torch.int64)
        allcs = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(allcs, cs)
        same = all(int(x.item()) == checksum for x in allcs)
    # ---- the same K steps with a CUDA-event pair around every kernel launch (on the launching stream): the per-kernel
    # durations behind the roofline objects.  The event records cost a few percent, so this pass is reported beside the
    # headline (ms_per_step_instrumented), not as the headline.
    pop.profile_enable(True)
    ms_instr, _ = timed(wl, K, barrier, torch)
    sampler.active.clear()
    prof = pop.profile_read()
    pop.profile_enable(False)

    # ---- end to end: host buffers through the C-ABI host entry points, copies inside the timed region ----
    # (a) rollout form: the B + C - 1 distinct planes of every path's frame stacks cross PCIe (pn_stage_rollout_host)
    # (b) stack form: the full [P, B, H, W, C] stacks cross PCIe (pn_stage_inputs_host)
    nsets = wl.nsets
    gen = torch.Generator(device="cpu")
    gen.manual_seed(77 + rank)
    hplanes = [torch.randint(0, 256, (P, B + cfg.C - 1, cfg.H, cfg.W), dtype=torch.uint8, generator=gen).pin_memory()
               for _ in range(nsets)]
    ha, htd, hr = wl.actions.cpu().pin_memory(), wl.td.cpu().pin_memory(), wl.r.cpu().pin_memory()
    hloss = torch.empty(P, dtype=torch.float32).pin_memory()
    pipe = PipelinedHostTrainer(pop)

    def e2e_run(batch, steps):
        pipe.prime(batch(0))
        for i in range(3):
            pipe.step(batch(i + 1), LR, loss_out=hloss)
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for i in range(steps):
            pipe.step(batch(4 + i) if i + 1 < steps else None, LR, loss_out=hloss)
            wl.gstep += 1
            wl.maybe_tournament()
        ev1.record()
        barrier()
        return maxms(ev0.elapsed_time(ev1))

    sampler.active.set()
    wl.gstep = TOURNAMENT_EVERY - max(1, K // 2)
    e2e_ms = e2e_run(lambda i: (hplanes[i % nsets], ha, htd, hr), K)
    sampler.active.clear()
    e2e_stack_ms = None
    if not args.no_stack_e2e:
        hf = [f.cpu().pin_memory() for f in wl.frames[:2]]
        Ks = max(5, K // 4)
        wl.gstep = 0
        e2e_stack_ms = e2e_run(lambda i: (hf[i % 2], ha, htd, hr), Ks) / Ks
        del hf

    # ---- the RMSProp kernel on its own: the whole slab in one launch, slabs L2-resident and evicted ----
    rms_alone = measure_rmsprop(pop, torch)

    # ---- 512 paths of the L=4 net in total, split over the GPUs (BASELINE config C5, strong scaling) ----
    c5 = None
    if not args.no_c5 and 512 % world == 0:
        masks_main = pop.masks.copy()
        wl2 = Workload("c3", 512 // world, rank, world, local, seed_off=100)
        for _ in range(W):
            wl2.step()
        barrier()
        ms5, nt5 = timed(wl2, K, barrier, torch)
        ms5 = maxms(ms5)
        cs5 = param_checksum(wl2.pop, torch)
        same5 = True
        if world > 1:
            t = torch.tensor([cs5], device=dev, dtype=torch.int64)
            allt = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allt, t)
            same5 = all(int(x.item()) == cs5 for x in allt)
        c5 = {"workload": "C5: 512 paths of the L=4 M=10 N=4 net in total (strong scaling: %d per GPU) x batch 32, "
                          "tournament every 1000 steps" % (512 // world),
              "paths_total": 512, "paths_per_gpu": 512 // world, "value": round(512 * B * K / (ms5 * 1e-3), 1),
              "unit": UNIT, "ms_per_step": round(ms5 / K, 4), "scaling": "strong", "tournaments_in_timed_region": nt5,
              "param_checksum": "%016x" % cs5, "ranks_identical": same5}
        wl2.close()
        del masks_main

    # ---- the same configuration with the reference's default network (constants.py:31 USE_LSTM = True): SURVEY 8f rank 1 ----
    lstm_obj = None
    if not args.no_lstm and world == 1:
        wl3 = Workload(args.config, P, rank, world, local, seed_off=200, lstm=True)
        for _ in range(W):
            wl3.step()
        barrier()
        ms6, nt6 = timed(wl3, K, barrier, torch)
        lstm_obj = {"workload": "%s with GameACPathNetLSTMNetwork (game_ac_network.py:303-521): BasicLSTMCell(256) over each path's "
                                "%d transitions, BPTT, W_lstm / b_lstm in the slab" % (args.config, B),
                    "value": round(P * B * K / (ms6 * 1e-3), 1), "unit": UNIT, "ms_per_step": round(ms6 / K, 4),
                    "tournaments_in_timed_region": nt6}
        wl3.close()

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
                traffic = json.load(open(tpath)).get(name, {}).get(dom)
            except Exception:
                traffic = None
        i8peak = None
        ipath = os.path.join(ROOT, "profiles", "int8_peak.json")
        if os.path.exists(ipath):
            try:
                i8peak = float(json.load(open(ipath))["i8_tops_dense_sustained"])
            except Exception:
                i8peak = None
        # Which roof binds follows from the roofline model: the kernel's arithmetic intensity (algorithmic flops over
        # algorithmic HBM bytes) against the machine balance of the MMA kind it runs (measured peak flops over measured
        # HBM bandwidth).  The layer-0 kernels (kind::i8, 107 flop/B at u8 frames) sit below the balance point of either
        # peak, i.e. on the HBM roof; the other view is kept in `tensor_view`.
        tensor_view = None
        note = ""
        if dom in ("conv_fwd_l0", "conv_wgrad_l0"):
            hb = layer0_hbm_bytes(dom, pop, B)
            tpeak = i8peak if i8peak else peaks["bf16_tflops_sustained"]
            intensity, balance = work / hb, tpeak * 1e12 / (peaks["hbm_gbs"] * 1e9)
            tensor_view = {"achieved": round(work / dur / 1e12, 3), "unit": "TFLOP/s",
                           "frac_of_bf16_peak": round(work / dur / 1e12 / peaks["bf16_tflops_sustained"], 5),
                           "frac_of_int8_peak": round(work / dur / 1e12 / i8peak, 5) if i8peak else None,
                           "flops_per_launch": work}
            note = ("arithmetic intensity %.0f flop/B against a machine balance of %.0f flop/B (%s peak %.0f T/s over %.1f GB/s): "
                    "the HBM roof binds; tensor_view and roofline_layer0 hold the tensor-pipe fractions"
                    % (intensity, balance, "int8" if i8peak else "bf16", tpeak, peaks["hbm_gbs"]))
            if intensity < balance:
                bound, work = "hbm", hb
        if bound == "tensor":
            ach, peak, runit = work / dur / 1e12, peaks["bf16_tflops_sustained"], "TFLOP/s"
            note = note or "fraction of the measured bf16 dense peak (sustained)"
        else:
            ach, peak, runit = work / dur / 1e9, peaks["hbm_gbs"], "GB/s"
        roof = {"kernel": dom, "bound": bound, "achieved": round(ach, 3), "peak": peak, "unit": runit,
                "frac": round(ach / peak, 5), "traffic": traffic, "peak_source": peak_src,
                "work_per_launch": work, "avg_launch_ms": round(prof[dom][0], 5), "tensor_view": tensor_view, "note": note}
        layer0 = []
        for kn in ("conv_fwd_l0", "conv_wgrad_l0"):
            if kn in prof:
                _, wk = kernel_work(kn, pop, B)
                hb = layer0_hbm_bytes(kn, pop, B)
                d = prof[kn][0] * 1e-3
                row = {"kernel": kn, "avg_launch_ms": round(prof[kn][0], 5), "flops_per_launch": wk, "hbm_bytes_per_launch": hb,
                       "tflops": round(wk / d / 1e12, 2), "frac_of_bf16_peak": round(wk / d / 1e12 / peaks["bf16_tflops_sustained"], 4),
                       "hbm_gbs": round(hb / d / 1e9, 1), "frac_of_hbm_peak": round(hb / d / 1e9 / peaks["hbm_gbs"], 4)}
                if i8peak:
                    row["frac_of_int8_peak"] = round(wk / d / 1e12 / i8peak, 4)
                    row["int8_peak_tops"] = i8peak
                layer0.append(row)
        tensor_rows = []
        for kn in sorted(prof):
            bd, wk = kernel_work(kn, pop, B)
            if bd == "tensor" and wk > 0:
                a = wk / (prof[kn][0] * 1e-3) / 1e12
                tensor_rows.append({"kernel": kn, "bound": "tensor", "achieved": round(a, 2), "peak": peaks["bf16_tflops_sustained"],
                                    "unit": "TFLOP/s", "frac": round(a / peaks["bf16_tflops_sustained"], 4),
                                    "avg_launch_ms": round(prof[kn][0], 5), "work_per_launch": wk})
        # RMSProp inside the step: one entry per launch kind, each with its own bytes and its own duration
        lf, lc = pop.lin_tensor_range()
        roof_rms = []
        for kn, byts, what in (("rms_update_lin", rms_bytes(pop, lf, lc, False), "W_lin / b_lin tensors (side stream)"),
                               ("rms_update_small", rms_bytes(pop, lf, lc, True), "conv + head tensors (compute stream)"),
                               ("rms_update", rms_bytes(pop, 0, 0, True), "whole slab")):
            if kn in prof:
                a = byts / (prof[kn][0] * 1e-3) / 1e9
                roof_rms.append({"kernel": kn, "what": what, "bound": "hbm", "bytes_per_launch": byts,
                                 "avg_launch_ms": round(prof[kn][0], 5), "achieved": round(a, 1), "peak": peaks["hbm_gbs"],
                                 "unit": "GB/s", "frac": round(a / peaks["hbm_gbs"], 4),
                                 "note": "inside the step the slabs are L2-resident; roofline_rmsprop_alone has both regimes"})
        if rms_alone:
            for row in rms_alone:
                row["peak"] = peaks["hbm_gbs"]
                row["frac"] = round(row["achieved"] / peaks["hbm_gbs"], 4)
        plane_bytes = P * (B + cfg.C - 1) * cfg.H * cfg.W
        stack_bytes = P * B * cfg.H * cfg.W * cfg.C
        line = {"metric": METRIC, "value": round(units * K / (ms_total * 1e-3), 1), "unit": UNIT, "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": round(ms_total / K, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
                "config": workload_config(name, world, P), "clocks": sampler.summary(),
                "tournaments_in_timed_region": ntour,
                "active_modules_per_layer": {"max": int(max(bin(int(m)).count("1") for m in pop.masks.ravel())),
                                             "mean": round(float(np.mean([bin(int(m)).count("1") for m in pop.masks.ravel()])), 2)},
                "frozen_tensors": int(pop.frozen.sum()),
                "e2e": {"value": round(units * K / (e2e_ms * 1e-3), 1), "unit": UNIT,
                        "h2d_bytes_per_step": plane_bytes + 3 * P * B * 4, "d2h_bytes_per_step": P * 4,
                        "ms_per_step": round(e2e_ms / K, 4),
                        "host_cpus_bound_to_gpu_numa_node": None if numa_cpus is None else len(numa_cpus),
                        "api": "PipelinedHostTrainer: pn_stage_rollout_host (pinned host buffers holding the B+C-1 distinct "
                               "planes of each path's rollout, copy stream, stacks rebuilt on the device) + "
                               "pn_train_step_staged + pn_rmsprop_apply; losses copied back every step",
                        "stack_form": None if e2e_stack_ms is None else {
                            "value": round(units / (e2e_stack_ms * 1e-3), 1), "ms_per_step": round(e2e_stack_ms, 4),
                            "h2d_bytes_per_step": stack_bytes + 3 * P * B * 4,
                            "api": "pn_stage_inputs_host: full [P,B,H,W,C] stacks cross PCIe"}},
                "gpu_launches": int(launches), "engines": pop.engine_paths(),
                "exchange": wl.trainer.tail.gradsum.path_name(),
                "param_checksum": "%016x" % checksum, "ranks_identical": same,
                "roofline": roof, "roofline_layer0": layer0, "roofline_rmsprop": roof_rms,
                "roofline_rmsprop_alone": rms_alone, "roofline_tensor_kernels": tensor_rows, "kernels": kern,
                "ms_per_step_instrumented": round(ms_instr / K, 4),
                "workspace_mb": round(pop.workspace_bytes / 1e6, 1), "c5_strong": c5, "lstm_variant": lstm_obj}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(name, target_seconds=12.0)
        else:
            line["cpu_baseline"] = None
        emit(line)
    wl.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def measure_rmsprop(pop, torch):
    """pn_rmsprop_apply over the whole slab, timed alone with CUDA events: (a) back to back, slabs L2-resident; (b) with the
    slabs evicted from L2 before every launch.  The launches are queued behind a device-side sleep so that the events
    bracket GPU time only (from Python a launch pair costs more host time than the kernels take)."""
    try:
        byts = rms_bytes(pop, 0, 0, True)
        g = pop.grads
        g.normal_()
        out = []
        for _ in range(5):
            pop.apply_gradients(1e-6)
        torch.cuda.synchronize()
        n = 20
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        torch.cuda._sleep(4_000_000)                       # ~2 ms: the queue fills while the GPU waits
        ev[0].record()
        for _ in range(n):
            pop.apply_gradients(1e-6)
        ev[1].record()
        torch.cuda.synchronize()
        t = ev[0].elapsed_time(ev[1]) / n * 1e-3
        out.append({"kernel": "pn_rmsprop_apply (whole slab: rms_sumsq + rms_update, programmatic dependent launch)",
                    "regime": "slabs L2-resident (back to back)",
                    "bound": "hbm", "bytes_per_launch": byts, "avg_launch_ms": round(t * 1e3, 5),
                    "achieved": round(byts / t / 1e9, 1), "unit": "GB/s"})
        flush = torch.empty(384 * 1024 * 1024, dtype=torch.uint8, device=pop.device)
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        torch.cuda._sleep(4_000_000)
        for a, b in evs:
            flush.fill_(1)                                  # evicts params / rms / grads from the 126 MB L2
            a.record()
            pop.apply_gradients(1e-6)
            b.record()
        torch.cuda.synchronize()
        t = sum(a.elapsed_time(b) for a, b in evs) / n * 1e-3
        out.append({"kernel": "pn_rmsprop_apply (whole slab: rms_sumsq + rms_update, programmatic dependent launch)",
                    "regime": "slabs evicted from L2 before every launch",
                    "bound": "hbm", "bytes_per_launch": byts, "avg_launch_ms": round(t * 1e3, 5),
                    "achieved": round(byts / t / 1e9, 1), "unit": "GB/s"})
        del flush
        # the same two kernels where HBM, not launch latency, is the bound: a 33.5 M-parameter synthetic slab (the largest
        # input of the handle-free entry point: 128 tensors x 128 chunks; 402 MB of slabs, far beyond L2)
        from pathnet_gym_b200.population import rmsprop_apply_noalloc
        n_t, per = 128, 128 * 2048
        nn = n_t * per
        p_, s_, g_ = (torch.randn(nn, device=pop.device), torch.ones(nn, device=pop.device), torch.randn(nn, device=pop.device))
        tensors = [dict(offset=i * per, length=per) for i in range(n_t)]
        for _ in range(3):
            rmsprop_apply_noalloc(p_, s_, g_, tensors, 1e-6)
        torch.cuda.synchronize()
        torch.cuda._sleep(2_000_000)
        ev[0].record()
        for _ in range(10):
            rmsprop_apply_noalloc(p_, s_, g_, tensors, 1e-6)
        ev[1].record()
        torch.cuda.synchronize()
        t = ev[0].elapsed_time(ev[1]) / 10 * 1e-3
        out.append({"kernel": "rms_sumsq + rms_update through pn_rmsprop_apply_noalloc (same kernel bodies)",
                    "regime": "33.5 M-parameter synthetic slab (402 MB of slabs, far beyond L2): HBM is the bound here, launch latency above",
                    "bound": "hbm", "bytes_per_launch": 20.0 * nn, "avg_launch_ms": round(t * 1e3, 5),
                    "achieved": round(20.0 * nn / t / 1e9, 1), "unit": "GB/s"})
        del p_, s_, g_
        return out
    except Exception as e:  # noqa: BLE001
        return [{"error": repr(e)}]


def oracle_cfg(name):
    from oracle import pathnet_oracle as po
    c = CONFIGS[name]
    return po.OracleConfig(L=c["L"], M=10, N=c["N"], H=84, W=84, C=4, A=c["A"])


def cpu_baseline(name, target_seconds=12.0, steps=None, warmup=1):
    """Oracle port timed on this box's host cores: one path at a time, all M modules gated."""
    from oracle import cpu_baseline as cb
    cores = os.cpu_count() or 1
    cfg = oracle_cfg(name)
    if steps is None:
        v, dt, _ = cb.time_cpu_baseline(cfg, BATCH, 4, warmup=warmup, threads=cores)
        steps = int(max(8, min(2000, target_seconds / (dt / 4))))
    v, dt, thr = cb.time_cpu_baseline(cfg, BATCH, steps, warmup=warmup, threads=cores)
    return {"value": round(v, 1), "unit": UNIT, "cores": thr, "kind": "port",
            "sample": "%d single-path updates of batch %d (%s net, all %d modules/layer evaluated and gated, torch-CPU "
                      "fp32, fwd+loss+bwd+clip+RMSProp), %.1f s; TensorFlow-1 reference itself is not installable"
                      % (steps, BATCH, name, cfg.M, dt)}


def tensorflow_reference_available():
    """if a TensorFlow with compat.v1 and the reference sources are present (a driver-side install
    under baseline/_ref), the real graph could be timed; probe only -- nothing in this image provides it."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    try:
        import importlib
        tf = importlib.import_module("tensorflow.compat.v1")
        ok = os.path.exists(os.path.join(ref, "game_ac_network.py"))
        return ("tensorflow %s importable; reference sources %s" % (getattr(tf, "__version__", "?"),
                                                                   "found" if ok else "absent under baseline/_ref"), ok)
    except Exception as e:  # noqa: BLE001
        return ("tensorflow.compat.v1 not importable (%s): oracle port timed instead" % type(e).__name__, False)


def run_reference(args):
    """Reference arm: the reference's own execution shape on the host cores (oracle port; the
    TF-1 reference cannot be installed here -- see DESIGN.md).  Each step = a bounded sample of
    the workload: 4 of the paths, one after another, batch 32."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline as cb
    cores = os.cpu_count() or 1
    K, W, per_step = args.steps, max(args.warmup, 1), 4
    name = args.config
    cfg = oracle_cfg(name)
    probe, have_tf = tensorflow_reference_available()
    if have_tf:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
            import make_tf_golden
            v, dt, thr = make_tf_golden.time_reference_graph(os.path.join(ROOT, "baseline", "_ref"), cfg, BATCH,
                                                             K * per_step, W * per_step)
            kind = "reference"
        except Exception as e:  # noqa: BLE001
            probe += "; running it failed (%r): oracle port timed instead" % (e,)
            have_tf = False
    if not have_tf:
        v, dt, thr = cb.time_cpu_baseline(cfg, BATCH, K * per_step, warmup=W * per_step, threads=cores)
        kind = "port"
    n = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    line = {"impl": "reference", "metric": METRIC, "value": round(v, 1), "unit": UNIT, "n_gpus": n, "steps": K,
            "warmup": W, "ms_per_step": round(dt / K * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (torch CPU)" if kind == "port" else "f32 (TensorFlow CPU)", "data": "synthetic",
            "config": workload_config(name, n, CONFIGS[name]["P"]),
            "cpu_baseline": {"value": round(v, 1), "unit": UNIT, "cores": thr, "kind": kind,
                             "sample": "each step = %d single-path updates of batch %d of the %s workload on the host "
                                       "cores (all M modules evaluated and gated, one path at a time as the reference "
                                       "runs them); the rate, not ms_per_step, is comparable with the GPU arm" % (per_step, BATCH, name)},
            "tensorflow_probe": probe,
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
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c5", action="store_true", help="skip the 512-path strong-scaling object")
    ap.add_argument("--no-lstm", action="store_true", help="skip the LSTM-variant object")
    ap.add_argument("--no-stack-e2e", action="store_true", help="skip the full-stack end-to-end form")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
