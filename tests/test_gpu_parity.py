"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the oracle
and the committed golden fixtures.  Nothing here reads /root/reference.

Tolerances: integer / genotype work is bit-exact.  Floating point follows north_star:
1e-3 relative, applied per tensor as max|x - ref| / max|ref| against the fp64 truth
(REL_TOL below); the fp32 CUDA-core kernels are expected to land near 1e-6.
"""
import json
import os
import zlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import pathnet_oracle as po          # the checker, never the thing measured
from tests import helpers as hp

pytestmark = pytest.mark.gpu

REL_TOL = 1e-3
_report = {}


def _note(key, val):
    _report[key] = float(val)
    out = os.path.join(hp.GOLDEN, "..", "..", "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_report.json"), "w") as f:
            json.dump(_report, f, indent=1, sort_keys=True)
    except OSError:
        pass


@pytest.fixture(scope="module")
def eng_mod(native_lib):
    from pathnet_gym_b200 import population
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    return population


def make_pop(eng_mod, cfg, P, B, params=None, masks=None, **kw):
    pop = eng_mod.PathNetPopulation(hp.to_engine_cfg(cfg), P, B, **kw)
    if params is not None:
        pop.load(pop.params, params)
    if masks is not None:
        pop.set_genotypes(masks)
    return pop


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


# ---------------------------------------------------------------------------
# genotype ops: bit-exact
# ---------------------------------------------------------------------------
def test_genotype_golden_bit_exact(eng_mod):
    g = np.load(os.path.join(hp.GOLDEN, "genotype_golden.npz"))
    for seed, row in zip(g["seeds"], g["rows"]):
        L, M, N, epoch, pid = [int(x) for x in row[:5]]
        m = eng_mod.geopath_init(1, L, M, N, int(seed), epoch, path_id0=pid)
        assert np.array_equal(m.cpu().numpy().view(np.uint32)[0], row[5:5 + L].astype(np.uint32))
        eng_mod.mutate_(m, M, N, int(seed), epoch + 1000, path_id0=pid)
        assert np.array_equal(m.cpu().numpy().view(np.uint32)[0], row[13:13 + L].astype(np.uint32))
        eng_mod.mutate_(m, M, N, int(seed), epoch + 2000, path_id0=pid)
        assert np.array_equal(m.cpu().numpy().view(np.uint32)[0], row[21:21 + L].astype(np.uint32))


@pytest.mark.parametrize("L,M,N,P", [(3, 10, 3, 512), (4, 10, 4, 256), (5, 32, 7, 97)])
def test_genotype_population_matches_oracle(eng_mod, L, M, N, P):
    seed = 0xABCDEF0123
    m = eng_mod.geopath_init(P, L, M, N, seed, 4, path_id0=3)
    ref = np.stack([po.get_geopath_masks(L, M, N, seed, 4, 3 + p) for p in range(P)])
    assert np.array_equal(m.cpu().numpy().view(np.uint32), ref)
    eng_mod.mutate_(m, M, N, seed, 9, path_id0=3)
    ref2 = np.stack([po.mutate_masks(ref[p], L, M, N, seed, 9, 3 + p) for p in range(P)])
    assert np.array_equal(m.cpu().numpy().view(np.uint32), ref2)


@pytest.mark.parametrize("P,B,groups", [(512, 3, 170), (64, 3, 1), (10, 4, 2), (256, 2, 128)])
def test_tournament_matches_oracle(eng_mod, P, B, groups):
    L, M, N, seed = 4, 10, 4, 77
    rng = np.random.default_rng(P)
    geno = np.stack([po.get_geopath_masks(L, M, N, seed, 0, p) for p in range(P)])
    scores = rng.integers(-5, 5, size=P).astype(np.float32)           # many ties
    scores[rng.integers(0, P, size=max(1, P // 16))] = -1000.0        # some paths have not reported
    fixed = np.asarray([0x3, 0x0, 0x210, 0x1], dtype=np.uint32)
    for epoch in (1, 2, 3):
        g_ref, eff_ref, sc_ref, win_ref = po.tournament(geno, scores, fixed, L, M, N, B, groups, seed, epoch)
        gd, sd, fd = dev(geno.view(np.int32)), dev(scores), dev(fixed.view(np.int32))
        eff, win = eng_mod.tournament_(gd, sd, fd, M, N, B, groups, seed, epoch)
        assert np.array_equal(gd.cpu().numpy().view(np.uint32), g_ref)
        assert np.array_equal(eff.cpu().numpy().view(np.uint32), eff_ref)
        assert np.array_equal(sd.cpu().numpy(), sc_ref)
        assert np.array_equal(win.cpu().numpy(), win_ref)
        geno, scores = g_ref, np.where(sc_ref == -1000.0, rng.integers(-5, 5, size=P), sc_ref).astype(np.float32)


# ---------------------------------------------------------------------------
# forward / loss / backward vs golden (fp64 truth)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["float_golden_tiny.npz", "float_golden_tiny4.npz"])
@pytest.mark.parametrize("as_float", [False, True])
def test_train_step_matches_golden(eng_mod, name, as_float):
    g = np.load(os.path.join(hp.GOLDEN, name))
    cfg = hp.cfg_from_golden(g["cfg"])
    P, B = g["frames"].shape[:2]
    params = po.init_params(cfg, 1)
    pop = make_pop(eng_mod, cfg, P, B, params, g["masks"])
    frames = dev(po.frames_to_float(g["frames"])) if as_float else dev(g["frames"])
    # acting forward
    pi, v = pop.forward(frames)
    e_pi, e_v = hp.rel_err(pi.cpu().numpy(), g["pi"]), hp.rel_err(v.cpu().numpy(), g["v"])
    assert e_pi < REL_TOL and e_v < REL_TOL
    loss, pi2, v2 = pop.train_step(frames, dev(g["actions"]), dev(g["td"]), dev(g["r"]), 0.01, want_outputs=True)
    assert torch.equal(pi, pi2) and torch.equal(v, v2)
    e_loss = hp.rel_err(loss.cpu().numpy(), g["loss"])
    assert e_loss < REL_TOL
    got = pop.dump(pop.grads)
    worst = 0.0
    for a, b, t in zip(got, hp.split_flat(cfg, g["grads"]), po.tensor_table(cfg)):
        e = hp.rel_err(a, b)
        worst = max(worst, e)
        assert e < REL_TOL, (t["name"], e)
    _note("%s%s.pi" % (name, ".f32" if as_float else ".u8"), e_pi)
    _note("%s%s.grad_worst" % (name, ".f32" if as_float else ".u8"), worst)
    # update with module 0 of every layer frozen; tiny4's td is scaled so the clip binds
    pop.set_frozen(g["frozen"])
    before_p, before_s = pop.dump(pop.params), pop.dump(pop.ms)
    pop.apply_gradients(float(g["lr"]))
    after_p, after_s = pop.dump(pop.params), pop.dump(pop.ms)
    refp, refs = hp.split_flat(cfg, g["params_after"]), hp.split_flat(cfg, g["ms_after"])
    for i, t in enumerate(po.tensor_table(cfg)):
        if g["frozen"][i]:
            assert np.array_equal(after_p[i], before_p[i]) and np.array_equal(after_s[i], before_s[i]), t["name"]
        else:
            # compare the UPDATE (delta), which is what the kernel computes, and the result
            dref = refp[i] - np.asarray(params[i], dtype=np.float64)
            dgot = after_p[i].astype(np.float64) - before_p[i].astype(np.float64)
            assert np.max(np.abs(dgot - dref)) <= REL_TOL * max(np.max(np.abs(dref)), 1e-12) + 1e-9, t["name"]
            assert hp.rel_err(after_p[i], refp[i]) < REL_TOL
            assert hp.rel_err(after_s[i], refs[i]) < REL_TOL, t["name"]
    pop.close()


def test_c1_matches_golden(eng_mod):
    """BASELINE config C1: L=3, M=10, N=3, 84x84x4, one path, B = LOCAL_T_MAX = 20."""
    g = np.load(os.path.join(hp.GOLDEN, "float_golden_c1.npz"))
    cfg = hp.cfg_from_golden(g["cfg"])
    frames, actions, td, r, masks = hp.synth_inputs(cfg, 1, 20, 1234)
    assert np.uint32(zlib.crc32(frames.tobytes())) == g["frames_crc"]
    params = po.init_params(cfg, 1)
    pop = make_pop(eng_mod, cfg, 1, 20, params, masks)
    # slab layout == oracle tensor table == SURVEY parameter count
    assert sum(t["length"] for t in pop.tensors) == 1697203
    loss, pi, v = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01, want_outputs=True)
    assert hp.rel_err(pi.cpu().numpy(), g["pi"]) < REL_TOL
    assert hp.rel_err(v.cpu().numpy(), g["v"]) < REL_TOL
    assert hp.rel_err(loss.cpu().numpy(), g["loss"]) < REL_TOL
    got = pop.dump(pop.grads)
    assert hp.rel_err(np.concatenate([x.ravel()[::97] for x in got]), g["grads_sample"]) < REL_TOL
    norms = np.asarray([np.sqrt(np.sum(x.astype(np.float64) ** 2)) for x in got])
    assert np.max(np.abs(norms - g["grad_norms"])) < REL_TOL * np.max(g["grad_norms"])
    _note("c1.grad_sample", hp.rel_err(np.concatenate([x.ravel()[::97] for x in got]), g["grads_sample"]))
    pop.set_frozen(g["frozen"])
    pop.apply_gradients(float(g["lr"]))
    assert hp.rel_err(np.concatenate([x.ravel()[::97] for x in pop.dump(pop.params)]), g["params_after_sample"]) < REL_TOL
    assert hp.rel_err(np.concatenate([x.ravel()[::97] for x in pop.dump(pop.ms)]), g["ms_after_sample"]) < REL_TOL
    pop.close()


# ---------------------------------------------------------------------------
# vs the oracle on fresh seeded inputs, edge cases
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("cfg,P,B", [(hp.TINY, 5, 1), (hp.TINY, 2, 33), (hp.TINY4, 3, 20), (hp.C3, 2, 3),
                                     (hp.TC20, 7, 5), (hp.TC20, 9, 1), (hp.C1, 5, 2)])
def test_population_vs_oracle(eng_mod, cfg, P, B):
    params = po.init_params(cfg, 11)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, P, B, 99)
    full = (1 << cfg.M) - 1
    masks = masks.copy()
    masks[0, :] = full                       # every module on (mutation can add modules)
    if P > 1:
        masks[1, 0] = 1 << (cfg.M - 1)       # single module in a layer
    if P > 2:
        masks[2, cfg.L - 1] = 0              # no module at all in the last layer -> h == 0
    if P > 3:
        masks[3, 0] = 0                      # no module in layer 0 -> its output is exactly zero
    if P > 4:
        masks[4, 0] = 0b101010101 & full     # five modules: two tensor-core passes
    pi_r, v_r, loss_r, G = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy")
    pop = make_pop(eng_mod, cfg, P, B, params, masks)
    loss, pi, v = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01, want_outputs=True)
    assert hp.rel_err(pi.cpu().numpy(), pi_r) < REL_TOL
    assert hp.rel_err(v.cpu().numpy(), v_r) < REL_TOL
    assert hp.rel_err(loss.cpu().numpy(), loss_r) < REL_TOL
    for a, b, t in zip(pop.dump(pop.grads), G, po.tensor_table(cfg)):
        assert hp.rel_err(a, b) < REL_TOL, t["name"]
    pop.close()


def test_tensor_core_layer0_matches_cuda_core_path(eng_mod):
    """C2-sized population: the tcgen05 layer-0 kernel (default for u8 frames) against the fp32
    CUDA-core kernels (flags=1) -- activations agree to fp32 rounding, ReLU masks almost always."""
    cfg, P, B = hp.C1, 16, 32
    params = po.init_params(cfg, 1)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, P, B, 77)
    masks = masks.copy()
    masks[1, 0] = (1 << cfg.M) - 1
    fd, ad, tdd, rd = dev(frames), dev(actions), dev(td), dev(r)
    out = []
    for flags in (0, 11):
        pop = make_pop(eng_mod, cfg, P, B, params, masks, flags=flags)
        loss, pi, v = pop.train_step(fd, ad, tdd, rd, 0.01, want_outputs=True)
        out.append((pi.clone(), v.clone(), loss.clone(), pop.grads.clone()))
        pop.close()
    (pi0, v0, l0, g0), (pi1, v1, l1, g1) = out
    _note("tc_vs_simt.pi", hp.rel_err(pi0.cpu().numpy(), pi1.cpu().numpy()))
    _note("tc_vs_simt.v", hp.rel_err(v0.cpu().numpy(), v1.cpu().numpy()))
    _note("tc_vs_simt.loss", hp.rel_err(l0.cpu().numpy(), l1.cpu().numpy()))
    _note("tc_vs_simt.grads", hp.rel_err(g0.cpu().numpy(), g1.cpu().numpy()))
    # forward operands are fp16 hi/lo (2^-22 per product), so the ReLU decisions agree with the fp32 kernels' on these
    # inputs and every tensor is compared on its own
    assert hp.rel_err(pi0.cpu().numpy(), pi1.cpu().numpy()) < 1e-5
    assert hp.rel_err(v0.cpu().numpy(), v1.cpu().numpy()) < 1e-4
    assert hp.rel_err(l0.cpu().numpy(), l1.cpu().numpy()) < 1e-5
    tt = po.tensor_table(cfg)
    pop = make_pop(eng_mod, cfg, 1, 1)
    for t, d in zip(tt, pop.tensors):
        sl = slice(d["offset"], d["offset"] + d["length"])
        assert hp.rel_err(g0[sl].cpu().numpy(), g1[sl].cpu().numpy()) < REL_TOL, t["name"]
    pop.close()


def test_pi_underflow_case(eng_mod):
    cfg = hp.TINY
    params = po.init_params(cfg, 3)
    names = [t["name"] for t in po.tensor_table(cfg)]
    # pi ~ (0.993, 6.7e-3, 2e-22, 2e-22, 9e-27): three entries sit below the 1e-20 clip while the
    # gradients of the others stay O(1e-2), so the comparison is meaningful in fp32
    params[names.index("b_fc2")] = np.asarray([50.0, 45.0, 0, 0, -10.0], dtype=np.float32)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, 1, 3, 5)
    actions[0, 0] = 4                         # an action whose probability is below the clip
    pi_r, v_r, loss_r, G = po.population_step(cfg, params, frames, masks, actions, td, r)
    pop = make_pop(eng_mod, cfg, 1, 3, params, masks)
    loss, pi, v = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01, want_outputs=True)
    assert float(pi.min()) < 1e-20 and torch.isfinite(loss).all()
    assert hp.rel_err(loss.cpu().numpy(), loss_r) < REL_TOL
    for a, b, t in zip(pop.dump(pop.grads), G, po.tensor_table(cfg)):
        assert np.all(np.isfinite(a)) and hp.rel_err(a, b) < REL_TOL, t["name"]
    pop.close()


def test_rmsprop_semantics(eng_mod):
    """inactive tensors decay ms with var unchanged; frozen untouched; clip binds per tensor; momentum slot."""
    cfg = hp.TINY
    params = po.init_params(cfg, 2)
    pop = make_pop(eng_mod, cfg, 1, 1, params, np.zeros((1, cfg.L), np.uint32), momentum_slot=True)
    tt = po.tensor_table(cfg)
    rng = np.random.default_rng(0)
    grads = [rng.normal(size=t["shape"]).astype(np.float32) * (50.0 if i % 3 == 0 else 0.01) for i, t in enumerate(tt)]
    grads[2] = np.zeros_like(grads[2])                    # merely inactive: zero gradient
    frozen = np.zeros(len(tt), np.uint8)
    frozen[4] = frozen[5] = 1
    pop.load(pop.grads, grads)
    pop.set_frozen(frozen)
    ms0 = [np.ones(t["shape"], np.float32) for t in tt]
    mom0 = [np.zeros(t["shape"], np.float32) for t in tt]
    P1, S1, Q1 = po.apply_gradients(params, ms0, mom0, grads, frozen, 7e-4, 0.99, 0.9, 0.1, 40.0)
    pop.apply_gradients(7e-4, 0.99, 0.9, 0.1, 40.0)
    P2, S2, Q2 = po.apply_gradients(P1, S1, Q1, grads, frozen, 5e-4, 0.99, 0.9, 0.1, 40.0)
    pop.apply_gradients(5e-4, 0.99, 0.9, 0.1, 40.0)
    gp, gs, gq = pop.dump(pop.params), pop.dump(pop.ms), pop.dump(pop.mom)
    for i, t in enumerate(tt):
        assert np.allclose(gp[i], P2[i], rtol=1e-5, atol=1e-7), t["name"]
        assert np.allclose(gs[i], S2[i], rtol=1e-5, atol=1e-7), t["name"]
        assert np.allclose(gq[i], Q2[i], rtol=1e-5, atol=1e-8), t["name"]
    assert np.array_equal(gp[2], params[2]) and np.allclose(gs[2], 0.99 * 0.99, rtol=1e-6)
    assert np.array_equal(gp[4], params[4]) and np.array_equal(gs[4], ms0[4])
    # gaps between tensors are never written
    flat = pop.ms.cpu().numpy()
    used = np.zeros(pop.slab_floats, bool)
    for t in pop.tensors:
        used[t["offset"]:t["offset"] + t["length"]] = True
    assert np.all(flat[~used] == 1.0)
    pop.close()


def test_reinit_unfrozen_and_fixed_path(eng_mod):
    cfg = hp.TINY
    params = po.init_params(cfg, 2)
    pop = make_pop(eng_mod, cfg, 1, 1, params, np.zeros((1, cfg.L), np.uint32))
    backup = pop.params.clone()
    pop.params.mul_(3.0)
    fixed = np.zeros((cfg.L, cfg.M))
    fixed[0, 1] = fixed[2, 3] = 1.0
    pop.set_fixed_path(fixed)
    assert np.array_equal(pop.frozen, po.frozen_flags(cfg, fixed))
    pop.reinit_unfrozen(backup)
    got = pop.dump(pop.params)
    for i, t in enumerate(po.tensor_table(cfg)):
        if pop.frozen[i]:
            assert np.array_equal(got[i], params[i] * np.float32(3.0)), t["name"]
        else:
            assert np.array_equal(got[i], params[i]), t["name"]       # heads are re-initialised too
    pop.close()


def test_host_entry_point_equals_device_entry_point(eng_mod):
    cfg = hp.TINY4
    params = po.init_params(cfg, 4)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, 3, 7, 42)
    pop = make_pop(eng_mod, cfg, 3, 8, params, masks)
    loss_d = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01)
    g_d = pop.grads.clone()
    pop.grads.zero_()
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    loss_h = pop.train_step_host(pin(frames), pin(actions), pin(td), pin(r), 0.01)
    assert torch.equal(loss_h, loss_d.cpu()) and torch.equal(pop.grads, g_d)
    pop.close()


def test_pipelined_host_trainer_equals_device_steps(eng_mod):
    """Staged double-buffered host path (copy stream) == the plain device-tensor step, three steps in a row."""
    from pathnet_gym_b200.trainer import PipelinedHostTrainer, PopulationTrainer
    cfg = hp.TC20
    params = po.init_params(cfg, 4)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    batches = [hp.synth_inputs(cfg, 4, 6, 50 + i) for i in range(3)]
    masks = batches[0][4]
    ref = make_pop(eng_mod, cfg, 4, 8, params, masks)
    tr = PopulationTrainer(ref)
    ref_losses = [tr.step(dev(f), dev(a), dev(td), dev(r), 7e-4).cpu() for (f, a, td, r, _) in batches]
    pop = make_pop(eng_mod, cfg, 4, 8, params, masks)
    pipe = PipelinedHostTrainer(pop)
    hb = [(pin(f), pin(a), pin(td), pin(r)) for (f, a, td, r, _) in batches]
    outs = [torch.empty(4).pin_memory() for _ in batches]
    pipe.prime(hb[0])
    for i in range(3):
        pipe.step(hb[i + 1] if i + 1 < 3 else None, 7e-4, loss_out=outs[i])
    torch.cuda.synchronize()
    for a, b in zip(outs, ref_losses):
        assert torch.equal(a, b)
    assert torch.equal(pop.params, ref.params) and torch.equal(pop.ms, ref.ms)
    pop.close(); ref.close()


def test_rollout_planes_staging_equals_full_stacks(eng_mod):
    """pn_stage_rollout_host (the B + C - 1 distinct planes of each rollout cross PCIe, the stacks are rebuilt on the
    device; game_state.py:62-66,75-81) == staging the full [P, B, H, W, C] stacks: bit-identical losses and weights.
    Covers a B_max-sized and a shorter batch, C1 geometry (tensor-core layer 0) and a tiny one."""
    from pathnet_gym_b200.trainer import PipelinedHostTrainer
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    for cfg, P, Bmax, Bs in ((hp.C1, 3, 8, (8, 5, 8)), (hp.TINY, 4, 6, (6, 1, 4))):
        params = po.init_params(cfg, 4)
        rng = np.random.Generator(np.random.PCG64(17))
        masks = hp.synth_inputs(cfg, P, 1, 3)[4]
        runs = []
        for use_planes in (False, True):
            pop = make_pop(eng_mod, cfg, P, Bmax, params, masks)
            pipe = PipelinedHostTrainer(pop)
            rng = np.random.Generator(np.random.PCG64(17))
            hb = []
            for B in Bs:
                planes = rng.integers(0, 256, size=(P, B + cfg.C - 1, cfg.H, cfg.W), dtype=np.uint8)
                stacks = np.stack([planes[:, b:b + cfg.C] for b in range(B)], axis=1).transpose(0, 1, 3, 4, 2)
                a = rng.integers(0, cfg.A, size=(P, B)).astype(np.int32)
                td = rng.normal(size=(P, B)).astype(np.float32)
                r = rng.normal(size=(P, B)).astype(np.float32)
                hb.append((pin(planes if use_planes else stacks), pin(a), pin(td), pin(r)))
            outs = [torch.empty(P).pin_memory() for _ in Bs]
            pipe.prime(hb[0])
            for i in range(len(Bs)):
                pipe.step(hb[i + 1] if i + 1 < len(Bs) else None, 7e-4, loss_out=outs[i])
            torch.cuda.synchronize()
            runs.append((outs, pop.params.clone(), pop.ms.clone()))
            pop.close()
        for a, b in zip(runs[0][0], runs[1][0]):
            assert torch.equal(a, b)
        assert torch.equal(runs[0][1], runs[1][1]) and torch.equal(runs[0][2], runs[1][2])


def test_error_paths(eng_mod):
    from pathnet_gym_b200 import PathNetError
    cfg = hp.TINY
    pop = eng_mod.PathNetPopulation(hp.to_engine_cfg(cfg), 2, 4)
    fr = torch.zeros(2, 4, cfg.H, cfg.W, cfg.C, dtype=torch.uint8, device="cuda")
    with pytest.raises(PathNetError):              # genotypes not set
        pop.forward(fr)
    with pytest.raises(PathNetError):              # bits >= M
        pop.set_genotypes(np.full((2, cfg.L), 1 << cfg.M, np.uint32))
    pop.set_genotypes(np.ones((2, cfg.L), np.uint32))
    with pytest.raises(PathNetError):              # B > B_max
        pop.forward(torch.zeros(2, 5, cfg.H, cfg.W, cfg.C, dtype=torch.uint8, device="cuda"))
    with pytest.raises(ValueError):
        pop.forward(fr[:, :, :-1])
    pop.close()


# ---------------------------------------------------------------------------
# BASELINE.json's full sizes: size-independent properties
# ---------------------------------------------------------------------------
def test_c2_full_size_properties(eng_mod):
    """C2: 64 paths x 32 transitions.  (1) a path's outputs do not depend on its neighbours;
    (2) the population gradient is the sum of the single-path gradients; (3) gathered == dense-gated:
    enabling extra modules whose weights are zero changes nothing; (4) frozen tensors are bit-identical
    after the update and inactive ones only decay ms."""
    cfg, P, B = hp.C1, 64, 32
    params = po.init_params(cfg, 1)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, P, B, 2024)
    pop = make_pop(eng_mod, cfg, P, B, params, masks)
    fd, ad, tdd, rd = dev(frames), dev(actions), dev(td), dev(r)
    loss, pi, v = pop.train_step(fd, ad, tdd, rd, 0.01, want_outputs=True)
    G = pop.grads.clone()
    assert torch.isfinite(G).all() and torch.isfinite(pi).all()
    assert torch.allclose(pi.sum(-1), torch.ones_like(pi.sum(-1)), atol=1e-5)
    # (1)+(2): three single-path populations
    one = make_pop(eng_mod, cfg, 1, B, params)
    acc = torch.zeros_like(G)
    for p in (0, 17, 63):
        one.set_genotypes(masks[p:p + 1])
        l1, pi1, v1 = one.train_step(fd[p:p + 1].contiguous(), ad[p:p + 1].contiguous(), tdd[p:p + 1].contiguous(),
                                     rd[p:p + 1].contiguous(), 0.01, want_outputs=True)
        assert torch.equal(pi1[0], pi[p]) and torch.equal(v1[0], v[p]) and torch.equal(l1[0], loss[p])
        acc += one.grads
    sub = make_pop(eng_mod, cfg, 3, B, params, masks[[0, 17, 63]])
    idx = torch.tensor([0, 17, 63], device="cuda")
    sub.train_step(fd[idx].contiguous(), ad[idx].contiguous(), tdd[idx].contiguous(), rd[idx].contiguous(), 0.01)
    assert hp.rel_err(sub.grads.cpu().numpy(), acc.cpu().numpy()) < 1e-5
    # oracle check of one full-size path (fp32 torch CPU finishes in seconds)
    pi_r, v_r, loss_r, G_r = po.population_step(cfg, params, frames[17:18], masks[17:18], actions[17:18], td[17:18],
                                                r[17:18], dtype="float64")
    assert hp.rel_err(pi[17].cpu().numpy(), pi_r[0]) < REL_TOL and hp.rel_err(loss[17:18].cpu().numpy(), loss_r) < REL_TOL
    one.set_genotypes(masks[17:18])
    one.train_step(fd[17:18].contiguous(), ad[17:18].contiguous(), tdd[17:18].contiguous(), rd[17:18].contiguous(), 0.01)
    for a, b, t in zip(one.dump(one.grads), G_r, po.tensor_table(cfg)):
        assert hp.rel_err(a, b) < REL_TOL, t["name"]
    # (3) zero-weight modules switched on change nothing in pi / v
    used = np.bitwise_or.reduce(masks, axis=0)
    pz = [x.copy() for x in params]
    tt = po.tensor_table(cfg)
    for i, t in enumerate(tt):
        if t["module"] >= 0 and not (int(used[t["layer"]]) >> t["module"]) & 1:
            pz[i][...] = 0
    free = [[j for j in range(cfg.M) if not (int(used[l]) >> j) & 1] for l in range(cfg.L)]
    if all(len(f) > 0 for f in free):
        pop.load(pop.params, pz)
        m2 = masks.copy()
        for l in range(cfg.L):
            m2[:, l] |= np.uint32(1 << free[l][0])
        pop.set_genotypes(m2)
        pi_b, v_b = pop.forward(fd)
        assert torch.equal(pi_b, pi) and torch.equal(v_b, v)
        pop.load(pop.params, params)
        pop.set_genotypes(masks)
    # (4) update: freeze path 0's modules
    fixed = po.gates_from_masks(cfg, masks[0])
    pop.set_fixed_path(fixed)
    pop.grads.copy_(G)
    p0, s0 = pop.params.clone(), pop.ms.clone()
    pop.apply_gradients(7.0016e-4)
    for i, t in enumerate(pop.tensors):
        sl = slice(t["offset"], t["offset"] + t["length"])
        if pop.frozen[i]:
            assert torch.equal(pop.params[sl], p0[sl]) and torch.equal(pop.ms[sl], s0[sl])
        elif float(G[sl].abs().max()) == 0.0:
            assert torch.equal(pop.params[sl], p0[sl]) and torch.allclose(pop.ms[sl], s0[sl] * 0.99, rtol=1e-6)
        else:
            assert not torch.equal(pop.params[sl], p0[sl])
    for q in (pop, one, sub):
        q.close()


# ---------------------------------------------------------------------------
# rollout bookkeeping
# ---------------------------------------------------------------------------
def test_sample_actions_bit_exact(eng_mod):
    rng = np.random.default_rng(11)
    for A in (3, 18, 31):
        z = rng.normal(size=(257, A)) * 3
        pi = np.exp(z - z.max(1, keepdims=True)); pi = (pi / pi.sum(1, keepdims=True)).astype(np.float32)
        pi[5] = 0; pi[5, A - 1] = 1.0                     # deterministic row
        pi[6] = 0; pi[6, 0] = 1.0
        got = eng_mod.sample_actions(dev(pi), 99, 4, id0=1000).cpu().numpy()
        want = po.sample_actions(pi, 99, 4, id0=1000)
        assert np.array_equal(got, want)
        assert got[5] == A - 1 and got[6] == 0


def test_nstep_returns_bit_exact(eng_mod):
    rng = np.random.default_rng(12)
    P, B = 37, 20
    rew = (rng.normal(size=(P, B)) * 2).astype(np.float32)
    val = rng.normal(size=(P, B)).astype(np.float32)
    boot = rng.normal(size=P).astype(np.float32)
    boot[::3] = 0.0                                       # terminal rollouts bootstrap from 0
    lengths = rng.integers(0, B + 1, size=P).astype(np.int32)
    for ln in (None, lengths):
        R, td = eng_mod.nstep_returns(dev(rew), dev(val), dev(boot), None if ln is None else dev(ln))
        Rw, tdw = po.nstep_returns_batch(rew, val, boot, ln)
        assert np.array_equal(R.cpu().numpy(), Rw) and np.array_equal(td.cpu().numpy(), tdw)


def test_padded_transitions_contribute_nothing(eng_mod):
    """action = -1 marks padding: loss and gradients equal those of the batch without the padded transitions."""
    cfg = hp.TINY
    params = po.init_params(cfg, 4)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, 2, 5, 21)
    n_real = [3, 5]
    act_pad = actions.copy()
    act_pad[0, 3:] = -1
    pop = make_pop(eng_mod, cfg, 2, 5, params, masks)
    loss = pop.train_step(dev(frames), dev(act_pad), dev(td), dev(r), 0.01).cpu().numpy()
    got = pop.dump(pop.grads)
    pop.close()
    # oracle: each path on its own real transitions, gradients summed over paths
    G = None
    for p in range(2):
        n = n_real[p]
        _, _, loss_r, Gp = po.population_step(cfg, params, frames[p:p + 1, :n], masks[p:p + 1], actions[p:p + 1, :n],
                                              td[p:p + 1, :n], r[p:p + 1, :n])
        assert hp.rel_err(loss[p:p + 1], loss_r) < REL_TOL
        G = Gp if G is None else [a + b for a, b in zip(G, Gp)]
    for a, b, t in zip(got, G, po.tensor_table(cfg)):
        assert hp.rel_err(a, b) < REL_TOL, t["name"]


def test_preprocess_frames_bit_exact(eng_mod):
    """pn_preprocess_frames against the oracle restatement (itself pinned against cv2): gray, resize, stack push, reset."""
    rng = np.random.default_rng(13)
    for (hin, win), (H, W, C) in [((210, 160), (84, 84, 4)), ((480, 640), (160, 120, 4)), ((50, 40), (21, 33, 3))]:
        n = 3
        obs = rng.integers(0, 256, (n, hin, win, 3), dtype=np.uint8)
        fr = rng.integers(0, 256, (n, H, W, C), dtype=np.uint8)
        reset = np.asarray([0, 1, 0], np.uint8)
        for rs in (None, reset):
            got = eng_mod.preprocess_frames_(dev(obs), dev(fr.copy()), None if rs is None else dev(rs)).cpu().numpy()
            assert np.array_equal(got, po.preprocess_frames(obs, fr, rs))


def test_rmsprop_range_plus_complement_equals_full(eng_mod):
    """pn_rmsprop_apply_range over the linear modules' rows + its complement == one pn_rmsprop_apply (per-tensor updates are
    independent, rmsprop_applier.py:92-106); also checks pn_lin_grads_region against the tensor table."""
    cfg = hp.TINY
    tt = po.tensor_table(cfg)
    rng = np.random.default_rng(3)
    params = po.init_params(cfg, 5)
    grads = [rng.normal(size=t["shape"]).astype(np.float32) * (30.0 if i % 4 == 0 else 0.05) for i, t in enumerate(tt)]
    frozen = np.zeros(len(tt), np.uint8); frozen[3] = 1
    pops = []
    for k in range(2):
        pop = make_pop(eng_mod, cfg, 1, 1, params, np.zeros((1, cfg.L), np.uint32))
        pop.load(pop.grads, grads)
        pop.set_frozen(frozen)
        pops.append(pop)
    first, count = pops[1].lin_tensor_range()
    names = [t["name"] for t in tt]
    assert names[first] == "W_lin[0]" and names[first + count - 1] == "b_lin[%d]" % (cfg.M - 1) and count == 2 * cfg.M
    off, cnt = pops[1].lin_grads_region()
    assert off == pops[1].tensors[first]["offset"] and off + cnt == pops[1].tensors[first + count]["offset"]
    for step in range(2):
        pops[0].apply_gradients(1e-2, 0.99, 0.0, 0.1, 40.0)
        pops[1].apply_gradients_range(first, count, 1e-2, 0.99, 0.0, 0.1, 40.0)
        pops[1].apply_gradients_range(first, count, 1e-2, 0.99, 0.0, 0.1, 40.0, complement=True)
    for slab in ("params", "ms"):
        a, b = pops[0].dump(getattr(pops[0], slab)), pops[1].dump(getattr(pops[1], slab))
        assert all(np.array_equal(x, y) for x, y in zip(a, b)), slab
    for pop in pops:
        pop.close()


def test_peer_sum_small_single_rank(eng_mod):
    """pn_peer_sum_small with world = 1 (the exchange buffer is ordinary device memory): the kernel's publish / flag /
    wait / sum sequence must leave the conv + head regions equal to their input, the linear region untouched, for
    several steps in a row (slot alternation, monotonic flags).  The multi-GPU behaviour is checked by
    tools/multigpu_check.py under torchrun."""
    cfg = hp.TINY
    pop = make_pop(eng_mod, cfg, 1, 1, po.init_params(cfg, 1), np.zeros((1, cfg.L), np.uint32))
    rng = np.random.default_rng(8)
    nbytes = pop.peer_buffer_bytes()
    buf = torch.zeros((nbytes + 3) // 4, dtype=torch.float32, device="cuda")
    off, cnt = pop.lin_grads_region()
    for step in range(5):
        g0 = torch.from_numpy(rng.normal(size=pop.slab_floats).astype(np.float32)).cuda()
        pop.grads.copy_(g0)
        pop.peer_sum_small([buf.data_ptr()], 0, 1, step)
        torch.cuda.synchronize()
        assert torch.equal(pop.grads, g0)
        slot = buf[256 + (step & 1) * (pop.slab_floats - cnt):][:pop.slab_floats - cnt]
        assert torch.equal(slot[:off], g0[:off]) and torch.equal(slot[off:], g0[off + cnt:])
    flags = buf[:64].view(torch.int32)
    assert int(flags[0]) == 5
    pop.close()


def test_peer_sum_parts_single_rank(eng_mod):
    """pn_peer_sum_part with world = 1: part 1 (everything but the first conv layer's tensors, exchanged early after
    pn_stream_wait_early_grads) and part 2 (the first conv layer) publish disjoint pieces of the same slots with their own
    flags; together they must leave the slab unchanged and the slots equal to the small regions, for several steps."""
    cfg = hp.TINY
    pop = make_pop(eng_mod, cfg, 1, 1, po.init_params(cfg, 1), np.zeros((1, cfg.L), np.uint32))
    rng = np.random.default_rng(9)
    buf = torch.zeros((pop.peer_buffer_bytes() + 3) // 4, dtype=torch.float32, device="cuda")
    off, cnt = pop.lin_grads_region()
    side = torch.cuda.Stream()
    late = pop.stream_wait_early_grads(side)
    assert 0 < late <= off
    for step in range(4):
        g0 = torch.from_numpy(rng.normal(size=pop.slab_floats).astype(np.float32)).cuda()
        pop.grads.copy_(g0)
        pop.peer_sum_part([buf.data_ptr()], 0, 1, step, 1)
        pop.peer_sum_part([buf.data_ptr()], 0, 1, step, 2)
        torch.cuda.synchronize()
        assert torch.equal(pop.grads, g0)
        slot = buf[256 + (step & 1) * (pop.slab_floats - cnt):][:pop.slab_floats - cnt]
        assert torch.equal(slot[:off], g0[:off]) and torch.equal(slot[off:], g0[off + cnt:])
    flags = buf[:256].view(torch.int32)
    assert int(flags[136]) == 4 and int(flags[160]) == 4 and int(flags[0]) == 0
    from pathnet_gym_b200 import PathNetError
    with pytest.raises(PathNetError):
        pop.peer_sum_part([buf.data_ptr()], 0, 1, 4, 3)
    pop.close()


def test_tensor_core_matches_cuda_core_path_l4(eng_mod):
    """L = 4 (third conv layer 3x3 / stride 1 on the tensor cores too): all tensor-core kernels against all CUDA-core
    kernels at a size where the raster tiles straddle transitions (81 rows per transition, 128-row tiles) and the weight
    gradient is split into chunks."""
    cfg, P, B = hp.C3, 6, 32
    params = po.init_params(cfg, 2)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, P, B, 78)
    fd, ad, tdd, rd = dev(frames), dev(actions), dev(td), dev(r)
    out = []
    for flags in (0, 11):
        pop = make_pop(eng_mod, cfg, P, B, params, masks, flags=flags)
        loss, pi, v = pop.train_step(fd, ad, tdd, rd, 0.01, want_outputs=True)
        out.append((pi.cpu().numpy(), v.cpu().numpy(), loss.cpu().numpy(), pop.dump(pop.grads)))
        pop.close()
    (pi0, v0, l0, g0), (pi1, v1, l1, g1) = out
    assert hp.rel_err(pi0, pi1) < 1e-4 and hp.rel_err(v0, v1) < 1e-4 and hp.rel_err(l0, l1) < 1e-4
    worst = max((hp.rel_err(a, b), t["name"]) for a, b, t in zip(g0, g1, po.tensor_table(cfg)))
    _note("tc_vs_simt_l4.worst_grad", worst[0])
    # per tensor (forward operands are fp16 hi/lo, so the ReLU decisions agree with the fp32 kernels' on these inputs)
    assert worst[0] < REL_TOL, worst


# ---------------------------------------------------------------------------
# population scale, per tensor, against the fp64 oracle -- with the fp32 oracle beside it
# ---------------------------------------------------------------------------
FRAGILE = 1e-5      # a ReLU decision may differ from the fp64 oracle's only where |pre| < FRAGILE * rms(pre of that module)


def _per_tensor_vs_fp64(eng_mod, cfg, P, B, seed, tag, masks=None, frozen_path=None, flags=0):
    """One population step against the fp64 oracle, every output and every gradient tensor on its own.

    A ReLU unit whose pre-activation is within fp32 rounding of zero can fall on either side in ANY fp32-level
    implementation (the fp32 evaluation of the oracle itself flips a few per million against fp64), and one flipped
    unit moves a small tensor's gradient by far more than 1e-3.  So the comparison has two parts:
      1. the CUDA path's ReLU decisions (pn_debug_read_relu_masks) equal the fp64 oracle's except on units with
         |pre| < FRAGILE * rms(pre) -- counted and recorded;
      2. against the fp64 oracle evaluated WITH the CUDA path's decisions, every tensor is within REL_TOL.
    The unconditioned comparison and the fp32 oracle's own distance from fp64 are recorded beside it."""
    params = po.init_params(cfg, seed)
    frames, actions, td, r, m0 = hp.synth_inputs(cfg, P, B, seed + 100)
    masks = m0 if masks is None else masks
    pre = [{} for _ in range(P)]
    pi64, v64, l64, G64 = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", dtype="float64",
                                             pre_out=pre)
    pi32, v32, l32, G32 = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", dtype="float32")
    pop = make_pop(eng_mod, cfg, P, B, params, masks, flags=flags)
    if frozen_path is not None:
        pop.set_fixed_path(frozen_path)
    loss, pi, v = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01, want_outputs=True)
    G = pop.dump(pop.grads)
    dec = [pop.relu_decisions(l, B) for l in range(cfg.L)]
    override, flips, total = [], [], 0
    for p in range(P):
        d = {}
        for (l, j), pr in pre[p].items():
            mk = dec[l][p, :, j].reshape(pr.shape)
            d[(l, j)] = mk
            diff = mk != (pr > 0)
            total += pr.size
            if diff.any():
                flips += [float(x) for x in np.abs(pr[diff]) / np.sqrt(np.mean(pr ** 2))]
        override.append(d)
    piF, vF, lF, GF = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", dtype="float64",
                                         relu_override=override)
    names = [t["name"] for t in po.tensor_table(cfg)]

    def errs(pi_, v_, l_, G_, ref):
        e = {"pi": hp.rel_err(pi_, ref[0]), "v": hp.rel_err(v_, ref[1]), "loss": hp.rel_err(l_, ref[2])}
        e.update({"grad/" + n: hp.rel_err(a, b) for n, a, b in zip(names, G_, ref[3])})
        return e

    out = (pi.cpu().numpy(), v.cpu().numpy(), loss.cpu().numpy(), G)
    same = errs(*out, (piF, vF, lF, GF))              # same ReLU decisions: arithmetic only
    plain = errs(*out, (pi64, v64, l64, G64))         # unconditioned
    f32 = errs(pi32, v32, l32, G32, (pi64, v64, l64, G64))
    pop.close()
    for key, e in (("cuda_same_decisions", same), ("cuda_unconditioned", plain), ("oracle_fp32", f32)):
        w = max(e, key=e.get)
        _report["%s.%s.worst" % (tag, key)] = [w, e[w]]
    _report[tag + ".relu_units"] = total
    _report[tag + ".relu_decisions_differing"] = len(flips)
    _report[tag + ".relu_differing_max_abs_pre_over_rms"] = max(flips) if flips else 0.0
    _note(tag + ".cuda_same_decisions.pi", same["pi"])
    return same, plain, f32, flips, total


def _check_population(same, flips, total):
    assert all(f < FRAGILE for f in flips), ("a ReLU decision differs on a well-conditioned unit", max(flips))
    assert len(flips) <= max(8, 1e-5 * total), (len(flips), total)
    bad = {k: x for k, x in same.items() if not x < REL_TOL}
    assert not bad, bad


@pytest.mark.parametrize("cfgname,P,B", [("C1", 16, 32), ("C3", 6, 32)])
def test_population_scale_per_tensor_vs_fp64(eng_mod, cfgname, P, B):
    """north_star's 1e-3, per tensor, at population scale: 16 paths x 32 of the C2 net, 6 x 32 of the C3 net."""
    cfg = getattr(hp, cfgname)
    same, plain, f32, flips, total = _per_tensor_vs_fp64(eng_mod, cfg, P, B, 3, "pop_%s_%dx%d" % (cfgname, P, B))
    _check_population(same, flips, total)


def test_c4_task_transfer_real_shape(eng_mod):
    """BASELINE config C4 as stated: the L=4 / M=10 / N=4 net with the 3-action head; task 1 runs tournaments (so
    mutations have already added and dropped modules), its winner's path is frozen (doom_pathnet.py:273-283), every
    other tensor is re-initialised (:289-293), and task 2 starts fresh genotypes OR'ed with the frozen path (:213-221)
    -- 4..8 active modules per layer.  Checks on that population: per-tensor parity of one step against the fp64
    oracle, frozen tensors bit-identical after the update, non-frozen ones equal to the oracle's clipped RMSProp."""
    from pathnet_gym_b200.controller import PopulationController
    cfg, P, B = hp.C4, 12, 32
    ecfg = hp.to_engine_cfg(cfg)
    pop = eng_mod.PathNetPopulation(ecfg, P, B)
    pop.init_params(5)
    ctl = PopulationController(pop, P, Bcand=3, seed=21)
    ctl.start_task()
    for rnd in range(6):                                        # task 1: a few rounds of tournaments + mutation
        ctl.report_scores(torch.arange(P, dtype=torch.float32).roll(rnd))
        ctl.tournament_step()
    fixed = ctl.end_task()                                      # freeze the winner's path, re-initialise the rest
    ctl.start_task()                                            # task 2 population: genotype | fixed path
    masks = pop.masks.copy()
    fixed_bits = po.masks_from_gates(fixed)
    assert np.all((masks & fixed_bits[None, :]) == fixed_bits[None, :])
    pc = np.array([[bin(int(m)).count("1") for m in row] for row in masks])
    assert pc.max() > 4 and pc.min() >= 4                       # the shapes the >4-module paths of every kernel see
    _report["c4.active_modules_per_layer.max"] = int(pc.max())
    params = pop.dump(pop.params)
    frames, actions, td, r, _ = hp.synth_inputs(cfg, P, B, 41)
    pre = [{} for _ in range(P)]
    pi64, v64, l64, G64 = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", pre_out=pre)
    loss, pi, v = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01, want_outputs=True)
    G = pop.dump(pop.grads)
    dec = [pop.relu_decisions(l, B) for l in range(cfg.L)]
    override, flips, total = [], [], 0
    for p in range(P):
        d = {}
        for (l, j), pr in pre[p].items():
            mk = dec[l][p, :, j].reshape(pr.shape)
            d[(l, j)] = mk
            diff = mk != (pr > 0)
            total += pr.size
            if diff.any():
                flips += [float(x) for x in np.abs(pr[diff]) / np.sqrt(np.mean(pr ** 2))]
        override.append(d)
    piF, vF, lF, GF = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", relu_override=override)
    same = {"pi": hp.rel_err(pi.cpu().numpy(), piF), "v": hp.rel_err(v.cpu().numpy(), vF),
            "loss": hp.rel_err(loss.cpu().numpy(), lF)}
    tt = po.tensor_table(cfg)
    same.update({"grad/" + t["name"]: hp.rel_err(a, b) for t, a, b in zip(tt, G, GF)})
    w = max(same, key=same.get)
    _report["c4.cuda_same_decisions.worst"] = [w, same[w]]
    _report["c4.relu_decisions_differing"] = len(flips)
    _note("c4.relu_units", total)
    _check_population(same, flips, total)
    # the masked update: frozen tensors untouched (bit-identical), the rest = clip_by_norm + ApplyRMSProp of the oracle
    frozen = po.frozen_flags(cfg, fixed)
    assert np.array_equal(pop.frozen, frozen) and frozen.sum() > 0
    p_before, ms_before = pop.params.clone(), pop.ms.clone()
    ms0 = pop.dump(pop.ms)
    pop.apply_gradients(7.0016e-4)
    want_p, want_ms, _ = po.apply_gradients(params, ms0, [np.zeros_like(x) for x in params], G, frozen, 7.0016e-4)
    got_p, got_ms = pop.dump(pop.params), pop.dump(pop.ms)
    for i, t in enumerate(pop.tensors):
        sl = slice(t["offset"], t["offset"] + t["length"])
        if frozen[i]:
            assert torch.equal(pop.params[sl], p_before[sl]) and torch.equal(pop.ms[sl], ms_before[sl]), tt[i]["name"]
        else:
            upd = np.abs(np.asarray(want_p[i], np.float64) - params[i]).max()
            assert np.abs(got_p[i] - want_p[i]).max() <= REL_TOL * max(upd, 1e-30), tt[i]["name"]
            assert hp.rel_err(got_ms[i], want_ms[i]) < 1e-5, tt[i]["name"]
    pop.close()


# ---------------------------------------------------------------------------
# the LSTM variant (GameACPathNetLSTMNetwork, game_ac_network.py:303-521)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("cfgname,P,B,zero_state", [("TC20_LSTM", 5, 7, False), ("TC20_LSTM", 3, 1, True),
                                                    ("C1_LSTM", 64, 20, False), ("C3_LSTM", 6, 20, False)])
def test_lstm_variant_vs_oracle(eng_mod, cfgname, P, B, zero_state):
    """Population-batched BasicLSTMCell(256) + BPTT against the fp64 oracle: pi, v, loss, every gradient tensor
    (W_lstm / b_lstm included), the state after the last step, and one clipped RMSProp update -- up to the reference's
    LOCAL_T_MAX = 20 steps on 64 paths of the C2 net."""
    cfg = getattr(hp, cfgname)
    params = po.init_params(cfg, 6)
    rng = np.random.default_rng(8)
    params[-1] = (rng.normal(size=params[-1].shape) * 0.1).astype(np.float32)          # (b_lstm is all zero at init)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, P, B, 60)
    st_in = None if zero_state else (rng.normal(size=(P, 2, cfg.hidden)) * 0.5).astype(np.float32)
    states = None if zero_state else [(st_in[p, 0], st_in[p, 1]) for p in range(P)]
    pre = [{} for _ in range(P)]
    pi64, v64, l64, G64 = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", pre_out=pre,
                                             lstm_states=states)
    pop = make_pop(eng_mod, cfg, P, B, params, masks)
    assert len(pop.tensors) == len(po.tensor_table(cfg)) and pop.tensors[-2]["kind"] == 8
    st_out = torch.zeros(P, 2, cfg.hidden, device="cuda")
    pop.bind_lstm_state(None if zero_state else dev(st_in), st_out)
    loss, pi, v = pop.train_step(dev(frames), dev(actions), dev(td), dev(r), 0.01, want_outputs=True)
    G = pop.dump(pop.grads)
    # evaluate the oracle with the CUDA path's ReLU decisions (see _per_tensor_vs_fp64)
    dec = [pop.relu_decisions(l, B) for l in range(cfg.L)]
    override, flips, total = [], [], 0
    for p in range(P):
        d = {}
        for (l, j), pr in pre[p].items():
            mk = dec[l][p, :, j].reshape(pr.shape)
            d[(l, j)] = mk
            diff = mk != (pr > 0)
            total += pr.size
            if diff.any():
                flips += [float(x) for x in np.abs(pr[diff]) / np.sqrt(np.mean(pr ** 2))]
        override.append(d)
    piF, vF, lF, GF = po.population_step(cfg, params, frames, masks, actions, td, r, impl="numpy", relu_override=override,
                                         lstm_states=states)
    names = [t["name"] for t in po.tensor_table(cfg)]
    same = {"pi": hp.rel_err(pi.cpu().numpy(), piF), "v": hp.rel_err(v.cpu().numpy(), vF),
            "loss": hp.rel_err(loss.cpu().numpy(), lF)}
    same.update({"grad/" + n: hp.rel_err(a, b) for n, a, b in zip(names, G, GF)})
    w = max(same, key=same.get)
    _report["lstm_%s_%dx%d.worst" % (cfgname, P, B)] = [w, same[w]]
    _note("lstm_%s_%dx%d.relu_decisions_differing" % (cfgname, P, B), len(flips))
    _check_population(same, flips, total)
    # final state of path 0 (one more oracle pass that returns it)
    so = {}
    po.forward_backward_numpy(cfg, params, po.frames_to_float(frames[0]), po.gates_from_masks(cfg, masks[0]), actions[0], td[0],
                              r[0], want_grads=False, lstm_state=None if zero_state else states[0], state_out=so)
    got = st_out[0].cpu().numpy()
    assert hp.rel_err(got[0], so[0]) < REL_TOL and hp.rel_err(got[1], so[1]) < REL_TOL
    # the update: W_lstm / b_lstm are ordinary tensors of the slab (clipped per tensor like every other)
    ms0 = pop.dump(pop.ms)
    pop.apply_gradients(7.0016e-4)
    fz = np.zeros(len(names), np.uint8)
    want_p, want_ms, _ = po.apply_gradients(params, ms0, [np.zeros_like(x) for x in params], G, fz, 7.0016e-4)
    got_p = pop.dump(pop.params)
    for i in (len(names) - 2, len(names) - 1):
        upd = np.abs(np.asarray(want_p[i], np.float64) - params[i]).max()
        assert np.abs(got_p[i] - want_p[i]).max() <= REL_TOL * max(upd, 1e-30), names[i]
    pop.close()
