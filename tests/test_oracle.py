"""CPU tests of the oracle: pinned against the reference-generated golden vectors
(genotype ops), known-answer vectors (Philox), and its own two restatements."""
import os

import numpy as np
import pytest

from oracle import pathnet_oracle as po
from tests import helpers as hp


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    assert po.philox4x32_10((0, 0, 0, 0), (0, 0)) == (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)
    assert po.philox4x32_10((0xffffffff,) * 4, (0xffffffff,) * 2) == (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)
    assert po.philox4x32_10((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == \
        (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)


def test_genotype_ops_match_reference_golden():
    """genotype_golden.npz was written by the real pathnet.get_geopath / pathnet.mutation
    (/root/reference/pathnet.py:50-63,78-87) driven by the Philox stream."""
    g = np.load(os.path.join(hp.GOLDEN, "genotype_golden.npz"))
    for seed, row in zip(g["seeds"], g["rows"]):
        L, M, N, epoch, pid = [int(x) for x in row[:5]]
        m0 = po.get_geopath_masks(L, M, N, int(seed), epoch, pid)
        assert np.array_equal(m0, row[5:5 + L].astype(np.uint32))
        m1 = po.mutate_masks(m0, L, M, N, int(seed), epoch + 1000, pid)
        assert np.array_equal(m1, row[13:13 + L].astype(np.uint32))
        m2 = po.mutate_masks(m1, L, M, N, int(seed), epoch + 2000, pid)
        assert np.array_equal(m2, row[21:21 + L].astype(np.uint32))


def test_get_geopath_has_N_distinct():
    for p in range(50):
        m = po.get_geopath_masks(4, 10, 4, 99, 3, p)
        assert all(bin(int(x)).count("1") == 4 for x in m)


def test_mutation_rate_is_plausible():
    # P(active bit mutates) = 2/(L*N) (pathnet.py:54-55)
    L, M, N, hits, tot = 3, 10, 3, 0, 0
    for p in range(400):
        m0 = po.get_geopath_masks(L, M, N, 5, 0, p)
        m1 = po.mutate_masks(m0, L, M, N, 5, 1, p)
        hits += int(np.any(m0 != m1))
        tot += 1
    assert 0.5 < hits / tot < 0.98


def test_tournament_semantics():
    L, M, N, P, B = 3, 10, 3, 12, 3
    geno = np.stack([po.get_geopath_masks(L, M, N, 1, 0, p) for p in range(P)])
    scores = np.arange(P, dtype=np.float32)
    fixed = np.asarray([1, 2, 4], dtype=np.uint32)
    g2, eff, sc2, win = po.tournament(geno, scores, fixed, L, M, N, B, 4, seed=11, epoch=5)
    perm = po.tournament_permutation(P, 11, 5)
    assert sorted(perm.tolist()) == list(range(P))
    for grp in range(4):
        mem = perm[grp * B:(grp + 1) * B]
        assert win[grp] == max(mem)                       # scores == ids
        assert np.array_equal(g2[win[grp]], geno[win[grp]])
        assert np.all(sc2[mem] == -1000.0)                # winner reset too (doom_pathnet.py:267)
    assert np.array_equal(eff, g2 | fixed[None, :])
    # a group with an unset score is left alone
    scores2 = scores.copy()
    scores2[perm[0]] = -1000.0
    g3, _, sc3, win3 = po.tournament(geno, scores2, fixed, L, M, N, B, 1, seed=11, epoch=5)
    assert win3[0] == -1 and np.array_equal(g3, geno) and np.array_equal(sc3, scores2)
    # ties: first arg-max in group order
    g4, _, _, win4 = po.tournament(geno, np.zeros(P, np.float32), fixed, L, M, N, B, 1, seed=11, epoch=5)
    assert win4[0] == perm[0]


@pytest.mark.parametrize("cfg", [hp.TINY, hp.TINY4])
def test_two_restatements_agree(cfg):
    """torch autograd on the literal all-M gated graph == NumPy closed form on gathered modules."""
    params = po.init_params(cfg, 3)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, 2, 4, 5)
    for p in range(2):
        gates = po.gates_from_masks(cfg, masks[p])
        xf = po.frames_to_float(frames[p])
        pi, v, ls, g = po.path_grads_torch(cfg, params, xf, gates, actions[p], td[p], r[p])
        pi2, v2, ls2, g2 = po.forward_backward_numpy(cfg, params, xf, gates, actions[p], td[p], r[p])
        assert np.allclose(pi, pi2, atol=1e-12) and np.allclose(v, v2, atol=1e-12) and abs(ls - ls2) < 1e-9
        for a, b in zip(g, g2):
            assert np.allclose(a, b, atol=1e-10)


def test_pi_underflow_clip_gradient():
    """pi < 1e-20: the log branch passes no gradient, the explicit pi factor still does."""
    cfg = hp.TINY
    params = po.init_params(cfg, 3)
    tt = [t["name"] for t in po.tensor_table(cfg)]
    params[tt.index("b_fc2")] = np.asarray([200.0, 0, 0, 0, -200.0], dtype=np.float32)
    frames, actions, td, r, masks = hp.synth_inputs(cfg, 1, 3, 5)
    gates = po.gates_from_masks(cfg, masks[0])
    xf = po.frames_to_float(frames[0])
    pi, v, ls, g = po.path_grads_torch(cfg, params, xf, gates, actions[0], td[0], r[0])
    pi2, v2, ls2, g2 = po.forward_backward_numpy(cfg, params, xf, gates, actions[0], td[0], r[0])
    assert pi.min() < 1e-20
    assert np.isfinite(ls) and abs(ls - ls2) < 1e-9
    for a, b in zip(g, g2):
        assert np.all(np.isfinite(a)) and np.allclose(a, b, atol=1e-10)


def test_float_goldens_reproduce():
    for name in ("float_golden_tiny.npz", "float_golden_tiny4.npz"):
        g = np.load(os.path.join(hp.GOLDEN, name))
        cfg = hp.cfg_from_golden(g["cfg"])
        params = po.init_params(cfg, 1)
        pi, v, loss, G = po.population_step(cfg, params, g["frames"], g["masks"], g["actions"], g["td"], g["r"])
        assert np.allclose(pi, g["pi"], atol=1e-12) and np.allclose(loss, g["loss"], atol=1e-9)
        assert np.allclose(np.concatenate([x.ravel() for x in G]), g["grads"], atol=1e-10)
        # the fp32 restatement stays within 1e-3 (relative to each tensor's max) of the fp64 truth
        _, _, _, G32 = po.population_step(cfg, params, g["frames"], g["masks"], g["actions"], g["td"], g["r"],
                                          dtype="float32")
        for a, b in zip(G32, hp.split_flat(cfg, g["grads"])):
            assert hp.rel_err(a, b) < 1e-4


def test_clip_and_rmsprop_semantics():
    g = np.asarray([30.0, 40.0], dtype=np.float32)                 # norm 50 > 40 -> scaled to norm 40
    gc = po.clip_by_norm(g, 40.0)
    assert np.allclose(gc, g * 0.8, rtol=1e-6)
    assert np.array_equal(po.clip_by_norm(np.zeros(4, np.float32), 40.0), np.zeros(4, np.float32))
    small = np.asarray([0.3, -0.4], dtype=np.float32)
    assert np.allclose(po.clip_by_norm(small, 40.0), small, rtol=1e-6)
    var, ms, mom = np.ones(3, np.float32), np.ones(3, np.float32), np.zeros(3, np.float32)
    v2, ms2, _ = po.apply_rms_prop(var, ms, mom, 1e-3, 0.99, 0.0, 0.1, np.zeros(3, np.float32))
    # merely-inactive (zero-gradient) tensors still decay ms
    assert np.allclose(ms2, 0.99) and np.array_equal(v2, var)
    P2, S2, _ = po.apply_gradients([var], [ms], [mom], [np.ones(3, np.float32)], [1], 1e-3)
    assert np.array_equal(P2[0], var) and np.array_equal(S2[0], ms)    # frozen: untouched


def test_lr_and_returns():
    lr0 = po.log_uniform(1e-4, 1e-2, 0.4226)
    assert abs(lr0 - 7.0016e-4) < 1e-7
    assert po.anneal_learning_rate(lr0, po.MAX_TIME_STEP * 2) == 0.0
    R, td = po.nstep_returns([1, 0, 5], [0.5, 0.5, 0.5], 2.0)
    assert abs(R[2] - (1 + 0.99 * 2.0)) < 1e-12 and abs(R[0] - (1 + 0.99 * (0 + 0.99 * R[2]))) < 1e-12
    assert np.allclose(td, R - 0.5)


def test_param_counts_match_survey():
    n = lambda cfg: sum(int(np.prod(t["shape"])) for t in po.tensor_table(cfg))
    assert n(hp.C1) == 1697203 and n(hp.C3) == 1047683
    assert hp.C1.last_lin == 648 and hp.C3.last_lin == 392


def test_choice_restatement_matches_numpy_randomstate():
    """choice_from_uniform is numpy's legacy RandomState.choice with the uniform made explicit: feed it the uniform
    the RandomState is about to draw and compare with what RandomState.choice returns (a3c_training_thread.py:89-90)."""
    rng = np.random.default_rng(5)
    for case in range(400):
        A = int(rng.integers(2, 19))
        z = rng.normal(size=A) * rng.choice([0.1, 1.0, 8.0])
        p = np.exp(z - z.max()); p = (p / p.sum()).astype(np.float32)
        rs = np.random.RandomState(1000 + case)
        st = rs.get_state()
        u = rs.random_sample()
        rs.set_state(st)
        want = rs.choice(range(A), p=p.astype(np.float64) / p.astype(np.float64).sum())
        got = po.choice_from_uniform(p, u)
        assert got == want, (case, A, u)


def test_nstep_returns_batch_padding():
    rng = np.random.default_rng(6)
    P, B = 5, 7
    rew = rng.normal(size=(P, B)).astype(np.float32) * 2
    val = rng.normal(size=(P, B)).astype(np.float32)
    boot = rng.normal(size=P).astype(np.float32)
    lengths = np.asarray([7, 3, 1, 0, 5], np.int32)
    R, td = po.nstep_returns_batch(rew, val, boot, lengths)
    for p in range(P):
        n = int(lengths[p])
        Rs, tds = po.nstep_returns(rew[p, :n], val[p, :n], boot[p])
        assert np.array_equal(R[p, :n], Rs.astype(np.float32)) and np.array_equal(td[p, :n], tds.astype(np.float32))
        assert not R[p, n:].any() and not td[p, n:].any()


def test_frame_preprocessing_restatement_matches_cv2():
    """bgr2gray_u8 / resize_linear_u8 are OpenCV's 8-bit fixed-point paths restated (game_state.py:41-45): compare
    with cv2 itself, bit for bit, on the geometries the reference sees."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    for (h, w), (dw, dh) in [((210, 160), (84, 84)), ((210, 160), (120, 160)), ((480, 640), (120, 160)), ((84, 84), (84, 84)),
                             ((100, 37), (84, 84))]:
        for k in range(3):
            img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
            if k == 1:
                img[:] = rng.integers(0, 256, (1, 1, 3))                 # flat colour
            g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
            assert np.array_equal(po.bgr2gray_u8(img), g)
            assert np.array_equal(po.resize_linear_u8(g, dw, dh), cv2.resize(g, (dw, dh)))
    obs = rng.integers(0, 256, (2, 50, 40, 3), dtype=np.uint8)
    fr = rng.integers(0, 256, (2, 21, 21, 4), dtype=np.uint8)
    out = po.preprocess_frames(obs, fr, reset=[0, 1])
    x0 = cv2.resize(cv2.cvtColor(obs[0], cv2.COLOR_BGR2GRAY), (21, 21))
    assert np.array_equal(out[0, :, :, :3], fr[0, :, :, 1:]) and np.array_equal(out[0, :, :, 3], x0)
    x1 = cv2.resize(cv2.cvtColor(obs[1], cv2.COLOR_BGR2GRAY), (21, 21))
    assert all(np.array_equal(out[1, :, :, c], x1) for c in range(4))
