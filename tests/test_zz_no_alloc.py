"""GPU test of pn_rmsprop_apply_noalloc, the update entry point for hosts that cannot allocate.

Some callers of this library have no way to obtain device memory at the moment they update the
weights: the GPU was handed to them fully reserved by a serving process, or they cannot afford the
device-wide synchronisation that cudaMalloc / cudaFree imply, or their host forbids plugins to own
device memory.  For them the library carries its scratch (one partial sum of squares per 2048-float
chunk) in a built-in per-device array instead of a handle's workspace.  This file checks, through
the C ABI, that the variant gives the same parameters and rms / momentum slots as the regular
pn_rmsprop_apply on two calls in a row (the second call finds the first call's leftovers in the
shared scratch), and that inputs beyond the built-in array's size are refused with a clear message.

Collected last on purpose (file name), so a run that stops at the first failure has already run
every other test.  Both entry points run the same kernel bodies (csrc/pn_rms_core.cuh), so their
results must be bit-identical; a mistake the two could share is caught by the second comparison,
against the oracle's clip_by_norm + ApplyRMSProp, within the suite's 1e-3 relative (REL_TOL) on the
update itself (measured ~1e-6).  A third test feeds the handle-free entry point a tensor table the
handle form never produces: offsets that are not multiples of 4 floats (scalar path), lengths that
are not multiples of 4 or of the 2048-float chunk, a tensor spanning many chunks, a non-default stream.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import pathnet_oracle as po
from tests import helpers as hp

pytestmark = pytest.mark.gpu

REL_TOL = 1e-3


def _grads(cfg, seed):
    rng = np.random.default_rng(seed)
    tt = po.tensor_table(cfg)
    g = [rng.normal(size=t["shape"]).astype(np.float32) * (30.0 if i % 4 == 0 else 0.02) for i, t in enumerate(tt)]
    g[3] = np.zeros_like(g[3])          # an unused module: zero gradient, rms still decays
    return g


@pytest.mark.parametrize("momentum", [0.0, 0.9])
def test_noalloc_matches_regular_entry_point_twice(native_lib, momentum):
    from pathnet_gym_b200 import population as eng
    assert torch.cuda.is_available()
    cfg = hp.TINY4
    params = po.init_params(cfg, 9)
    pop = eng.PathNetPopulation(hp.to_engine_cfg(cfg), 1, 1, momentum_slot=momentum != 0.0)
    pop.load(pop.params, params)
    frozen = np.zeros(len(pop.tensors), np.uint8)
    frozen[6] = frozen[7] = 1
    pop.set_frozen(frozen)
    # independent copies of the slabs for the handle-free path
    p2, s2 = pop.params.clone(), pop.ms.clone()
    q2 = pop.mom.clone() if pop.mom is not None else None
    host_p = [np.asarray(x, np.float32) for x in params]
    host_ms = [np.ones_like(x) for x in host_p]
    host_q = [np.zeros_like(x) for x in host_p]
    for call, (seed, lr) in enumerate([(1, 7.0016e-4), (2, 3.0e-4)]):
        pop.load(pop.grads, _grads(cfg, seed))
        before = pop.params.clone()
        pop.apply_gradients(lr, 0.99, momentum, 0.1, 40.0)                       # existing entry point
        eng.rmsprop_apply_noalloc(p2, s2, pop.grads, pop.tensors, lr, 0.99, momentum, 0.1, 40.0,
                                  mom=q2, frozen=frozen)                          # new entry point
        torch.cuda.synchronize()
        for i, t in enumerate(pop.tensors):
            sl = slice(t["offset"], t["offset"] + t["length"])
            d_ref = (pop.params[sl] - before[sl]).double().cpu().numpy()
            d_new = (p2[sl] - before[sl]).double().cpu().numpy()
            if frozen[i]:
                assert np.all(d_new == 0.0) and torch.equal(s2[sl], pop.ms[sl])
                continue
            assert np.max(np.abs(d_new - d_ref)) <= REL_TOL * max(np.max(np.abs(d_ref)), 1e-30), (call, i)
            assert hp.rel_err(s2[sl].cpu().numpy(), pop.ms[sl].cpu().numpy()) < REL_TOL, (call, i)
            if q2 is not None:
                assert hp.rel_err(q2[sl].cpu().numpy(), pop.mom[sl].cpu().numpy()) < REL_TOL, (call, i)
            # same kernel bodies, same order of operations: not merely close
            assert torch.equal(p2[sl], pop.params[sl]) and torch.equal(s2[sl], pop.ms[sl]), (call, i)
            if q2 is not None:
                assert torch.equal(q2[sl], pop.mom[sl]), (call, i)
        assert hp.rel_err(p2.cpu().numpy(), pop.params.cpu().numpy()) < REL_TOL
        # and against the CPU oracle (clip_by_norm + ApplyRMSProp, rmsprop_applier.py:79-106), which shares no code with either
        want_p, want_ms, want_q = po.apply_gradients(host_p, host_ms, host_q, pop.dump(pop.grads), frozen, lr, 0.99, momentum, 0.1, 40.0)
        for i, t in enumerate(pop.tensors):
            sl = slice(t["offset"], t["offset"] + t["length"])
            upd = max(float(np.abs(want_p[i] - host_p[i]).max()), 1e-30)
            got = p2[sl].cpu().numpy().reshape(want_p[i].shape)
            assert np.abs(got - want_p[i]).max() <= REL_TOL * upd, (call, i)
            assert hp.rel_err(s2[sl].cpu().numpy().reshape(want_ms[i].shape), want_ms[i]) < 1e-5, (call, i)
        host_p, host_ms, host_q = want_p, want_ms, want_q
    # the update really happened (a skipped update would also sit inside 1e-3 of the parameters)
    sl = slice(pop.tensors[0]["offset"], pop.tensors[0]["offset"] + pop.tensors[0]["length"])
    assert float((s2[sl] - 1.0).abs().max()) > 1e-3
    pop.close()


def test_noalloc_odd_tables_and_user_stream(native_lib):
    """Tensor tables the handle form never produces, on a non-default stream, twice in a row, against the oracle."""
    from pathnet_gym_b200 import population as eng
    rng = np.random.default_rng(5)
    # (offset, length): unaligned start; not a multiple of 4; many chunks with a ragged tail; a single element
    table = [(3, 1001), (1100, 6), (1200, 5 * 2048 + 777), (12300, 1), (12400, 2048), (14501, 4097)]
    n = 19000
    hp_, hs_, hg_ = rng.normal(size=n).astype(np.float32), np.ones(n, np.float32), None
    p = torch.from_numpy(hp_.copy()).cuda()
    s = torch.from_numpy(hs_.copy()).cuda()
    q = torch.zeros(n, dtype=torch.float32, device="cuda")
    frozen = np.array([0, 0, 0, 0, 1, 0], np.uint8)
    tensors = [dict(offset=o, length=l) for o, l in table]
    st = torch.cuda.Stream()
    hq_ = np.zeros(n, np.float32)
    for call in range(2):
        hg_ = (rng.normal(size=n) * (25.0 if call == 0 else 0.01)).astype(np.float32)
        g = torch.from_numpy(hg_).cuda()
        torch.cuda.synchronize()
        with torch.cuda.stream(st):
            eng.rmsprop_apply_noalloc(p, s, g, tensors, 1e-3, 0.99, 0.5, 0.1, 40.0, mom=q, frozen=frozen)
        st.synchronize()
        args = ([hp_[o:o + l] for o, l in table], [hs_[o:o + l] for o, l in table], [hq_[o:o + l] for o, l in table],
                [hg_[o:o + l] for o, l in table], frozen, 1e-3, 0.99, 0.5, 0.1, 40.0)
        wp, ws, wq = po.apply_gradients(*args)
        got_p, got_s = p.cpu().numpy(), s.cpu().numpy()
        touched = np.zeros(n, bool)
        for (o, l), a, b, c, fz in zip(table, wp, ws, wq, frozen):
            upd = max(float(np.abs(a - hp_[o:o + l]).max()), 1e-30)
            assert np.abs(got_p[o:o + l] - a).max() <= (0.0 if fz else REL_TOL * upd), (call, o)
            assert hp.rel_err(got_s[o:o + l], b) < 1e-5, (call, o)
            hp_[o:o + l], hs_[o:o + l], hq_[o:o + l] = a, b, c
            touched[o:o + l] = True
        assert np.array_equal(got_p[~touched], hp_[~touched])          # the gaps between tensors are never written


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch(native_lib):
    from pathnet_gym_b200 import PathNetError, population as eng
    buf = torch.zeros(64, dtype=torch.float32, device="cuda")
    too_many = [dict(offset=0, length=4) for _ in range(129)]
    with pytest.raises(PathNetError, match="128"):
        eng.rmsprop_apply_noalloc(buf, buf.clone(), buf.clone(), too_many, 1e-3)
    too_big = [dict(offset=0, length=16385 * 2048)]
    with pytest.raises(PathNetError, match="16384"):
        eng.rmsprop_apply_noalloc(buf, buf.clone(), buf.clone(), too_big, 1e-3)
    with pytest.raises(PathNetError, match="momentum"):
        eng.rmsprop_apply_noalloc(buf, buf.clone(), buf.clone(), [dict(offset=0, length=64)], 1e-3, momentum=0.5)
    # a valid small call on the same buffers still works afterwards
    g = torch.ones(64, dtype=torch.float32, device="cuda")
    p, s = buf.clone(), torch.ones(64, dtype=torch.float32, device="cuda")
    eng.rmsprop_apply_noalloc(p, s, g, [dict(offset=0, length=64)], 1e-3)
    torch.cuda.synchronize()
    assert float(p.max()) < 0.0 and abs(float(s[0]) - 1.0) < 1e-6      # g=1: ms stays 1, var decreases
