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
every other test.  Tolerance: the suite's 1e-3 relative (REL_TOL), applied to the update itself
(delta of the parameters) and to the slots.
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
        assert hp.rel_err(p2.cpu().numpy(), pop.params.cpu().numpy()) < REL_TOL
    # the update really happened (a skipped update would also sit inside 1e-3 of the parameters)
    sl = slice(pop.tensors[0]["offset"], pop.tensors[0]["offset"] + pop.tensors[0]["length"])
    assert float((s2[sl] - 1.0).abs().max()) > 1e-3
    pop.close()


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
