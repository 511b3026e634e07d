"""world_size-2 gloo tests (CPU) of the host-side sharding logic used at N > 1."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pathnet_gym_b200 import parallel


def test_shard_range_partitions_population():
    for P in (1, 7, 64, 512):
        for ws in (1, 2, 3, 8):
            got, prev = [], 0
            for r in range(ws):
                lo, hi = parallel.shard_range(P, r, ws)
                assert lo == prev and hi >= lo
                prev = hi
                got.append(hi - lo)
            assert prev == P and max(got) - min(got) <= 1
    assert parallel.shard_range(512, 3, 8) == (192, 256)


def _worker(rank, ws, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        # gradient slab: rank r contributes r+1 everywhere -> sum = ws(ws+1)/2
        g = torch.full((1000,), float(rank + 1))
        parallel.allreduce_sum_(g)
        ok = bool(torch.all(g == ws * (ws + 1) / 2))
        # scores of uneven slices gathered in rank order
        P = 5
        lo, hi = parallel.shard_range(P, rank, ws)
        local = torch.arange(lo, hi, dtype=torch.float32)
        sizes = [parallel.shard_range(P, r, ws) for r in range(ws)]
        allv = parallel.allgather_concat(local, [b - a for a, b in sizes])
        ok = ok and torch.equal(allv, torch.arange(P, dtype=torch.float32))
        ok = ok and parallel.world() == (rank, ws)
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_gloo_world_size_2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


def test_checkpoint_roundtrip_and_tf_names(tmp_path):
    """tensors in get_vars() order, slots, genotypes, frozen path; TF-Saver name mapping
    (game_ac_network.py:125-148: variables are created in get_vars() order inside scope net_<i>)."""
    import numpy as np
    from oracle import pathnet_oracle as po
    from pathnet_gym_b200 import checkpoint as ck
    from tests import helpers as hp
    cfg = hp.TINY
    tt = po.tensor_table(cfg)
    assert ck.tensor_names(cfg) == [t["name"] for t in tt]
    names = ck.tf_saver_names(cfg)
    assert names[0] == "net_0/Variable" and names[1] == "net_0/Variable_1" and len(names) == len(tt)
    params = po.init_params(cfg, 3)
    rms = [np.full(t["shape"], 1.0 + i, np.float32) for i, t in enumerate(tt)]
    geno = np.asarray([[3, 5, 6], [9, 10, 12]], np.uint32)
    fixed = np.zeros((cfg.L, cfg.M), np.float32); fixed[0, 1] = 1
    path = str(tmp_path / "ck.npz")
    ck.save_checkpoint(path, cfg, params, rms, None, geno, fixed, global_t=1234, extra={"scores": np.arange(2.0)})
    got = ck.load_checkpoint(path, cfg)
    assert got["global_t"] == 1234 and got["momentum"] is None
    assert all(np.array_equal(a, b) for a, b in zip(got["params"], params))
    assert all(np.array_equal(a, b) for a, b in zip(got["rms"], rms))
    assert np.array_equal(got["genotypes"], geno) and np.array_equal(got["fixed_path"], fixed)
    assert np.array_equal(got["extra"]["scores"], np.arange(2.0))
    # a reference Saver dump maps back onto the slab order through its variable names
    tfvars = {nm: p for nm, p in zip(names, params)}
    back = ck.from_tf_variables(cfg, tfvars)
    assert all(np.array_equal(a, b) for a, b in zip(back, params))
    import pytest
    with pytest.raises(ValueError):
        ck.load_checkpoint(path, hp.TINY4)
    # the LSTM network appends W_lstm / b_lstm (game_ac_network.py:502-506), saved by TF under the cell's own names
    lcfg = hp.TC20_LSTM
    ltt = po.tensor_table(lcfg)
    assert ck.tensor_names(lcfg) == [t["name"] for t in ltt] and ck.tensor_names(lcfg)[-2:] == ["W_lstm", "b_lstm"]
    lnames = ck.tf_saver_names(lcfg, 2)
    assert lnames[-2:] == ["net_2/basic_lstm_cell/kernel", "net_2/basic_lstm_cell/bias"] and lnames[-3] == "net_2/Variable_%d" % (len(ltt) - 3)
    lparams = po.init_params(lcfg, 5)
    lpath = str(tmp_path / "ck_lstm.npz")
    ck.save_checkpoint(lpath, lcfg, lparams, lparams)
    lgot = ck.load_checkpoint(lpath, lcfg)
    assert all(np.array_equal(a, b) for a, b in zip(lgot["params"], lparams))
    with pytest.raises(ValueError):                         # a feed-forward configuration refuses the LSTM checkpoint
        ck.load_checkpoint(lpath, po.OracleConfig(**{**lcfg.__dict__, "lstm": False}))
