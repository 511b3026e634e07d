"""Multi-GPU correctness of the gradient exchange (skips when the box has fewer than 2 GPUs).

Each test launches tools/multigpu_check.py under torchrun on 2 GPUs (one process per GPU, NCCL, rendezvous on
127.0.0.1): every rank holds different paths, and the overlapped exchange of trainer.OverlappedGradSum -- the linear
region all-reduced on a side stream underneath the conv backward kernels, the small regions summed by the peer-memory
kernel (or by NCCL) -- must give the same gradient slab as ONE plain all-reduce of the whole slab; after a few
complete steps (exchange + clipped RMSProp) the parameter slabs of all ranks must be bit-identical, and the exchange
kernels' health word must be clean."""
import json
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(nproc, env_extra, port):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", **env_extra)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "multigpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, r.stdout[-2000:] + r.stderr[-2000:]
    return json.loads(lines[-1])


@pytest.mark.parametrize("mode", ["default", "nccl_only", "peer_lin", "peer_split", "lstm"])
def test_overlapped_exchange_equals_plain_allreduce(native_lib, mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = {"default": {}, "nccl_only": {"PN_PEER_SUM": "0"}, "peer_lin": {"PN_PEER_LIN": "1"}, "peer_split": {"PN_PEER_SPLIT": "1"},
           "lstm": {"PN_CHECK_LSTM": "1"}}[mode]
    out = _run(2, env, 29610 + {"default": 0, "nccl_only": 1, "peer_lin": 2, "peer_split": 3, "lstm": 4}[mode])
    assert out["world"] == 2
    assert out["overlapped_vs_plain_allreduce_rel_err"] < 1e-6
    assert out["params_identical_on_all_ranks"] is True
    assert out["trained_vs_plain_allreduce_training_rel_err"] < 1e-6
    if mode == "nccl_only":
        assert out["small_regions_via_peer_memory"] is False
    if mode == "peer_lin" and out["peer_setup_error"] is None:
        assert out["linear_region_via_peer_memory"] is True
