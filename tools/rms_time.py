"""GPU time of pn_rmsprop_apply over the whole slab (launches queued behind a device-side sleep): python tools/rms_time.py [L N A]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
L, N, A = [int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (4, 4, 18))]
pop = PathNetPopulation(PathNetConfig(L=L, N=N, A=A), 4, 4); pop.init_params(1)
pop.set_genotypes(geopath_init(4, L, 10, N, 7, 0).cpu().numpy().view("uint32"))
for row in bench.measure_rmsprop(pop, torch): print(json.dumps(row))
pop.close()

# the same kernels where HBM, not launch latency, is the bound: a 33.5 M-parameter synthetic slab (the no-alloc entry point's
# largest input: 128 tensors x 128 chunks), slabs far larger than L2
from pathnet_gym_b200.population import rmsprop_apply_noalloc
n_t, per = 128, 128 * 2048
n = n_t * per
p = torch.randn(n, device="cuda"); s = torch.ones(n, device="cuda"); g = torch.randn(n, device="cuda")
tensors = [dict(offset=i * per, length=per) for i in range(n_t)]
for _ in range(3): rmsprop_apply_noalloc(p, s, g, tensors, 1e-6)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda._sleep(2_000_000)
e0.record()
for _ in range(10): rmsprop_apply_noalloc(p, s, g, tensors, 1e-6)
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 10 * 1e-3
print(json.dumps({"kernel": "rms_sumsq + rms_update on a 33.5 M-parameter slab (402 MB of slabs, > L2)", "bytes_per_launch": 20.0 * n,
                  "avg_launch_ms": round(t * 1e3, 4), "achieved_GBps": round(20.0 * n / t / 1e9, 1)}))
