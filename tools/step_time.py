"""Un-instrumented step time of any configuration: python tools/step_time.py L M N P B [A]   (C3 = 4 10 4 256 32)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
from pathnet_gym_b200.trainer import PopulationTrainer

L, M, N, P, B = [int(x) for x in sys.argv[1:6]]
A = int(sys.argv[6]) if len(sys.argv) > 6 else 18
cfg = PathNetConfig(L=L, M=M, N=N, A=A, lstm=bool(os.environ.get("PN_LSTM")))      # PN_LSTM=1: the LSTM variant
pop = PathNetPopulation(cfg, P, B); pop.init_params(1)
pop.set_genotypes(geopath_init(P, L, M, N, 7, 0).cpu().numpy().view("uint32"))
dev = pop.device
frames = [torch.randint(0, 256, (P, B, cfg.H, cfg.W, cfg.C), dtype=torch.uint8, device=dev) for _ in range(2)]
actions = torch.randint(0, A, (P, B), dtype=torch.int32, device=dev)
td, r = torch.randn(P, B, device=dev), torch.randn(P, B, device=dev)
tr = PopulationTrainer(pop)
for i in range(5): tr.step(frames[i & 1], actions, td, r, 1e-4)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 30
e0.record()
for i in range(K): tr.step(frames[i & 1], actions, td, r, 1e-4)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
if os.environ.get("PN_KERNELS"):
    pop.profile_enable(True)
    for i in range(20): tr.step(frames[i & 1], actions, td, r, 1e-4)
    torch.cuda.synchronize()
    for n, (a, c) in sorted(pop.profile_read().items(), key=lambda kv: -kv[1][0] * kv[1][1]): print("%-22s %7.1f us x %d" % (n, a * 1e3, c // 20))
    pop.profile_enable(False)
print(json.dumps({"L": L, "M": M, "N": N, "P": P, "B": B, "A": A, "ms_per_step": round(ms, 4), "env_steps_per_s": round(P * B / ms * 1e3, 1)}))
pop.close()
