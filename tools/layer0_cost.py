"""Layer-0 kernel time as a function of the number of active modules per path (uniform populations):
python tools/layer0_cost.py [P] [B]   -> one line per ns = 1..8 with conv_fwd_l0 / conv_wgrad_l0 in us."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
from pathnet_gym_b200.trainer import PopulationTrainer

P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
cfg = PathNetConfig(L=4, N=4)
pop = PathNetPopulation(cfg, P, B); pop.init_params(1)
base = geopath_init(P, 4, 10, 4, 7, 0).cpu().numpy().view("uint32").copy()
dev = pop.device
frames = [torch.randint(0, 256, (P, B, 84, 84, 4), dtype=torch.uint8, device=dev) for _ in range(2)]
actions = torch.randint(0, 18, (P, B), dtype=torch.int32, device=dev)
td, r = torch.randn(P, B, device=dev), torch.randn(P, B, device=dev)
tr = PopulationTrainer(pop)
rows = []
cases = [("ns=%d" % n, np.full(P, (1 << n) - 1, np.uint32)) for n in range(1, 9)]
mix = np.full(P, 0b111, np.uint32); mix[::10] = 0b11111
cases.append(("mix 90% ns=3 / 10% ns=5", mix))
for name, col in cases:
    m = base.copy(); m[:, 0] = col
    pop.set_genotypes(m)
    for i in range(3): tr.step(frames[i & 1], actions, td, r, 1e-4)
    pop.profile_enable(True)
    for i in range(10): tr.step(frames[i & 1], actions, td, r, 1e-4)
    torch.cuda.synchronize()
    pr = pop.profile_read(); pop.profile_enable(False)
    rows.append((name, pr["conv_fwd_l0"][0] * 1e3, pr["conv_wgrad_l0"][0] * 1e3))
    print("%-26s fwd %7.1f us   wgrad %7.1f us" % rows[-1], flush=True)
pop.close()
