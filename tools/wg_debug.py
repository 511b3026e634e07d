"""PN_TC_DEBUG=8 python tools/wg_debug.py [L N P B]: where the roles of conv0_wgrad_tc_kernel wait (cycles, mean / max over CTAs)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
L, N, P, B = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (4, 4, 256, 32))]
cfg = PathNetConfig(L=L, N=N)
pop = PathNetPopulation(cfg, P, B); pop.init_params(1)
pop.set_genotypes(geopath_init(P, L, 10, N, 7, 0).cpu().numpy().view("uint32"))
dev = pop.device
frames = torch.randint(0, 256, (P, B, 84, 84, 4), dtype=torch.uint8, device=dev)
actions = torch.randint(0, 18, (P, B), dtype=torch.int32, device=dev)
td, r = torch.randn(P, B, device=dev), torch.randn(P, B, device=dev)
for i in range(3): pop.train_step(frames, actions, td, r)
torch.cuda.synchronize()
buf = (C.c_longlong * (148 * 8))()
n = pop.lib.pn_debug_read(pop._h, buf, 148 * 8)
a = np.array(buf[:n]).reshape(-1, 8)
names = ["mma: wait frame", "mma: wait dq", "mma: wait acc_empty", "mma: total", "quant: wait dq_empty", "flush: wait acc_full", "flush: work", "quant: total"]
for i, nm in enumerate(names): print("%-24s mean %9.0f  max %9.0f" % (nm, a[:, i].mean(), a[:, i].max()))
pop.close()
