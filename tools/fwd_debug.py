"""PN_TC_DEBUG=7|4 python tools/fwd_debug.py [L N P B]: where the roles of conv0_fwd_tc_kernel spend their cycles
(mean / max over CTAs).  Mode 7: epilogue group 0, warp 0; mode 4: producer and epilogue waits."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
L, N, P, B = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (4, 3, 256, 32))]
mode = int(os.environ.get("PN_TC_DEBUG", "7"))
cfg = PathNetConfig(L=L, N=N)
pop = PathNetPopulation(cfg, P, B); pop.init_params(1)
pop.set_genotypes(geopath_init(P, L, 10, N, 7, 0).cpu().numpy().view("uint32"))
dev = pop.device
frames = torch.randint(0, 256, (P, B, 84, 84, 4), dtype=torch.uint8, device=dev)
for i in range(3): pop.forward(frames)
torch.cuda.synchronize()
buf = (C.c_longlong * (148 * 8))()
n = pop.lib.pn_debug_read(pop._h, buf, 148 * 8)
a = np.array(buf[:n]).reshape(-1, 8)
names = {7: ["epi: wait path", "epi: wait acc_full", "epi: tmem ld", "epi: math + mask store", "epi: total", "epi: y store", "-", "-"],
         4: ["producer: wait frame_empty", "producer: total", "-", "-", "-", "-", "epi: wait acc_full", "epi: total"]}[mode]
for i, nm in enumerate(names):
    if nm != "-": print("%-28s mean %9.0f  max %9.0f" % (nm, a[:, i].mean(), a[:, i].max()))
pop.close()
