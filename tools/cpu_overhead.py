"""How much host time does one training step cost?  torchrun --nproc-per-node N tools/cpu_overhead.py
Prints per-step wall time (GPU-synchronised over the whole loop) and the host time spent issuing the step."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from pathnet_gym_b200.population import PathNetConfig, PathNetPopulation, geopath_init
from pathnet_gym_b200.trainer import PopulationTrainer

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg, P, B = PathNetConfig(), 64, 32
pop = PathNetPopulation(cfg, P, B, device=local); pop.init_params(1)
pop.set_genotypes(geopath_init(P, cfg.L, cfg.M, 3, 7, 0, device=local).cpu().numpy().view("uint32"))
dev = pop.device
frames = torch.randint(0, 256, (P, B, cfg.H, cfg.W, cfg.C), dtype=torch.uint8, device=dev)
actions = torch.randint(0, cfg.A, (P, B), dtype=torch.int32, device=dev)
td, r = torch.randn(P, B, device=dev), torch.randn(P, B, device=dev)
tr = PopulationTrainer(pop)
for _ in range(20): tr.step(frames, actions, td, r, 1e-4)
torch.cuda.synchronize()
K = 40
t0 = time.perf_counter(); host = 0.0
for _ in range(K):
    a = time.perf_counter(); tr.step(frames, actions, td, r, 1e-4); host += time.perf_counter() - a
issue_done = time.perf_counter() - t0
torch.cuda.synchronize()
wall = time.perf_counter() - t0
# split of the host time
a = time.perf_counter()
for _ in range(K): pop.train_step(frames, actions, td, r, 0.01)
t_train = (time.perf_counter() - a) / K
torch.cuda.synchronize()
a = time.perf_counter()
for _ in range(K): tr.tail(1e-4)
t_tail = (time.perf_counter() - a) / K
torch.cuda.synchronize()
if rank == 0:
    print(json.dumps({"world": world, "wall_us_per_step": wall / K * 1e6, "host_issue_us_per_step": host / K * 1e6,
                      "loop_issue_done_us_per_step": issue_done / K * 1e6, "host_train_step_us": t_train * 1e6, "host_tail_us": t_tail * 1e6}))
pop.close()
if world > 1: dist.destroy_process_group()
