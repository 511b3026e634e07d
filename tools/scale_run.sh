#!/bin/bash
# usage: tools/scale_run.sh N  -> gpurun_out/scale_nN.json (bench.py at N GPUs, driver-style launch)
N=$1
if [ "$N" = "1" ]; then
  python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
fi
tail -c 400 gpurun_out/scale_n$N.json | head -c 10 > /dev/null
python - <<PY
import json
d=json.loads(open("gpurun_out/scale_n$N.json").read().strip().splitlines()[-1])
print("N=%d value=%.0f ms=%.4f instr=%.4f e2e=%.0f clocks=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["ms_per_step_instrumented"], d["e2e"]["value"], d["clocks"]))
PY
