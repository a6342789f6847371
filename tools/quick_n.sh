#!/bin/bash
# quick EP=N timings for a few values of an environment switch: tools/quick_n.sh N VAR v1 v2 ...
n=$1; var=$2; shift; shift
for v in "$@"; do
  env $var=$v python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29720 bench.py --gpus $n --steps 20 --warmup 5 --no-ll --no-cpu --no-parity --breakdown > gpurun_out/qn_${var}_$v.json 2> gpurun_out/qn_${var}_$v.err || { echo "$var=$v failed"; tail -5 gpurun_out/qn_${var}_$v.err; continue; }
  python - <<PY
import json
j=json.loads([l for l in open("gpurun_out/qn_${var}_$v.json") if l.startswith("{")][0])
k=j["roofline"]["kernels"]
print("N=$n $var=$v step %.1f us nosync %.1f us | dispatch send %.1f us (+wait %.1f) | combine send %.1f us (send+reduce %.1f) | step_frac %.3f" % (j["ms_per_step"]*1e3, j["host_sync_free"]["ms_per_step"]*1e3, k["dispatch"]["send_kernel_ms"]*1e3, k["dispatch"]["send_plus_wait_ms"]*1e3, k["combine"]["send_kernel_ms"]*1e3, k["combine"]["send_plus_reduce_ms"]*1e3, j["roofline"]["step_frac"]))
PY
  grep "rank 0\] step breakdown" gpurun_out/qn_${var}_$v.err | head -1
done
