#!/bin/bash
# tools/sweep_n.sh N "ENV1=a ENV2=b" "ENV3=c" ... : short EP=N runs (no LL, no CPU arm, no parity) of bench.py under each
# environment setting; prints step / push-kernel times.  Experiments only: the committed numbers come from plain bench.py.
n=$1; shift
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  out=gpurun_out/sweep_${n}_$i
  env $envs timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29720+i)) bench.py --gpus $n --steps 30 --warmup 5 --no-ll --no-cpu --no-parity --breakdown > $out.json 2> $out.err || { echo "[$envs] failed"; tail -5 $out.err; continue; }
  python - "$envs" $out.json <<'PY'
import json, sys
j = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][0])
k = j["roofline"]["kernels"]
print("[%s] step %.1f us nosync %.1f us | dispatch send %.1f us (+wait %.1f) | combine send %.1f us (send+reduce %.1f) | step_frac %.3f" % (
    sys.argv[1], j["ms_per_step"] * 1e3, j["host_sync_free"]["ms_per_step"] * 1e3, k["dispatch"]["send_kernel_ms"] * 1e3,
    k["dispatch"]["send_plus_wait_ms"] * 1e3, k["combine"]["send_kernel_ms"] * 1e3, k["combine"]["send_plus_reduce_ms"] * 1e3,
    j["roofline"]["step_frac"]))
PY
  grep "rank 0\] step breakdown" $out.err | head -1
done
