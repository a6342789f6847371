#!/bin/bash
# quick EP=1 kernel timings for a few values of an environment switch: tools/quick_n1.sh VAR v1 v2 ...
var=$1; shift
for v in "$@"; do
  env $var=$v python bench.py --steps 30 --warmup 5 --no-ll --no-cpu --no-parity > gpurun_out/q_$v.json 2> gpurun_out/q_$v.err || { echo "$var=$v failed"; tail -3 gpurun_out/q_$v.err; continue; }
  python - <<PY
import json
j=json.load(open("gpurun_out/q_$v.json"))
k=j["roofline"]["kernels"]
print("$var=$v step %.1f us nosync %.1f us dispatch %.2f us (%.2f) combine %.2f us (%.2f)" % (j["ms_per_step"]*1e3, j["host_sync_free"]["ms_per_step"]*1e3, k["dispatch"]["ms"]*1e3, k["dispatch"]["frac"], k["combine"]["ms"]*1e3, k["combine"]["frac"]))
PY
done
