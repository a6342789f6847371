#!/bin/bash
# one 8-GPU measurement pass: bench at N=8 and N=4 (parity inside), phase breakdown, in-kernel trace
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29555 bench.py --gpus 8 --breakdown > gpurun_out/s3_bench_n8.json 2> gpurun_out/s3_bench_n8.err; echo "bench8 rc=$?"
timeout 150 $TR --nproc-per-node 4 --master-port 29556 bench.py --gpus 4 --breakdown > gpurun_out/s3_bench_n4.json 2> gpurun_out/s3_bench_n4.err; echo "bench4 rc=$?"
timeout 100 $TR --nproc-per-node 8 --master-port 29557 tools/trace_report.py > gpurun_out/s3_trace_n8.txt 2>&1; echo "trace rc=$?"
grep -i "breakdown" gpurun_out/s3_bench_n8.err | head -3
python - <<'PY'
import json
for n in (8, 4):
    try:
        j = json.loads([l for l in open("gpurun_out/s3_bench_n%d.json" % n) if l.startswith("{")][0])
        print(n, j["ms_per_step"], j["host_sync_free"]["ms_per_step"], j["parity"], j["ll"]["p50_us"], j["ll"]["sweep_p50_us"],
              j["roofline"]["step_frac"], {k: (v["send_kernel_ms"], v["frac"]) for k, v in j["roofline"]["kernels"].items()})
    except Exception as e:
        print(n, "failed", e)
PY
