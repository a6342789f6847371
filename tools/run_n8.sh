#!/bin/bash
# one 8-GPU measurement pass: bench at N=8 and N=4 (parity inside), two knob variants, production-call tests over CUDA IPC at world 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29555 bench.py --gpus 8 --breakdown > gpurun_out/f_bench_n8.json 2> gpurun_out/f_bench_n8.err; echo "bench8 rc=$?"
timeout 150 $TR --nproc-per-node 4 --master-port 29556 bench.py --gpus 4 --breakdown > gpurun_out/f_bench_n4.json 2> gpurun_out/f_bench_n4.err; echo "bench4 rc=$?"
bash tools/sweep_n.sh 8 "B200EP_HT_PIECE_ROWS=4" "B200EP_HT_DISPATCH_CTAS=2"
timeout 150 $TR --nproc-per-node 8 --master-port 29558 tests/mp_worker.py gpu > gpurun_out/f_mp_gpu_w8.log 2>&1; echo "mp rc=$?"; grep -c "GPU-MP-OK" gpurun_out/f_mp_gpu_w8.log
grep -h "breakdown" gpurun_out/f_bench_n8.err | head -2
python - <<'PY'
import json
for n in (8, 4):
    try:
        j = json.loads([l for l in open("gpurun_out/f_bench_n%d.json" % n) if l.startswith("{")][0])
        print(n, j["ms_per_step"], j["host_sync_free"]["ms_per_step"], j["parity"], j["ll"]["p50_us"], j["ll"]["sweep_p50_us"],
              j["roofline"]["step_frac"], {k: (v["send_kernel_ms"], v["frac"]) for k, v in j["roofline"]["kernels"].items()}, j["grouped"])
    except Exception as e:
        print(n, "failed", e)
PY
