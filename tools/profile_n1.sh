#!/bin/bash
# EP=1 evidence for profiles/: the plain bench line, then (separately, never timed) an ncu launch list of the same command
# and one --set full capture of the kernels of the step and of the grouped-layout calls.  Run under gpurun, one GPU.
mkdir -p gpurun_out
python bench.py > gpurun_out/p_bench_n1.json 2> gpurun_out/p_bench_n1.err; echo "bench rc=$?"
CMD="python bench.py --steps 5 --warmup 3 --no-ll --no-cpu --no-parity"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 120 --csv --log-file gpurun_out/p_launches_n1.csv $CMD > gpurun_out/p_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'layout_kernel|dispatch_local|combine_local|perm_|gather_local' -s 12 -c 14 -f -o gpurun_out/p_prof_n1 $CMD > gpurun_out/p_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/p_prof_n1.ncu-rep 2>/dev/null
cut -c1-400 gpurun_out/p_bench_n1.json
