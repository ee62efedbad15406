#!/bin/bash
# Round-2 evidence capture (run under gpurun on one B200): tests, bench lines, ncu launch list, ncu full captures.
# Every ncu pass is preceded by the same command run plainly (exit code checked by &&), as the profiling recipe asks.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider 2>&1 | tail -4 > gpurun_out/pytest_gpu_r2.txt
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r2_reference.json 2>/dev/null
for w in preprocess search classify models gemm attention decode n4; do python bench_paths.py $w 2>/dev/null; done > gpurun_out/paths_r2.jsonl
python bench_paths.py search --gallery 10000000 2>/dev/null > gpurun_out/paths_search10m_r2.jsonl
S="python bench.py --steps 2 --warmup 3 --batches-per-step 1 --cpu-seconds 1 --skip-search --skip-configs"
$S > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 276 -c 92 --csv --log-file gpurun_out/launches_r2.csv $S > gpurun_out/ncu_launches.log 2>&1
$S > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|attention|layernorm|convert" -s 2 -c 8 -o gpurun_out/step_r2 $S > gpurun_out/ncu_step.log 2>&1
python bench_paths.py preprocess > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none -k regex:"preprocess_kernel" -s 2 -c 1 -o gpurun_out/k1_resize_r2 python bench_paths.py preprocess > gpurun_out/ncu_k1.log 2>&1
python bench_paths.py search --qs 64,4096 > gpurun_out/plain4.log 2>&1 && \
ncu --set full --clock-control none -k regex:"topk_tc|merge_topk" -s 2 -c 2 -o gpurun_out/search_q64_r2 python bench_paths.py search --qs 64 > gpurun_out/ncu_s1.log 2>&1
ncu --set full --clock-control none -k regex:"topk_tc" -s 2 -c 2 -o gpurun_out/search_q4096_r2 python bench_paths.py search --qs 4096 > gpurun_out/ncu_s2.log 2>&1
ls -la gpurun_out
