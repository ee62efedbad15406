#!/bin/bash
# Round-1 evidence capture (run under gpurun on one B200): tests, bench lines, ncu launch list, ncu full captures.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider 2>&1 | tail -3 > gpurun_out/pytest_gpu.txt
python bench.py > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1_reference.json 2>/dev/null
for w in preprocess search classify models gemm attention decode n4; do python bench_paths.py $w 2>/dev/null; done > gpurun_out/paths_r1.jsonl
python bench.py --steps 2 --warmup 3 --cpu-seconds 1 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 273 -c 91 --csv --log-file gpurun_out/launches_r1.csv \
    python bench.py --steps 2 --warmup 3 --cpu-seconds 1 > gpurun_out/ncu_launches.log 2>&1
python bench.py --steps 2 --warmup 3 --cpu-seconds 1 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|attention|layernorm|convert" -s 2 -c 8 -o gpurun_out/step_r1 \
    python bench.py --steps 2 --warmup 3 --cpu-seconds 1 > gpurun_out/ncu_step.log 2>&1
python bench_paths.py preprocess > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none -k regex:"preprocess_kernel" -s 2 -c 1 -o gpurun_out/k1_resize_r1 python bench_paths.py preprocess > gpurun_out/ncu_k1.log 2>&1
python bench_paths.py search --qs 1,4096 > gpurun_out/plain4.log 2>&1 && \
ncu --set full --clock-control none -k regex:"topk_tc|merge_topk" -s 2 -c 2 -o gpurun_out/search_q1_r1 python bench_paths.py search --qs 1 > gpurun_out/ncu_s1.log 2>&1
ncu --set full --clock-control none -k regex:"topk_tc" -s 2 -c 1 -o gpurun_out/search_q4096_r1 python bench_paths.py search --qs 4096 > gpurun_out/ncu_s2.log 2>&1
python scratch_tools/attn_one.py 128 577 16 > gpurun_out/plain5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"attention_tc" -s 2 -c 1 -o gpurun_out/attn_l14_r1 python scratch_tools/attn_one.py 128 577 16 > gpurun_out/ncu_attn.log 2>&1
ls -la gpurun_out
