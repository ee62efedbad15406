"""Writes profiles/README.md from the captured JSON lines of the current round (run after capture_r2.sh + summarise.py):
    python profiles/make_readme.py"""
import collections
import csv
import json
from pathlib import Path

P = Path(__file__).resolve().parent
TAG = "r2"


def jl(name):
    f = P / name
    return [json.loads(l) for l in f.read_text().splitlines() if l.strip().startswith("{")] if f.exists() else []


def one(name):
    x = jl(name)
    return x[-1] if x else None


paths, s10 = jl(f"paths_{TAG}.jsonl"), jl(f"paths_search10m_{TAG}.jsonl")
b, ref = one(f"bench_{TAG}.json"), one(f"bench_{TAG}_reference.json")
peaks = json.loads((P.parent / "MEASURED_PEAKS.json").read_text()) if (P.parent / "MEASURED_PEAKS.json").exists() else {}
L = []
L.append(f"# profiles - round 2 (B200, sm_100a, driver 580, CUDA 12.9)\n")
L.append("Produced by `profiles/capture_r2.sh` on one B200 (`gpurun`), summarised with `profiles/summarise.py gpurun_out r2`, this file by "
         "`profiles/make_readme.py`.  Fractions are of the measured peaks in `MEASURED_PEAKS.json` "
         f"(HBM copy {peaks.get('hbm_gbs')} GB/s, bf16 {peaks.get('bf16_tflops')} TFLOP/s burst / {peaks.get('bf16_tflops_sustained')} sustained).  "
         "Round-1 files (`*_r1*`) are kept for comparison where the kernel still exists.\n")
L.append("| file | what |\n|---|---|")
L.append(f"| `bench_{TAG}.json` / `bench_{TAG}_reference.json` | `python bench.py --steps 20 --warmup 5` and `--impl reference` lines |")
L.append(f"| `bench_{TAG}_n2.json`, `_n4.json`, `_n8.json` | the same line under torchrun on 2 / 4 / 8 GPUs of one box (NCCL): weak-scaling embed, strong-scaling search and C2 |")
L.append(f"| `bench_{TAG}_two_in_flight.json`, `bench_{TAG}_single_stream.json` | the headline part of the line twice on one box, back to back: two batches in flight (the default) and `--single-stream` |")
L.append(f"| `paths_{TAG}.jsonl`, `paths_search10m_{TAG}.jsonl` | `bench_paths.py` per-kernel lines (search sweep on a 1.25 M-row shard and on 10 M rows) |")
L.append(f"| `launches_{TAG}.csv` | ncu `gpu__time_duration.sum` for every launch of one 512-image pass (cold cache, serialised: compare shares) |")
L.append(f"| `step_{TAG}_ncu.csv` | `--set full` metrics of the per-layer kernels (GEMM shapes, attention_short_kernel, LayerNorm) |")
L.append(f"| `k1_resize_{TAG}_ncu.csv`, `k1_ident_{TAG}_ncu.csv`, `search_q64_{TAG}_ncu.csv`, `search_q4096_{TAG}_ncu.csv` | same for K1 (640x480 -> 224; identity) and K10 / K12 |")
L.append(f"| `gemm_traffic_{TAG}.json` | dram bytes per GEMM launch (feeds `roofline.traffic` in bench.py) |")
L.append(f"| `sass_summary_{TAG}.txt` | `cuobjdump -sass` instruction counts per kernel: tcgen05 / TMEM / TMA mnemonics, no legacy HMMA anywhere |")
L.append(f"| `pytest_gpu_{TAG}.txt` | tail of `pytest -m gpu` on the same box |\n")
if b:
    r = b["roofline"]
    L.append(f"## Headline (`bench_{TAG}.json`): steady state, 20 steps of 20 x 512 images\n")
    L.append(f"* value **{b['value']:.0f} images/s** (device-resident batches), e2e **{b['e2e']['value']:.0f} images/s** (pinned host in, host out, "
             f"{b['e2e']['h2d_bytes_per_step'] / 1e6:.0f} MB H2D + {b['e2e']['d2h_bytes_per_step'] / 1e6:.0f} MB D2H per step inside the timed region), "
             f"{b['ms_per_step']:.1f} ms per step ({b['ms_per_step'] / b['config']['batches_per_step']:.2f} ms per 512-image batch), timed region "
             f"{b['ms_per_step'] * b['steps'] / 1e3:.2f} s, {b['gpu_launches']} kernel launches in it.")
    L.append(f"* clocks during the timed region: {b['clocks']} - the run sits at the 1 kW power cap, so the regime is `{r['regime']}`.")
    L.append(f"* dominant kernels (tcgen05 GEMMs, separate event pass after the timed region): {r['achieved']:.0f} TFLOP/s = **{r['frac_vs_sustained']:.3f} of the "
             f"sustained / {r['frac_vs_burst']:.3f} of the burst bf16 peak**; whole step (algorithmic 8.8176 GFLOP per image): "
             f"{r['whole_step']['frac_vs_sustained']:.3f} / {r['whole_step']['frac_vs_burst']:.3f}; CUDA-event share of a step {r['share_of_step']}; "
             f"dram traffic per GEMM launch (ncu) {(r['traffic'] or 0) / 1e6:.0f} MB.")
    if ref:
        L.append(f"* CPU arm (`--impl reference`, oracle port, {ref['cpu_baseline']['cores']} host threads): {ref['value']:.0f} images/s; in-run CPU baseline "
                 f"{b['cpu_baseline']['value']:.0f} images/s.")
    f_ = P / f"launches_{TAG}.csv"
    if f_.exists():
        rows = list(csv.reader(open(f_)))
        hi = [i for i, r_ in enumerate(rows) if r_ and r_[0] == "ID"][0]
        ki, vi = rows[hi].index("Kernel Name"), rows[hi].index("Metric Value")
        agg, tot = collections.Counter(), 0.0
        for r_ in rows[hi + 2:]:
            if len(r_) <= vi:
                continue
            t, n = float(r_[vi].replace(",", "")), r_[ki]
            key = "GEMM" if "gemm_tc" in n else "attention" if "attention" in n else "LayerNorm" if ("layernorm" in n or "preln" in n) else \
                "preprocess" if ("preprocess" in n or "convert" in n or "pull_desc" in n) else "other"
            agg[key] += t; tot += t
        L.append(f"* ncu launch list (`launches_{TAG}.csv`) shares of one pass: " + ", ".join(f"{k} {v / tot:.3f}" for k, v in agg.most_common()) +
                 " - agrees with the CUDA-event shares above.\n")
    two, single = one(f"bench_{TAG}_two_in_flight.json"), one(f"bench_{TAG}_single_stream.json")
    if two and single:
        L.append(f"* two batches in flight against one, same box, back to back (`bench_{TAG}_two_in_flight.json`, `bench_{TAG}_single_stream.json`): "
                 f"{two['value']:.0f} vs {single['value']:.0f} images/s device-resident ({two['value'] / single['value'] - 1:+.1%}), "
                 f"{two['e2e']['value']:.0f} vs {single['e2e']['value']:.0f} end to end ({two['e2e']['value'] / single['e2e']['value'] - 1:+.1%}).\n")
    s = b.get("search")
    if s:
        L.append(f"## Search block of the same line (BASELINE configs[4] on {s['n_gpus']} GPU: {s['rows_per_gpu']} rows per GPU)\n")
        L.append(f"* Q = 4096, k = 100: **{s['ms']:.2f} ms** = {s['queries_per_s']:.0f} queries/s; {s['roofline']['frac']:.3f} of the burst tensor peak "
                 f"({s['roofline']['frac_vs_sustained']:.3f} of sustained); phases: local top-k {s['local_topk_ms']:.2f} ms, all-gather {s['allgather_ms']:.3f} ms, merge {s['merge_ms']:.3f} ms.")
        L.append("* small batches on the same shard (HBM-bound): " + ", ".join(f"Q={x['Q']}: {x['ms']:.2f} ms = {x['hbm_frac']:.2f} of HBM" for x in s["small_batch_sweep"]["runs"]) + ".")
        if "mid_batch_sweep" in s:
            L.append("* a few query tiles, same shard: " + ", ".join(f"Q={x['Q']}: {x['ms']:.2f} ms = {x['tensor_frac_of_burst']:.2f} of the burst tensor peak / {x['hbm_frac']:.2f} of HBM" for x in s["mid_batch_sweep"]["runs"]) + ".")
        if "cpu_baseline" in s:
            L.append(f"* CPU baseline (reference per-request search, extrapolated): {s['cpu_baseline']['value']:.3f} queries/s on {s['cpu_baseline']['cores']} threads.\n")
    c = b.get("configs") or {}
    if c:
        L.append("## Other configs of the same line\n")
        if "c1_vit_b32_batch1" in c:
            L.append(f"* C1 ViT-B/32 batch 1: {c['c1_vit_b32_batch1']['ms_per_image_device_resident']:.3f} ms device-resident, {c['c1_vit_b32_batch1']['ms_per_image_host_to_host']:.3f} ms host to host.")
        for key, nm in (("c3_vit_b16_batch256_classify1000", "C3 ViT-B/16 batch 256"), ("c4_vit_l14_336_batch256", "C4 ViT-L/14@336 batch 256")):
            if key in c:
                x = c[key]
                L.append(f"* {nm}: {x['images_per_s']:.0f} images/s = {x['frac_of_sustained_peak']:.3f} of sustained / {x['frac_of_burst_peak']:.3f} of burst"
                         + (f"; classify 256 x 1000 labels: {x['classify_us']:.0f} us" if "classify_us" in x else "") + ".")
        if "k1_preprocess" in c:
            L.append("* K1 alone: " + ", ".join(f"{x['src']}: {x['hbm_frac']:.3f} of HBM" for x in c["k1_preprocess"]["runs"]) + ".\n")
L.append("## Multi-GPU (same bench line under torchrun)\n\n| GPUs | embed images/s (weak) | e2e images/s | C2 strong (512 images in total) | search ms (10 M rows sharded) | local / all-gather / merge ms | Q=1/16/64/128 of HBM on the local shard |\n|---|---|---|---|---|---|---|")
for n in (1, 2, 4, 8):
    d = b if n == 1 else one(f"bench_{TAG}_n{n}.json")
    if not d:
        continue
    s = d.get("search") or {}
    c2 = (d.get("configs") or {}).get("c2_strong_scaling")
    sw = "/".join(f"{x['hbm_frac']:.2f}" for x in s.get("small_batch_sweep", {}).get("runs", []))
    L.append(f"| {n} | {d['value']:.0f} | {d['e2e']['value']:.0f} | {(str(round(c2['images_per_s'])) + ' images/s, ' + str(round(c2['ms_per_batch'], 2)) + ' ms') if c2 else '(= headline batch)'} | "
             f"{s.get('ms', float('nan')):.2f} | {s.get('local_topk_ms', 0):.2f} / {s.get('allgather_ms', 0):.3f} / {s.get('merge_ms', 0):.3f} | {sw} |")
L.append("\nThe all-gather column is dominated by the ranks' skew at the end of a 5-35 ms kernel (the collective itself moves Q*k*8 = 3.3 MB per GPU: 30 us back to back).\n")
L.append("## Per-shape GEMM (`paths_r2.jsonl`, M = 25,600 token rows, kernel timed alone)\n\n"
         "Burst = 20 launches from idle clocks; steady = a ~0.4 s loop after a warm-up loop (the power-capped regime a long pass runs in), next to "
         "cuBLASLt (torch `F.linear`: bias epilogue, bf16 output, no activation / residual) on the same operands in the same regime.  Epilogue 0 = fp32 store, "
         "1 = bias + bf16, 2 = bias + activation + bf16, 3 = bias + fp32 residual reduce-add.\n\n"
         "| shape | N | K | epilogue | burst us | burst TFLOP/s | of burst peak | steady us | steady TFLOP/s | of sustained peak | cuBLASLt steady TFLOP/s | ours / cuBLASLt |\n|---|---|---|---|---|---|---|---|---|---|---|---|")
for d in paths:
    if d.get("path") == "gemm_tc_kernel":
        if "steady_us" in d:
            L.append(f"| {d['shape']} | {d['N']} | {d['K']} | {d['epilogue']} | {d['us']:.1f} | {d['tflops']:.0f} | {d['frac_burst']:.3f} | {d['steady_us']:.1f} | {d['steady_tflops']:.0f} | "
                     f"{d['frac_sustained']:.3f} | {d['cublaslt_bias_bf16_steady_tflops']:.0f} | {d['vs_cublaslt']:.2f} |")
        else:
            L.append(f"| {d['shape']} | {d['N']} | {d['K']} | {d['epilogue']} | {d['us']:.1f} | {d['tflops']:.0f} | {d['frac_burst']:.3f} | | | {d['frac_sustained']:.3f} | | |")
L.append("\nThe fp32 residual stream is what separates epilogue 3 from the library figure: the same shapes with a bf16 output (`*_bf16` rows) sit within a few per cent of it, "
         "and the K = N = 768 out-projection with 196 MB of traffic per 30 GFLOP (39 MB in, 78 MB read-modify-write) is bound by HBM, not by the tensor pipe.\n")
L.append("\n## K5 attention (`paths_r2.jsonl`)\n\n| tower | B x S x heads | kernel | us | of HBM (algorithmic bytes) | of the SFU exp2 rate |\n|---|---|---|---|---|---|")
for d in paths:
    if d.get("path") == "K5 attention":
        L.append(f"| {d['model']} | {d['B']} x {d['S']} x {d['heads']} | {d['kernel']} | {d['us']:.1f} | {d['hbm_frac']:.3f} | {d['frac_of_sfu_exp2_peak']:.3f} |")
L.append("\nS = 50 (ViT-B/32): round 1's mma.sync kernel took 50 us per layer; `attention_short_kernel` (tcgen05, four small CTAs per SM - the TMEM limit) takes 32.  "
         "What it waits for is the hand-off chain of a tile (TMA -> MMA -> softmax -> MMA -> store), not bytes: boxes of exactly 50 rows instead of 64 (28 % fewer bytes) changed nothing.\n")
for title, rows_ in (("1.25 M-row shard (the per-GPU share of C5)", [d for d in paths if d.get("path") == "K10+K12 gallery search"]), ("10 M rows on one GPU", s10)):
    if rows_:
        L.append(f"## K10 + K12 gallery search, {title}, k = 100\n\n| Q | ms | queries/s | of burst bf16 peak | of HBM peak |\n|---|---|---|---|---|")
        for d in rows_:
            L.append(f"| {d['Q']} | {d['ms']:.3f} | {d['queries_per_s']:.0f} | {d['roofline']['tensor_frac']:.3f} | {d['roofline']['hbm_frac']:.3f} |")
        L.append("")
L.append("Round 1 on the 1.25 M-row shard: Q = 1 / 16 / 64 / 128 at 0.70 / 0.58 / 0.35 / 0.23 of HBM; the maxima-sharing bound, the two-pass first tile and the per-query "
         "candidate lists are what moved them (DESIGN.md §4).  At Q = 4096 the kernel is paced by the survivor extraction in the epilogue warps, not by the tensor pipe: with the scan "
         "switched off it runs at 0.89 of the burst peak (3.54 ms), which is the headroom left.\n")
L.append("## K1 preprocess against the HBM roofline (`paths_r2.jsonl`)\n\n| source | batch | ms | images/s | of HBM peak |\n|---|---|---|---|---|")
for d in paths:
    if d.get("path") == "K1 preprocess":
        L.append(f"| {d['src']} | {d['batch']} | {d['ms']:.3f} | {d['images_per_s']:.0f} | {d['roofline']['frac']:.3f} |")
L.append("\nIdentity sizes write two bytes for every byte they read; a pure write stream (`torch.zero_` on 617 MB) tops out at 3.9 TB/s on this part, so 0.91 of the copy peak is the "
         "ceiling for this mix.  Down-scaling stays bound by issue slots and barrier stalls (ncu `k1_resize_r2_ncu.csv`), not by HBM.\n")
other = [d for d in paths if d.get("path", "").startswith(("embed", "K11", "P0", "N4"))]
if other:
    L.append("## Other paths (`paths_r2.jsonl`)\n")
    for d in other:
        keys = [k for k in ("model", "batch", "src", "host_threads", "ms", "us", "images_per_s", "texts_per_s", "rows_per_s", "frac_of_tensor_peak_sustained") if k in d]
        L.append("* " + d["path"] + ": " + ", ".join(f"{k} {round(d[k], 3) if isinstance(d[k], float) else d[k]}" for k in keys))
(P / "README.md").write_text("\n".join(L) + "\n")
print("wrote README.md", len(L), "lines")
