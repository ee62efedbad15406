#!/usr/bin/env python
"""Headline benchmark (BASELINE.json configs[1]): CLIP ViT-B/32 image embedding, batch 512 synthetic 224x224
uint8 images per GPU, preprocess + ViT encode in bf16, data-parallel (weak scaling, no collective).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--model ViT-B-32] [--src 224x224]

One step = one pass of the hot path (K1 preprocess -> K2..K9 encoder -> L2 normalise) over one batch.
`value`   images/s with the decoded uint8 batch already resident in HBM (CUDA events, max over ranks);
`e2e`     the same through the backend's host-buffer call `embed_stream`: pinned host uint8 -> H2D -> kernels -> D2H float32
          every step, the upload of batch i+1 overlapping the encode of batch i;
`roofline` the tcgen05 GEMM kernel: algorithmic FLOPs of all its launches / their summed CUDA-event time;
`cpu_baseline` the oracle (CPU restatement of the reference path) timed on this box's host cores.
`--impl reference` times that CPU path alone (the reference has no GPU-independent build of its ONNX path here).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from lumen_b200.vit_config import VIT_CONFIGS, synthetic_weights  # noqa: E402

METRIC = "clip_image_embed_throughput"
UNIT = "images/s"


def peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d["hbm_gbs"], "bf16_sustained": d["bf16_tflops_sustained"], "bf16_burst": d["bf16_tflops"], "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_sustained": 1400.0, "bf16_burst": 1590.0, "src": "fallback"}


def measured_gemm_traffic():
    """dram__bytes_read+write per GEMM launch from the committed ncu --set full capture (None if absent)."""
    p = ROOT / "profiles" / "gemm_traffic_r1.json"
    try:
        return json.loads(p.read_text())["dram_bytes_per_launch_mean"]
    except Exception:  # noqa: BLE001
        return None


def gemm_flops_per_image(cfg: dict) -> float:
    n = (cfg["image_size"] // cfg["patch"]) ** 2
    S, d, L, mlp, D, P = n + 1, cfg["width"], cfg["layers"], cfg["mlp_dim"], cfg["embed_dim"], cfg["patch"]
    kpad = (3 * P * P + 63) // 64 * 64
    return 2.0 * n * kpad * d + L * (2.0 * S * d * 3 * d + 2.0 * S * d * d + 4.0 * S * d * mlp) + 2.0 * d * D


def total_flops_per_image(cfg: dict) -> float:
    n = (cfg["image_size"] // cfg["patch"]) ** 2
    S, d, L, mlp, D, P = n + 1, cfg["width"], cfg["layers"], cfg["mlp_dim"], cfg["embed_dim"], cfg["patch"]
    return 2.0 * n * 3 * P * P * d + L * (2.0 * S * d * 3 * d + 4.0 * S * S * d + 2.0 * S * d * d + 4.0 * S * d * mlp) + 2.0 * d * D


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def workload_config(args, cfg: dict, src_hw, world: int) -> dict:
    return {"workload": f"{args.model} image embedding: batch {args.batch} per GPU of {src_hw[0]}x{src_hw[1]} uint8 RGB -> preprocess (Pillow-exact) + ViT encode -> "
                        f"{cfg['embed_dim']}-d unit vectors", "global_batch": args.batch * world,
            "parallelism": f"dp{world} (replicated weights, no collective)"}


# ---------------------------------------------------------------------------------------------- CPU (oracle) arm
def cpu_reference_rate(cfg: dict, weights: dict, src_hw: tuple[int, int], budget_s: float, max_images: int, seed: int = 1234):
    """Times the oracle (reference closure + fp32 tower restatement) on host cores; returns (img/s, images, threads)."""
    import torch
    from oracle import clip_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rng = np.random.default_rng(seed)
    chunk = 32
    imgs = list(rng.integers(0, 256, (chunk, src_hw[0], src_hw[1], 3), dtype=np.uint8))
    O.vit_encode(weights, O.preprocess_batch(imgs[:2], cfg["image_size"], 0), cfg)        # warm-up (thread pool, oneDNN)
    done, t0 = 0, time.perf_counter()
    while done < max_images and (time.perf_counter() - t0) < budget_s:
        O.vit_encode(weights, O.preprocess_batch(imgs, cfg["image_size"], 0), cfg)
        done += chunk
    dt = time.perf_counter() - t0
    return done / dt, done, threads


def run_reference(args, cfg, weights, src_hw):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rates = []
    for _ in range(args.warmup + args.steps):
        r, n, threads = cpu_reference_rate(cfg, weights, src_hw, budget_s=max(4.0, 60.0 / max(1, args.steps + args.warmup)), max_images=128)
        rates.append((r, n))
    timed = rates[args.warmup:]
    value = float(np.mean([r for r, _ in timed]))
    sample = f"{timed[0][1]} images per step ({src_hw[0]}x{src_hw[1]} uint8), fp32, batches of 32, oracle port of the reference CPU path"
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * timed[0][1] / value, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": dict(workload_config(args, cfg, src_hw, max(1, args.gpus)), sample=sample),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------- GPU arm
def run_b200(args, cfg, weights, src_hw):
    import torch
    import torch.distributed as dist
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    from lumen_b200.engine import pack_images

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    B = args.batch
    be = B200ClipBackend(SyntheticResources(model_name=args.model), device_preference=f"cuda:{local}", max_batch_size=B, weights=weights)
    be.initialize()
    eng = be.engine
    rng = np.random.default_rng(1234 + rank)
    imgs = list(rng.integers(0, 256, (B, src_hw[0], src_hw[1], 3), dtype=np.uint8))
    per_img = (src_hw[0] * src_hw[1] * 3 + 15) & ~15
    pinned = torch.empty(per_img * B + 16, dtype=torch.uint8, pin_memory=True)
    _, offsets, hw, total = pack_images(imgs, pinned.numpy())
    src_dev = pinned.cuda(non_blocking=False)
    out_dev = torch.empty((B, cfg["embed_dim"]), dtype=torch.float32, device=f"cuda:{local}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident throughput (`value`)
    for _ in range(args.warmup):
        eng.embed_u8(src_dev, offsets, hw, 0, out=out_dev)
    eng.sync()
    eng.profile_read()                                   # drop records of the warm-up
    launches0 = eng.launches
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    prof_steps = 0
    for i in range(args.steps):
        sampled = i % 4 == 0                             # per-kernel CUDA events on every 4th step of the timed region
        eng.profile(sampled)
        prof_steps += sampled
        eng.embed_u8(src_dev, offsets, hw, 0, out=out_dev)
    ev1.record()
    barrier()
    ms = reduce_max(ev0.elapsed_time(ev1))
    clocks = sampler.stop() if sampler else None
    launches = eng.launches - launches0
    prof = eng.profile_read()
    eng.profile(False)
    eng.sync()

    # ---- end to end through the backend's host-buffer call (`e2e`)
    pinned2 = torch.empty_like(pinned).pin_memory()            # two host batches alternate, as a feeder thread would fill them
    pinned2.copy_(pinned)
    feed = lambda n: ((pinned if i % 2 == 0 else pinned2, total, offsets, hw) for i in range(n))  # noqa: E731
    for emb in be.embed_stream(feed(max(2, args.warmup // 2))):
        pass
    barrier()
    t0 = time.perf_counter()
    ev0.record()
    n_out = 0
    for emb in be.embed_stream(feed(args.steps)):
        n_out += emb.shape[0]
    ev1.record()
    barrier()
    assert n_out == B * args.steps
    e2e_ms = reduce_max(ev0.elapsed_time(ev1))
    e2e_wall_ms = reduce_max((time.perf_counter() - t0) * 1e3)
    assert emb.shape == (B, cfg["embed_dim"]) and np.isfinite(emb).all()

    if rank == 0:
        pk = peaks()
        images = B * world * args.steps
        value = images / (ms / 1e3)
        gemm_ms = prof["gemm"]["ms"]
        gemm_flops = gemm_flops_per_image(cfg) * B * prof_steps          # rank 0's launches on the sampled steps
        achieved = gemm_flops / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else 0.0
        step_ms = ms / args.steps
        share = {k: round(v["ms"] / (ms * prof_steps / args.steps), 4) for k, v in prof.items()}
        cpu_rate, cpu_n, cpu_threads = cpu_reference_rate(cfg, weights, src_hw, budget_s=args.cpu_seconds, max_images=512)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": dict(workload_config(args, cfg, src_hw, world),
                           l2="per-step working set ~0.9 GB (activations + 77 MB input) exceeds the 126 MB L2; no explicit flush",
                           timing="CUDA events on the launch stream, barrier+synchronize both sides, max over ranks"),
            "e2e": {"value": images / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(total), "d2h_bytes_per_step": int(B * cfg["embed_dim"] * 4),
                    "wall_clock_value": images / (e2e_wall_ms / 1e3), "call": "B200ClipBackend.embed_stream (pinned host batches in, host float32 out; the upload of batch i+1 overlaps the encode of batch i)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05, all encoder GEMMs)", "achieved": achieved, "peak": pk["bf16_sustained"],
                         "unit": "TFLOP/s", "frac": achieved / pk["bf16_sustained"], "peak_source": pk["src"] + " (sustained: kernel timed inside a long step)",
                         "traffic": measured_gemm_traffic() if args.model == "ViT-B-32" and B == 512 else None,
                         "traffic_note": "mean dram bytes per launch over the 4 per-layer GEMM shapes, ncu --set full (profiles/step_r1_ncu.csv)",
                         "launches": prof["gemm"]["launches"], "event_sampled_steps": prof_steps, "avg_launch_ms": gemm_ms / max(1, prof["gemm"]["launches"]),
                         "share_of_step": share, "end_to_end_frac_of_tensor_peak": total_flops_per_image(cfg) * value / world / 1e12 / pk["bf16_sustained"]},
            "cpu_baseline": {"value": cpu_rate, "unit": UNIT, "cores": cpu_threads, "kind": "port",
                             "sample": f"{cpu_n} images ({src_hw[0]}x{src_hw[1]}), oracle = reference preprocess closure restated + fp32 tower (torch CPU), batches of 32"},
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    be.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="ViT-B-32", choices=list(VIT_CONFIGS))
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--src", default="224x224", help="source image HxW")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    cfg = VIT_CONFIGS[args.model]
    src_hw = tuple(int(x) for x in args.src.lower().split("x"))
    weights = synthetic_weights(cfg, seed=0)
    if args.impl == "reference":
        run_reference(args, cfg, weights, src_hw)
    else:
        run_b200(args, cfg, weights, src_hw)


if __name__ == "__main__":
    main()
