#!/usr/bin/env python
"""Benchmark of the lumen-clip image path on B200 (BASELINE.json: "CLIP images/sec (preproc+ViT encode) & top-k queries/s at
1/2/4/8 B200").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N ...

Headline (`value`): BASELINE configs[1] - CLIP ViT-B/32, batches of 512 synthetic 224x224 uint8 images per GPU, preprocess + ViT
encode in bf16, data-parallel with replicated weights (weak scaling, no collective).  One STEP = `--batches-per-step` (default 20)
back-to-back passes of the hot path (K1 preprocess -> K2..K9 tower -> L2 normalise) over 512-image batches, so that the K timed
steps last seconds, not milliseconds: the number is taken in the power-capped steady state, after W >= 3 warm-up steps of the same
work.  Everything else BASELINE.json's metric and configs name rides on the same JSON line:

  value / ms_per_step   images/s with the decoded uint8 batches resident in HBM (CUDA events on the launch stream, max over ranks)
  e2e                   the same through `B200ClipBackend.embed_stream`: pinned host uint8 -> H2D -> kernels -> D2H float32 for
                        every batch inside the timed region (upload of batch i+1 overlaps the encode of batch i)
  roofline              the tcgen05 GEMM kernels: algorithmic FLOPs / CUDA-event time per launch, taken in a separate pass AFTER
                        the timed region (so the events perturb nothing); fraction of the burst and of the sustained measured peak
  search                BASELINE configs[4]: 10 M x 512 fp16 gallery row-sharded over the N ranks, 4096 queries, top-100, NCCL
                        all-gather of the per-GPU candidates + merge (strong scaling); local-shard sweep Q in {1,16,64,128} against HBM
  configs               C1 batch-1 latency, C3 ViT-B/16 batch 256 + zero-shot classify over 1000 labels, C4 ViT-L/14@336 batch 256,
                        K1 on real photo sizes (N = 1 only), C2 strong scaling (512 images in total over N GPUs, N > 1)
  cpu_baseline          the oracle (CPU restatement of the reference path) timed on this box's host cores, bounded samples

`--impl reference` times that CPU path alone (the reference's ONNX Runtime path cannot run here: DESIGN.md §8).
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
GALLERY_ROWS, GALLERY_DIM, SEARCH_Q, SEARCH_K = 10_000_000, 512, 4096, 100


def peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d["hbm_gbs"], "bf16_sustained": d["bf16_tflops_sustained"], "bf16_burst": d["bf16_tflops"],
                "sm_mhz_sustained": (d.get("clocks_under_load") or {}).get("sm_mhz_median"), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_sustained": 1400.0, "bf16_burst": 1590.0, "sm_mhz_sustained": 1300.0, "src": "fallback"}


def measured_gemm_traffic():
    """dram__bytes_read+write per GEMM launch from the committed ncu --set full capture (None if absent)."""
    for name in ("gemm_traffic_r2.json", "gemm_traffic_r1.json"):
        try:
            d = json.loads((ROOT / "profiles" / name).read_text())
            return d["dram_bytes_per_launch_mean"], name
        except Exception:  # noqa: BLE001
            continue
    return None, None


def gemm_flops_per_image(cfg: dict) -> float:
    """FLOPs the GEMM kernels EXECUTE per image (the roofline of the dominant kernel is computed from what it did): K is padded
    to 64 in the patch embedding, and the last layer computes keys / values for every token but everything else for the CLS
    row only (lumen_b200/csrc/api.cu:encode_chunk; B200_CLS_LAST=0 turns that off)."""
    n = (cfg["image_size"] // cfg["patch"]) ** 2
    S, d, L, mlp, D, P = n + 1, cfg["width"], cfg["layers"], cfg["mlp_dim"], cfg["embed_dim"], cfg["patch"]
    kpad = (3 * P * P + 63) // 64 * 64
    layer = 2.0 * S * d * 3 * d + 2.0 * S * d * d + 4.0 * S * d * mlp
    last = layer if os.environ.get("B200_CLS_LAST", "1") == "0" else 2.0 * S * d * 2 * d + 2.0 * d * d + 2.0 * d * d + 4.0 * d * mlp
    return 2.0 * n * kpad * d + (L - 1) * layer + last + 2.0 * d * D


def total_flops_per_image(cfg: dict) -> float:
    """ALGORITHMIC FLOPs per image of the full tower (8.8176 G for ViT-B/32), whatever shortcut the kernels take."""
    n = (cfg["image_size"] // cfg["patch"]) ** 2
    S, d, L, mlp, D, P = n + 1, cfg["width"], cfg["layers"], cfg["mlp_dim"], cfg["embed_dim"], cfg["patch"]
    return 2.0 * n * 3 * P * P * d + L * (2.0 * S * d * 3 * d + 4.0 * S * S * d + 2.0 * S * d * d + 4.0 * S * d * mlp) + 2.0 * d * D


class ClockSampler:
    """nvidia-smi clocks / power / throttle reasons sampled every 50 ms while a timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            time.sleep(0.15)                       # let the first sample arrive before the region starts
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.monotonic(), [x.strip() for x in line.split(",")]))

    def mark(self) -> float:
        return time.monotonic()

    def stop(self, t0: float | None = None, t1: float | None = None) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        time.sleep(0.1)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, pw, mx, reasons = [], [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, r in self.rows:
            if (t0 is not None and ts < t0) or (t1 is not None and ts > t1 + 0.05):
                continue                           # only samples taken while the timed region ran
            try:
                sm.append(float(r[0])); mx = float(r[1]); pw.append(float(r[2]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": float(min(sm)) if sm else None, "sm_max_mhz": mx,
                "power_w_median": float(np.median(pw)) if pw else None, "reasons": sorted(reasons), "samples": len(sm), "period_ms": 50}


def workload_config(args, cfg: dict, src_hw, world: int) -> dict:
    """The SAME dict for the GPU arm and the reference arm."""
    return {"workload": f"{args.model} image embedding, BASELINE configs[1]: batches of {args.batch} {src_hw[0]}x{src_hw[1]} uint8 RGB images per GPU -> "
                        f"preprocess (Pillow-exact) + ViT encode (bf16) -> {cfg['embed_dim']}-d unit vectors; one step = {args.batches_per_step} "
                        f"such batches back to back",
            "batch": args.batch, "batches_per_step": args.batches_per_step, "global_batch": args.batch * world,
            "images_per_step": args.batch * args.batches_per_step * world,
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


def cpu_search_rate(budget_s: float, rows: int = 1_000_000, queries: int = 8):
    """The reference's per-request search (BioCLIPModelManager.classify_image, restated in oracle.bioclip_classify_image) on a
    `rows` x 512 fp32 subsample; returns (queries/s on the subsample, queries timed, rows)."""
    import torch
    from oracle import clip_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    g = torch.nn.functional.normalize(torch.randn((rows, GALLERY_DIM), generator=torch.Generator().manual_seed(42)), dim=1).numpy()
    q = torch.nn.functional.normalize(torch.randn((queries, GALLERY_DIM), generator=torch.Generator().manual_seed(43)), dim=1).numpy()
    O.bioclip_classify_image(q[0], g, SEARCH_K)
    done, t0 = 0, time.perf_counter()
    while done < queries and (time.perf_counter() - t0) < budget_s:
        O.bioclip_classify_image(q[done], g, SEARCH_K)
        done += 1
    return done / (time.perf_counter() - t0), done, rows


def run_reference(args, cfg, weights, src_hw):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = max(1, args.gpus)
    per_step = 128                                          # bounded sample of the step's images (the step itself is 10,240 per GPU)
    rates = []
    for _ in range(args.warmup + args.steps):
        r, n, threads = cpu_reference_rate(cfg, weights, src_hw, budget_s=max(3.0, 90.0 / max(1, args.steps + args.warmup)), max_images=per_step)
        rates.append((r, n))
    timed = rates[args.warmup:] or rates
    value = float(np.mean([r for r, _ in timed]))
    sample = (f"{timed[0][1]} images per step, a bounded sample of the step's {args.batch * args.batches_per_step} ({src_hw[0]}x{src_hw[1]} uint8), fp32, "
              f"batches of 32, oracle port of the reference CPU path (preprocess closure + tower) on {threads} host threads")
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * timed[0][1] / value, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args, cfg, src_hw, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "note": "a step of this arm is the bounded sample named in cpu_baseline.sample; ms_per_step is the time that sample took"}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------- GPU arm
class Dist:
    """The little bit of torch.distributed the bench needs (barrier, max over ranks)."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1 and not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.dev = f"cuda:{self.local}"

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce_max(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, iters: int, warm: int = 3) -> float:
        """ms per call: warm-up, barrier + synchronize both sides, CUDA events on the launch stream, max over ranks."""
        torch = self.torch
        for _ in range(warm):
            fn()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        self.barrier()
        return self.reduce_max(e0.elapsed_time(e1)) / iters

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def device_batches(torch, dev, rng, B, src_hw, copies=2):
    """`copies` different packed uint8 batches resident in HBM (+ their pinned host twins for the e2e leg)."""
    from lumen_b200.engine import pack_images
    per_img = (src_hw[0] * src_hw[1] * 3 + 15) & ~15
    out = []
    for _ in range(copies):
        imgs = list(rng.integers(0, 256, (B, src_hw[0], src_hw[1], 3), dtype=np.uint8))
        pinned = torch.empty(per_img * B + 16, dtype=torch.uint8, pin_memory=True)
        _, offsets, hw, total = pack_images(imgs, pinned.numpy())
        out.append({"pinned": pinned, "dev": pinned.to(dev), "offsets": offsets, "hw": hw, "total": total})
    return out


def bench_search(D: Dist, pk: dict, quick: bool) -> dict:
    """configs[4]: the 10 M x 512 gallery row-sharded over the ranks (strong scaling), 4096 queries, top-100."""
    torch = D.torch
    from lumen_b200.engine import Engine
    from lumen_b200.managers import ShardedGallery
    eng = Engine({}, device=D.local)
    rows_total = GALLERY_ROWS // (8 if quick else 1)
    lo, hi = ShardedGallery.shard_bounds(rows_total, D.world, D.rank)
    gen = torch.Generator(device=D.dev).manual_seed(42 + D.rank)
    g = torch.empty((hi - lo, GALLERY_DIM), dtype=torch.float16, device=D.dev)
    for s in range(0, hi - lo, 1 << 20):                    # generated on the device, shard by shard, chunked
        blk = torch.randn((min(1 << 20, hi - lo - s), GALLERY_DIM), device=D.dev, generator=gen)
        g[s:s + blk.shape[0]] = torch.nn.functional.normalize(blk, dim=1).half()
    del blk
    gal = ShardedGallery(eng, g, index_base=lo)
    qgen = torch.Generator(device=D.dev).manual_seed(43)
    q_all = torch.nn.functional.normalize(torch.randn((SEARCH_Q, GALLERY_DIM), device=D.dev, generator=qgen), dim=1).half()
    ms = D.timed(lambda: gal.search(q_all, SEARCH_K), iters=8, warm=3)
    # phases, from events inside the call (a second pass: the first one carries no instrumentation)
    ph = np.zeros(3)
    for _ in range(5):
        ev = []
        gal.search(q_all, SEARCH_K, events=ev)
        torch.cuda.synchronize()
        ph += [ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])]
    ph = [D.reduce_max(float(x) / 5) for x in ph]
    n_local = hi - lo
    flops_gpu = 2.0 * SEARCH_Q * n_local * GALLERY_DIM
    out = {"workload": f"BASELINE configs[4]: {rows_total} x {GALLERY_DIM} fp16 gallery (generated per shard on the device) row-sharded over "
                       f"{D.world} GPU(s), {SEARCH_Q} queries, top-{SEARCH_K}, NCCL all-gather of the per-GPU candidates + merge on every rank",
           "scaling": "strong", "rows_total": rows_total, "rows_per_gpu": n_local, "Q": SEARCH_Q, "k": SEARCH_K, "n_gpus": D.world,
           "ms": ms, "queries_per_s": SEARCH_Q / ms * 1e3,
           "local_topk_ms": ph[0], "allgather_ms": ph[1], "merge_ms": ph[2],
           "roofline": {"bound": "tensor", "tflops_per_gpu": flops_gpu / ms / 1e9, "peak": pk["bf16_burst"], "unit": "TFLOP/s",
                        "frac": flops_gpu / ms / 1e9 / pk["bf16_burst"], "frac_vs_sustained": flops_gpu / ms / 1e9 / pk["bf16_sustained"],
                        "kernel_only_frac": flops_gpu / ph[0] / 1e9 / pk["bf16_burst"],
                        "hbm_frac": (n_local * GALLERY_DIM * 2) / ms / 1e6 / pk["hbm_gbs"],
                        "note": "4,093 flop/B: the tensor pipe binds at Q = 4096; a kernel timed alone for tens of ms -> burst peak"},
           "allgather_bytes_per_gpu": SEARCH_Q * SEARCH_K * 8}
    # small-batch sweep on this rank's shard (no collective): the gallery read binds, graded against HBM
    sweep = []
    for Q in (1, 16, 64, 128):
        q = q_all[:Q].contiguous()
        t = D.timed(lambda: eng.topk(q, g, SEARCH_K, lo), iters=10, warm=3)
        byts = n_local * GALLERY_DIM * 2 + Q * GALLERY_DIM * 2 + Q * SEARCH_K * 12
        sweep.append({"Q": Q, "rows": n_local, "ms": t, "queries_per_s": Q / t * 1e3, "hbm_gbs": byts / t / 1e6, "hbm_frac": byts / t / 1e6 / pk["hbm_gbs"]})
    out["small_batch_sweep"] = {"what": "K10+K12 on this rank's shard, k = 100; algorithmic bytes = rows*D*2 + Q*D*2 + Q*k*12", "peak_gbs": pk["hbm_gbs"], "runs": sweep}
    # batches of a few query tiles (256 .. 1024 queries): between the two roofs, both fractions reported
    mid = []
    for Q in (256, 512, 1024):
        q = q_all[:Q].contiguous()
        t = D.timed(lambda: eng.topk(q, g, SEARCH_K, lo), iters=8, warm=3)
        fl = 2.0 * Q * n_local * GALLERY_DIM
        mid.append({"Q": Q, "rows": n_local, "ms": t, "queries_per_s": Q / t * 1e3, "tensor_frac_of_burst": fl / t / 1e9 / pk["bf16_burst"],
                    "hbm_frac": (n_local * GALLERY_DIM * 2) / t / 1e6 / pk["hbm_gbs"]})
    out["mid_batch_sweep"] = {"what": "K10+K12 on this rank's shard, k = 100, a few query tiles", "runs": mid}
    eng.close()
    del g, gal
    torch.cuda.empty_cache()
    return out


def bench_tower(D: Dist, pk: dict, name: str, B: int, iters: int) -> tuple[dict, object]:
    torch = D.torch
    from lumen_b200.engine import Engine
    cfg = VIT_CONFIGS[name]
    eng = Engine(cfg, device=D.local, max_batch=B)
    eng.load_weights(synthetic_weights(cfg, 0))
    S = cfg["image_size"]
    bt = device_batches(torch, D.dev, np.random.default_rng(1), B, (S, S), copies=1)[0]
    out = torch.empty((B, cfg["embed_dim"]), dtype=torch.float32, device=D.dev)
    # two batches in flight, like the headline: engine and twin on two streams that fork from / join the timed stream
    engs, outs, side = [eng, eng.twin()], [out, torch.empty_like(out)], [torch.cuda.Stream(device=D.dev), torch.cuda.Stream(device=D.dev)]

    def pair():
        cur = torch.cuda.current_stream(D.dev)
        for k in range(2):
            side[k].wait_stream(cur)
            with torch.cuda.stream(side[k]):
                engs[k].embed_u8(bt["dev"], bt["offsets"], bt["hw"], 0, out=outs[k])
        for k in range(2):
            cur.wait_stream(side[k])

    ms = D.timed(pair, iters=max(2, iters // 2), warm=2) / 2
    ms_single = D.timed(lambda: eng.embed_u8(bt["dev"], bt["offsets"], bt["hw"], 0, out=out), iters=max(2, iters // 2), warm=1)
    ips = B / ms * 1e3
    tf = total_flops_per_image(cfg) * ips / 1e12
    return ({"model": name, "batch": B, "image_size": S, "ms_per_batch": ms, "ms_per_batch_one_in_flight": ms_single, "images_per_s": ips,
             "gflop_per_image": total_flops_per_image(cfg) / 1e9, "tensor_tflops": tf, "frac_of_burst_peak": tf / pk["bf16_burst"],
             "frac_of_sustained_peak": tf / pk["bf16_sustained"]}, (eng, out))


def bench_configs(D: Dist, pk: dict, be, batches, args) -> dict:
    """The other configs BASELINE.json names, on one GPU."""
    torch = D.torch
    out = {}
    eng = be.engine
    # C1: ViT-B/32, batch 1 (the reference's own CPU-runnable case): device-resident and through the host-buffer call
    one = device_batches(torch, D.dev, np.random.default_rng(3), 1, (224, 224), copies=1)[0]
    o1 = torch.empty((1, 512), dtype=torch.float32, device=D.dev)
    ms = D.timed(lambda: eng.embed_u8(one["dev"], one["offsets"], one["hw"], 0, out=o1), iters=100, warm=10)
    t0 = time.perf_counter()
    for _ in range(50):
        be.embed_packed(one["pinned"], one["total"], one["offsets"], one["hw"])
    host_ms = (time.perf_counter() - t0) / 50 * 1e3
    out["c1_vit_b32_batch1"] = {"ms_per_image_device_resident": ms, "ms_per_image_host_to_host": host_ms,
                                "floor_ms": 176e6 / (pk["hbm_gbs"] * 1e9) * 1e3, "note": "weight-streaming / kernel-latency bound: 89 kernels, 176 MB of bf16 weights"}
    # C3: ViT-B/16 batch 256 + zero-shot classify over 1000 labels
    c3, (e3, emb3) = bench_tower(D, pk, "ViT-B-16", 256, iters=8)
    e3.set_labels(np.random.default_rng(7).standard_normal((1000, 512)).astype(np.float32))
    c3["classify_us"] = D.timed(lambda: e3.classify(emb3, 5), iters=50, warm=5) * 1e3
    c3["classify_note"] = "256 images x 1000 labels x 512, softmax(100 s), top-5: 262 MFLOP, launch-latency bound"
    out["c3_vit_b16_batch256_classify1000"] = c3
    e3.close()
    # C4: ViT-L/14 at 336, batch 256
    c4, (e4, _) = bench_tower(D, pk, "ViT-L-14-336", 256, iters=4)
    out["c4_vit_l14_336_batch256"] = c4
    e4.close()
    torch.cuda.empty_cache()
    # K1 alone on photo-sized sources (the headline's 224x224 sources take the identity path)
    import ctypes as C
    from lumen_b200 import _lib
    from lumen_b200.engine import pack_images
    k1 = []
    for (h, w), B in (((224, 224), 2048), ((480, 640), 512), ((1080, 1920), 128)):
        img = np.random.default_rng(5).integers(0, 256, (h, w, 3), dtype=np.uint8)
        buf, off, hw, total = pack_images([img] * B)
        src = torch.from_numpy(buf[:total]).to(D.dev)
        dst = torch.zeros((B * 49, 3072), dtype=torch.bfloat16, device=D.dev)

        def run():
            _lib.check(eng.lib.b200clip_preprocess(eng._h, C.c_void_p(src.data_ptr()), off.ctypes.data_as(C.c_void_p), hw.ctypes.data_as(C.c_void_p),
                                                   B, 0, 1, 1, C.c_void_p(dst.data_ptr()), eng._stream()))
        t = D.timed(run, iters=20, warm=3)
        byts = B * (h * w * 3 + 3 * 224 * 224 * 2)
        k1.append({"src": f"{h}x{w}", "batch": B, "ms": t, "images_per_s": B / t * 1e3, "hbm_gbs": byts / t / 1e6, "hbm_frac": byts / t / 1e6 / pk["hbm_gbs"]})
        del src, dst
    out["k1_preprocess"] = {"what": "K1 alone -> 224 patch-major bf16; algorithmic bytes = H*W*3 + 3*224*224*2 per image", "peak_gbs": pk["hbm_gbs"], "runs": k1}
    return out


def run_b200(args, cfg, weights, src_hw):
    D = Dist()
    torch = D.torch
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    world, rank, B, R = D.world, D.rank, args.batch, args.batches_per_step
    pk = peaks()
    be = B200ClipBackend(SyntheticResources(model_name=args.model), device_preference=D.dev, max_batch_size=B, weights=weights)
    be.initialize()
    eng = be.engine
    batches = device_batches(torch, D.dev, np.random.default_rng(1234 + rank), B, src_hw, copies=2)
    out_dev = torch.empty((B, cfg["embed_dim"]), dtype=torch.float32, device=D.dev)
    # two batches in flight, as B200ClipBackend.embed_stream runs them: even batches on the engine, odd ones on its twin (same weights,
    # own activations; b200clip_clone), each on a stream of its own, so that one batch's memory-bound kernels run under the other's GEMMs
    twin = not args.single_stream
    engs = [eng, eng.twin()] if twin else [eng, eng]
    outs = [out_dev, torch.empty_like(out_dev)] if twin else [out_dev, out_dev]
    side = [torch.cuda.Stream(device=D.dev), torch.cuda.Stream(device=D.dev)] if twin else None

    def step():
        if not twin:
            for j in range(R):
                bt = batches[j & 1]
                eng.embed_u8(bt["dev"], bt["offsets"], bt["hw"], 0, out=out_dev)
            return
        cur = torch.cuda.current_stream(D.dev)
        for st in side:
            st.wait_stream(cur)
        for j in range(R):
            k = j & 1
            bt = batches[k]
            with torch.cuda.stream(side[k]):
                engs[k].embed_u8(bt["dev"], bt["offsets"], bt["hw"], 0, out=outs[k])
        for st in side:
            cur.wait_stream(st)

    # ---- device-resident throughput (`value`): W warm-up steps, then K timed steps; no instrumentation inside
    for _ in range(args.warmup):
        step()
    eng.sync()
    launches0 = eng.launches + (engs[1].launches if twin else 0)
    sampler = ClockSampler(D.local) if rank == 0 else None
    D.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_a = time.monotonic()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    D.barrier()
    t_b = time.monotonic()
    ms = D.reduce_max(ev0.elapsed_time(ev1))
    clocks = sampler.stop(t_a, t_b) if sampler else None
    launches = eng.launches + (engs[1].launches if twin else 0) - launches0
    eng.sync()

    # ---- roofline pass (one engine, one stream: every kernel timed on its own), after the timed region: CUDA events around every kernel launch of two more steps
    eng.profile_read()
    eng.profile(True)
    prof_batches = 0
    for _ in range(2):
        for j in range(R):
            bt = batches[j & 1]
            eng.embed_u8(bt["dev"], bt["offsets"], bt["hw"], 0, out=out_dev)
        prof_batches += R
    prof = eng.profile_read()
    eng.profile(False)
    eng.sync()

    # ---- end to end through the backend's host-buffer call (`e2e`): same step, every batch crosses PCIe both ways
    feed = lambda n: ((batches[i & 1]["pinned"], batches[i & 1]["total"], batches[i & 1]["offsets"], batches[i & 1]["hw"]) for i in range(n))  # noqa: E731
    for emb in be.embed_stream(feed(max(2, min(args.warmup, 3) * R))):
        pass
    D.barrier()
    t0 = time.perf_counter()
    ev0.record()
    n_out = 0
    for emb in be.embed_stream(feed(args.steps * R)):
        n_out += emb.shape[0]
    ev1.record()
    D.barrier()
    assert n_out == B * R * args.steps
    e2e_ms = D.reduce_max(ev0.elapsed_time(ev1))
    e2e_wall_ms = D.reduce_max((time.perf_counter() - t0) * 1e3)
    assert emb.shape == (B, cfg["embed_dim"]) and np.isfinite(emb).all()

    # ---- C2 strong scaling (512 images in total over the ranks), the search, and on one GPU the remaining configs
    extra = {}
    if world > 1 and B % world == 0:
        sb = B // world
        sub = {k: (v[:sb] if k in ("offsets", "hw") else v) for k, v in batches[0].items()}
        so = out_dev[:sb]
        # long enough (a few hundred ms after as much warm-up) to sit in the steady state of THIS workload, whatever ran before it
        t = D.timed(lambda: eng.embed_u8(sub["dev"], sub["offsets"], sub["hw"], 0, out=so), iters=160, warm=60)
        extra["c2_strong_scaling"] = {"images_total": B, "images_per_gpu": sb, "ms_per_batch": t, "images_per_s": B / t * 1e3,
                                      "token_rows_per_gpu": sb * ((cfg["image_size"] // cfg["patch"]) ** 2 + 1)}
    total_bytes = int(batches[0]["total"])
    if not args.skip_configs and world == 1:
        extra.update(bench_configs(D, pk, be, batches, args))
    be.close()
    del batches, out_dev
    torch.cuda.empty_cache()
    search = None if args.skip_search else bench_search(D, pk, args.quick)

    if rank == 0:
        images = B * R * world * args.steps
        value = images / (ms / 1e3)
        step_ms = ms / args.steps
        gemm_ms, gemm_n = prof["gemm"]["ms"], max(1, prof["gemm"]["launches"])
        achieved = gemm_flops_per_image(cfg) * B * prof_batches / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else 0.0
        ev_total = sum(v["ms"] for v in prof.values())
        share = {k: round(v["ms"] / ev_total, 4) for k, v in prof.items()} if ev_total > 0 else {}
        # which measured peak the timed region ran under: SM clocks near the maximum -> burst, pulled down by the power cap -> sustained
        sm, smx = (clocks or {}).get("sm_mhz"), (clocks or {}).get("sm_max_mhz")
        regime = "sustained" if (sm and smx and sm < 0.93 * smx) else "burst"
        peak = pk["bf16_sustained"] if regime == "sustained" else pk["bf16_burst"]
        traffic, traffic_src = measured_gemm_traffic()
        e2e_tf = total_flops_per_image(cfg) * value / world / 1e12
        cpu_rate, cpu_n, cpu_threads = cpu_reference_rate(cfg, weights, src_hw, budget_s=args.cpu_seconds, max_images=512)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args, cfg, src_hw, world),
            "notes": {"l2": "per-batch working set ~0.9 GB (activations + 77 MB input, two input batches alternate) exceeds the 126 MB L2; no explicit flush",
                      "timing": "CUDA events on the launch stream, barrier+synchronize both sides, max over ranks; no per-kernel events inside the timed region",
                      "in_flight": ("two batches at a time (engine + twin sharing the weights, two streams that fork from / join the timed stream every step)"
                                    if twin else "one batch at a time")},
            "e2e": {"value": images / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": total_bytes * R, "d2h_bytes_per_step": int(B * cfg["embed_dim"] * 4) * R,
                    "wall_clock_value": images / (e2e_wall_ms / 1e3),
                    "call": "B200ClipBackend.embed_stream (pinned host batches in, host float32 out; the upload of batch i+1 overlaps the encode of batch i, two batches in flight on engine + twin)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "kernel": "gemm_tc2_kernel / gemm_tc_kernel (tcgen05, all encoder GEMMs)", "achieved": achieved, "peak": peak,
                         "unit": "TFLOP/s", "frac": achieved / peak, "regime": regime,
                         "peak_source": f"{pk['src']} ({regime}: timed region {ms / 1e3:.2f} s, SM clock median {sm} of {smx} MHz)",
                         "frac_vs_burst": achieved / pk["bf16_burst"], "frac_vs_sustained": achieved / pk["bf16_sustained"],
                         "traffic": traffic if args.model == "ViT-B-32" and B == 512 else None,
                         "traffic_note": f"mean dram bytes per launch over the per-layer GEMM shapes, ncu --set full (profiles/{traffic_src})",
                         "launches": prof["gemm"]["launches"], "launches_per_step": prof["gemm"]["launches"] * R // max(1, prof_batches),
                         "avg_launch_ms": gemm_ms / gemm_n, "measured": f"separate pass of {prof_batches} batches after the timed region, events around every launch",
                         "share_of_step": share,
                         "whole_step": {"tensor_tflops": e2e_tf, "frac_vs_burst": e2e_tf / pk["bf16_burst"], "frac_vs_sustained": e2e_tf / pk["bf16_sustained"]}},
            "cpu_baseline": {"value": cpu_rate, "unit": UNIT, "cores": cpu_threads, "kind": "port",
                             "sample": f"{cpu_n} images ({src_hw[0]}x{src_hw[1]}), oracle = reference preprocess closure restated + fp32 tower (torch CPU), batches of 32"},
            "clocks": clocks,
        }
        if search is not None:
            if world == 1 and not args.quick:
                qps, nq, rows = cpu_search_rate(budget_s=max(6.0, args.cpu_seconds))
                search["cpu_baseline"] = {"value": qps * rows / search["rows_total"], "unit": "queries/s", "cores": os.cpu_count(), "kind": "port", "extrapolated": True,
                                          "sample": f"{nq} queries, each the whole of BioCLIPModelManager.classify_image (row norms, dot product, full argsort) on a "
                                                    f"{rows} x 512 fp32 subsample: {qps:.2f} queries/s there, scaled linearly to {search['rows_total']} rows"}
            line["search"] = search
        if extra:
            line["configs"] = extra
        print(json.dumps(line), flush=True)
    D.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="ViT-B-32", choices=list(VIT_CONFIGS))
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--batches-per-step", type=int, default=20, help="512-image batches per step (a step of ~0.1 s keeps 20 steps in the steady state)")
    ap.add_argument("--src", default="224x224", help="source image HxW")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--single-stream", action="store_true", help="one batch in flight instead of two (A/B against the default)")
    ap.add_argument("--skip-search", action="store_true", help="leave the gallery-search block out (development runs)")
    ap.add_argument("--skip-configs", action="store_true", help="leave the C1/C3/C4/K1 block out (development runs)")
    ap.add_argument("--quick", action="store_true", help="1.25 M-row gallery instead of 10 M (development runs; the line says so)")
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
