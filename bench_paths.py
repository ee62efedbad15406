#!/usr/bin/env python
"""Secondary measurements for the other SURVEY.md §8 rows (bench.py stays the headline contract):

    python bench_paths.py preprocess   K1 at 224x224 / 640x480 / 1920x1080 -> 224 against the HBM roofline
    python bench_paths.py search       K10+K12: Q sweep over an N x 512 fp16 gallery, k = 100 (both rooflines)
    python bench_paths.py classify     K11: 256 images x 1000 labels (C3), microseconds
    python bench_paths.py models       ViT-B/16 (C3) and ViT-L/14@336 (C4) embedding throughput
    torchrun --nproc-per-node G bench_paths.py search --sharded   C5: gallery row-sharded over G GPUs, NCCL all-gather + merge

Each line printed is JSON.  CUDA events on the launch stream; every timed loop is preceded by >= 3 warm-up runs; inputs
larger than L2 unless noted."""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from bench import gemm_flops_per_image, peaks, total_flops_per_image  # noqa: E402
from lumen_b200.vit_config import VIT_CONFIGS, synthetic_weights  # noqa: E402
from lumen_b200._lib import check as _lib_check  # noqa: E402


def timed(fn, iters, warm=3):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench_preprocess(args):
    import torch
    from lumen_b200.engine import Engine, pack_images
    pk = peaks()
    eng = Engine({"patch": 32}, device=0, image_size=224)
    rng = np.random.default_rng(1234)
    for (h, w), B in (((224, 224), 2048), ((480, 640), 512), ((1080, 1920), 128)):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        imgs = [img] * B                                   # same pixels, distinct copies in the packed buffer
        buf, off, hw, total = pack_images(imgs)
        src = torch.from_numpy(buf[:total]).cuda()
        out = torch.zeros((B * 49, 3072), dtype=torch.bfloat16, device="cuda")
        import ctypes as C
        from lumen_b200 import _lib

        def run():
            _lib.check(eng.lib.b200clip_preprocess(eng._h, C.c_void_p(src.data_ptr()), off.ctypes.data_as(C.c_void_p),
                                                   hw.ctypes.data_as(C.c_void_p), B, 0, 1, 1, C.c_void_p(out.data_ptr()), eng._stream()))
        ms = timed(run, 20)
        bytes_alg = B * (h * w * 3 + 3 * 224 * 224 * 2)
        gbs = bytes_alg / ms / 1e6
        print(json.dumps({"path": "K1 preprocess", "src": f"{h}x{w}", "batch": B, "ms": ms, "images_per_s": B / ms * 1e3,
                          "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                                       "algorithmic_bytes_per_image": h * w * 3 + 3 * 224 * 224 * 2}}), flush=True)
    eng.close()


def bench_search(args):
    import torch
    import torch.distributed as dist
    from lumen_b200.engine import Engine
    from lumen_b200.managers import ShardedGallery
    pk = peaks()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine({}, device=local)
    N_total, D, k = args.gallery, 512, 100
    lo, hi = ShardedGallery.shard_bounds(N_total, world, rank)
    gen = torch.Generator(device="cuda").manual_seed(42 + rank)
    g = torch.empty((hi - lo, D), dtype=torch.float16, device="cuda")
    for s in range(0, hi - lo, 1 << 20):                   # generate the shard on the device, chunked
        blk = torch.randn((min(1 << 20, hi - lo - s), D), device="cuda", generator=gen)
        g[s:s + blk.shape[0]] = torch.nn.functional.normalize(blk, dim=1).half()
    gal = ShardedGallery(eng, g, index_base=lo)
    qs = [int(x) for x in args.qs.split(",")] if args.qs else ([4096] if args.sharded else [1, 16, 64, 128, 512, 4096])
    for Q in qs:
        q = torch.nn.functional.normalize(torch.randn((Q, D), device="cuda", generator=torch.Generator(device="cuda").manual_seed(43)), dim=1).half()
        ms = timed(lambda: gal.search(q, k), 5 if Q >= 512 else 10)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        if rank == 0:
            n_local = hi - lo
            flops = 2.0 * Q * n_local * D
            byts = n_local * D * 2 + Q * D * 2 + Q * k * 12
            tf, gbs = flops / ms / 1e9, byts / ms / 1e6
            print(json.dumps({"path": "K10+K12 gallery search", "Q": Q, "N_total": N_total, "N_per_gpu": n_local, "k": k, "n_gpus": world, "ms": ms,
                              "queries_per_s": Q / ms * 1e3,
                              "roofline": {"bound": "tensor" if Q > 222 else "hbm", "tensor_tflops": tf, "tensor_frac": tf / pk["bf16_burst"],
                                           "hbm_gbs": gbs, "hbm_frac": gbs / pk["hbm_gbs"], "peaks": pk}}), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def bench_classify(args):
    import torch
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    eng.set_labels(np.random.default_rng(7).standard_normal((1000, 512)).astype(np.float32))
    q = torch.nn.functional.normalize(torch.randn((256, 512), device="cuda"), dim=1)
    ms = timed(lambda: eng.classify(q, 5), 50)
    print(json.dumps({"path": "K11 classify (C3 numerics)", "B": 256, "labels": 1000, "D": 512, "k": 5, "us": ms * 1e3,
                      "note": "launch-latency bound: 262 MFLOP, 2 MB of labels"}), flush=True)
    eng.close()


def bench_models(args):
    import torch
    from lumen_b200.engine import Engine, pack_images
    pk = peaks()
    for name, B in (("ViT-B-16", 256), ("ViT-L-14", 256), ("ViT-L-14-336", 256), ("ViT-B-32", 1)):
        cfg = VIT_CONFIGS[name]
        eng = Engine(cfg, device=0, max_batch=B)
        eng.load_weights(synthetic_weights(cfg, 0))
        S = cfg["image_size"]
        imgs = list(np.random.default_rng(1).integers(0, 256, (B, S, S, 3), dtype=np.uint8))
        buf, off, hw, total = pack_images(imgs)
        src = torch.from_numpy(buf[:total]).cuda()
        out = torch.empty((B, cfg["embed_dim"]), dtype=torch.float32, device="cuda")
        ms = timed(lambda: eng.embed_u8(src, off, hw, 0, out=out), 5 if B > 1 else 50)
        ips = B / ms * 1e3
        print(json.dumps({"path": "embed (K1..K9)", "model": name, "batch": B, "ms": ms, "images_per_s": ips,
                          "frac_of_tensor_peak_sustained": total_flops_per_image(cfg) * ips / 1e12 / pk["bf16_sustained"],
                          "gflop_per_image": total_flops_per_image(cfg) / 1e9}), flush=True)
        eng.close()


def bench_gemm(args):
    """Every encoder GEMM shape of ViT-B/32 at batch 512 through b200clip_gemm_bf16, per epilogue: a burst figure (20 launches) and
    a steady-state one (a ~0.4 s loop after a warm-up loop, the regime a long pass runs in), the latter next to cuBLASLt on the
    same operands (torch F.linear: bias epilogue, bf16 output - no activation, no residual) as a yardstick of what the shape allows."""
    import torch
    from lumen_b200 import _lib
    from lumen_b200.engine import Engine
    pk = peaks()
    eng = Engine({}, device=0)

    def loop(fn, iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(iters): fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / iters

    shapes = [("patch", 25088, 768, 3072, _lib.EPI_F32), ("qkv", 25600, 2304, 768, _lib.EPI_BIAS_BF16), ("out_proj", 25600, 768, 768, _lib.EPI_BIAS_RES_F32),
              ("fc1", 25600, 3072, 768, _lib.EPI_BIAS_ACT_BF16), ("fc2", 25600, 768, 3072, _lib.EPI_BIAS_RES_F32), ("qkv_f32", 25600, 2304, 768, _lib.EPI_F32),
              ("out_proj_bf16", 25600, 768, 768, _lib.EPI_BIAS_BF16), ("fc2_bf16", 25600, 768, 3072, _lib.EPI_BIAS_BF16)]
    for name, M, N, K, epi in shapes:
        A = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        W = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        biasb = bias.to(torch.bfloat16)
        out = torch.zeros((M, N), dtype=torch.float32 if epi in (_lib.EPI_F32, _lib.EPI_BIAS_RES_F32) else torch.bfloat16, device="cuda")
        mine = lambda: eng.gemm(A, W, bias, epi, 0, out=out)
        ms = timed(mine, 20)
        fl = 2.0 * M * N * K
        iters = max(200, int(400 / (fl / 1.2e12)))
        loop(mine, iters)
        ms_s = loop(mine, iters)
        cub = lambda: torch.nn.functional.linear(A, W, biasb)
        loop(cub, iters // 2)
        ms_c = loop(cub, iters)
        print(json.dumps({"path": "gemm_tc_kernel", "shape": name, "M": M, "N": N, "K": K, "epilogue": epi, "us": ms * 1e3, "tflops": fl / ms / 1e9,
                          "frac_burst": fl / ms / 1e9 / pk["bf16_burst"], "steady_us": ms_s * 1e3, "steady_tflops": fl / ms_s / 1e9,
                          "frac_sustained": fl / ms_s / 1e9 / pk["bf16_sustained"], "cublaslt_bias_bf16_steady_tflops": fl / ms_c / 1e9,
                          "vs_cublaslt": ms_c / ms_s}), flush=True)
    eng.close()


def bench_attention(args):
    """K5 alone through b200clip_attention on the three towers' shapes (the kernel is picked by shape)."""
    import torch
    from lumen_b200.engine import Engine
    pk = peaks()
    eng = Engine({}, device=0)
    for name, B, S, heads in (("ViT-B-32", 512, 50, 12), ("ViT-B-16", 256, 197, 12), ("ViT-L-14", 256, 257, 16), ("ViT-L-14-336", 128, 577, 16)):
        d = heads * 64
        qkv = torch.randn(B * S, 3 * d, device="cuda").to(torch.bfloat16)
        out = torch.empty((B * S, d), dtype=torch.bfloat16, device="cuda")
        ms = timed(lambda: eng.attention(qkv, B, S, d, heads, out=out), 20)
        fl, ex = 4.0 * B * heads * S * S * 64, float(B) * heads * S * S
        byt = B * S * 4 * d * 2
        print(json.dumps({"path": "K5 attention", "model": name, "B": B, "S": S, "heads": heads, "us": ms * 1e3,
                          "kernel": "attention_tc_kernel (tcgen05)" if S > 64 else "attention_short_kernel (tcgen05)",
                          "tflops_useful": fl / ms / 1e9, "gexp_per_s": ex / ms / 1e6,
                          "frac_of_sfu_exp2_peak": ex / ms / 1e6 / (148 * 16 * 1.92), "hbm_gbs": byt / ms / 1e6,
                          "hbm_frac": byt / ms / 1e6 / pk["hbm_gbs"]}), flush=True)
    eng.close()


def bench_decode(args):
    """P0 / N2: JPEG bytes -> device RGB (host-thread Huffman + device IDCT / upsampling / colour), against Pillow on one core."""
    import io
    import time
    import torch
    from PIL import Image
    from lumen_b200.engine import Engine
    pk = peaks()
    eng = Engine({}, device=0)
    rng = np.random.default_rng(0)
    for (h, w, B) in ((480, 640, 128), (1080, 1920, 32)):
        datas = []
        for i in range(B):
            yy, xx = np.mgrid[0:h, 0:w]
            a = np.stack([128 + 90 * np.sin(xx / (20.0 + i)) + rng.normal(0, 10, (h, w)), 128 + 90 * np.cos(yy / (13.0 + i)) + rng.normal(0, 10, (h, w)),
                          128 + 60 * np.sin((xx + yy) / 23.0) + rng.normal(0, 10, (h, w))], -1).clip(0, 255).astype(np.uint8)
            buf = io.BytesIO(); Image.fromarray(a).save(buf, "JPEG", quality=85, subsampling=2); datas.append(buf.getvalue())
        t0 = time.perf_counter()
        for d in datas[:8]: np.asarray(Image.open(io.BytesIO(d)).convert("RGB"))
        pil_ms = (time.perf_counter() - t0) / 8 * 1e3
        for threads in (1, 0):
            def run():
                eng.decode_jpeg(datas, host_threads=threads)
            for _ in range(2): run()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5): run()
            torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) / 5 * 1e3
            print(json.dumps({"path": "P0 jpeg decode", "src": f"{w}x{h} 4:2:0 q85", "batch": B, "host_threads": threads or os.cpu_count(),
                              "ms": ms, "images_per_s": B / ms * 1e3, "mpix_per_s": B * h * w / ms / 1e3, "jpeg_mb": sum(map(len, datas)) / 1e6,
                              "pillow_one_core_images_per_s": 1e3 / pil_ms, "timing": "wall clock around the call (host Huffman + H2D + 2 kernels)"}), flush=True)
        if (h, w) == (480, 640):
            # bytes -> embeddings (image_batch_to_vectors): device decode + K1 + ViT-B/32, against the reference's order of work (PIL decode on one core)
            cfg = VIT_CONFIGS["ViT-B-32"]
            e2 = Engine(cfg, device=0, max_batch=B)
            e2.load_weights(synthetic_weights(cfg, 0))
            for _ in range(2): e2.embed_bytes(datas)
            t0 = time.perf_counter()
            for _ in range(5): e2.embed_bytes(datas)
            ms = (time.perf_counter() - t0) / 5 * 1e3
            print(json.dumps({"path": "P0+K1..K9 jpeg bytes -> embeddings", "model": "ViT-B-32", "src": f"{w}x{h} 4:2:0 q85", "batch": B, "ms": ms,
                              "images_per_s": B / ms * 1e3, "host_threads": os.cpu_count(),
                              "note": "host-bound: Huffman decoding on the host cores; the device part of this batch takes about 2 ms"}), flush=True)
            e2.close()
    eng.close()


def bench_n4(args):
    """N4: text tower (label-embedding bootstrap) and gallery store (fp32 rows -> unit-norm fp16 rows on the device)."""
    import tempfile
    import time
    import torch
    from lumen_b200.engine import Engine
    from lumen_b200.gallery import DeviceGallery
    from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_text_weights, synthetic_token_ids
    pk = peaks()
    cfg = TEXT_CONFIGS["ViT-B-32"]
    d, L, mlp, D, S = cfg["width"], cfg["layers"], cfg["mlp_dim"], cfg["embed_dim"], cfg["context_length"]
    flops_per_text = L * (2 * S * d * 3 * d + 2 * S * d * d + 4 * S * d * mlp + 4 * S * S * d) + 2 * d * D
    eng = Engine({}, device=0)
    eng.init_text(cfg, max_batch=1024)
    eng.load_weights(synthetic_text_weights(cfg, 0))
    for B in (1, 64, 1024):
        ids = torch.from_numpy(synthetic_token_ids(cfg, B, seed=1)).cuda()
        out = torch.empty((B, D), dtype=torch.float32, device="cuda")
        def run():
            _lib_check(eng.lib.b200clip_encode_text(eng._h, ctypes.c_void_p(ids.data_ptr()), B, ctypes.c_void_p(out.data_ptr()), eng._stream()))
        ms = timed(run, 10 if B >= 64 else 50)
        tps = B / ms * 1e3
        print(json.dumps({"path": "N4 text tower (encode_text)", "model": "ViT-B-32 text", "batch": B, "ms": ms, "texts_per_s": tps,
                          "gflop_per_text": flops_per_text / 1e9, "frac_of_tensor_peak_sustained": flops_per_text * tps / 1e12 / pk["bf16_sustained"]}), flush=True)
    # gallery pack: HBM-bound, 4 B in + 2 B out per element
    N, Dg = 2_000_000, 512
    src = torch.randn((N, Dg), device="cuda")
    g = DeviceGallery(eng, Dg, capacity=N)
    def pack():
        g.rows = 0
        g.append(src, normalize=True)
    ms = timed(pack, 10)
    gbs = N * Dg * 6 / ms / 1e6
    print(json.dumps({"path": "N4 gallery_pack (device fp32 -> unit-norm fp16)", "rows": N, "dim": Dg, "ms": ms, "rows_per_s": N / ms * 1e3,
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"]}}), flush=True)
    # the whole path from the reference's file format: fp32 .npy (mmap) -> pinned staging -> device shard
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "g.npy")
        host = np.random.default_rng(0).standard_normal((500_000, Dg)).astype(np.float32)
        np.save(path, host)
        DeviceGallery.from_npy(eng, path); torch.cuda.synchronize()
        t0 = time.perf_counter()
        DeviceGallery.from_npy(eng, path); torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"path": "N4 gallery from_npy (fp32 .npy mmap -> device fp16)", "rows": host.shape[0], "dim": Dg, "ms": ms,
                          "host_gbs": host.nbytes / ms / 1e6, "note": "bound by the host copy into pinned staging and PCIe"}), flush=True)
    eng.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["preprocess", "search", "classify", "models", "gemm", "attention", "decode", "n4"])
    ap.add_argument("--gallery", type=int, default=1_250_000, help="total gallery rows (all GPUs)")
    ap.add_argument("--sharded", action="store_true")
    ap.add_argument("--qs", default="", help="comma-separated query counts for the search sweep")
    a = ap.parse_args()
    {"preprocess": bench_preprocess, "search": bench_search, "classify": bench_classify, "models": bench_models, "gemm": bench_gemm,
     "attention": bench_attention, "decode": bench_decode, "n4": bench_n4}[a.what](a)
