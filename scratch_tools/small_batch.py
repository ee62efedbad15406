"""Latency of embed_u8 for small batches (the serving regime): CUDA-event time per call and wall time per call."""
import sys, time, numpy as np, torch
sys.path.insert(0, ".")
from lumen_b200 import VIT_CONFIGS, synthetic_weights
from lumen_b200.engine import Engine, pack_images
cfg = VIT_CONFIGS["ViT-B-32"]
eng = Engine(cfg, device=0, max_batch=64); eng.load_weights(synthetic_weights(cfg, 0))
for B in (1, 2, 4, 8, 16, 32, 64):
    imgs = list(np.random.default_rng(1).integers(0, 256, (B, 224, 224, 3), dtype=np.uint8))
    buf, off, hw, total = pack_images(imgs); src = torch.from_numpy(buf[:total]).cuda()
    out = torch.empty((B, 512), dtype=torch.float32, device="cuda")
    for _ in range(10): eng.embed_u8(src, off, hw, 0, out=out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(50): eng.embed_u8(src, off, hw, 0, out=out)
    b.record(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"B={B:3d}  gpu {a.elapsed_time(b) / 50:.3f} ms  host-issue {(t1 - t0) / 50 * 1e3:.3f} ms  wall {(t2 - t0) / 50 * 1e3:.3f} ms  -> {B / (a.elapsed_time(b) / 50) * 1e3:.0f} img/s")
