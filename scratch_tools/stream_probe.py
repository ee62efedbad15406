import sys, time, numpy as np, torch
sys.path.insert(0, '/root/repo')
from lumen_b200 import VIT_CONFIGS, synthetic_weights
from lumen_b200.backend import B200ClipBackend, SyntheticResources
from lumen_b200.engine import pack_images
cfg = VIT_CONFIGS["ViT-B-32"]
be = B200ClipBackend(SyntheticResources(model_name="ViT-B-32"), device_preference="cuda:0", max_batch_size=512, weights=synthetic_weights(cfg, 0))
be.initialize(); eng = be.engine
B = 512
imgs = list(np.random.default_rng(0).integers(0, 256, (B, 224, 224, 3), dtype=np.uint8))
per = (224 * 224 * 3 + 15) & ~15
pinned = torch.empty(per * B + 16, dtype=torch.uint8, pin_memory=True)
_, off, hw, total = pack_images(imgs, pinned.numpy())
dev = pinned.cuda(); out = torch.empty((B, 512), device="cuda")
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("compute only ms", t(lambda: eng.embed_u8(dev, off, hw, 0, out=out)))
print("h2d only ms", t(lambda: dev.copy_(pinned, non_blocking=True)))
t0 = time.perf_counter(); eng.embed_u8(dev, off, hw, 0, out=out); print("host enqueue ms", (time.perf_counter() - t0) * 1e3); torch.cuda.synchronize()
feed = lambda n: ((pinned, total, off, hw) for _ in range(n))
list(be.embed_stream(feed(3)))
torch.cuda.synchronize(); t0 = time.perf_counter(); list(be.embed_stream(feed(20))); torch.cuda.synchronize()
print("stream ms/batch", (time.perf_counter() - t0) / 20 * 1e3)
# direct overlap probe: kernels on stream A, H2D on stream B, no cross dependencies
A, Bs = torch.cuda.Stream(), torch.cuda.Stream()
dev2 = torch.empty_like(dev)
def both():
    with torch.cuda.stream(Bs):
        dev2.copy_(pinned, non_blocking=True)
    with torch.cuda.stream(A):
        eng.embed_u8(dev, off, hw, 0, out=out)
print("independent copy + compute ms", t(both))
def onlyA():
    with torch.cuda.stream(A):
        eng.embed_u8(dev, off, hw, 0, out=out)
print("compute on stream A ms", t(onlyA))
