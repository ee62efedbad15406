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
stamps = []
def feed(n):
    for i in range(n):
        stamps.append(("feed", i, time.perf_counter()))
        yield (pinned, total, off, hw)
list(be.embed_stream(feed(3)))
stamps.clear(); torch.cuda.synchronize(); t0 = time.perf_counter()
for j, e in enumerate(be.embed_stream(feed(8))):
    stamps.append(("out", j, time.perf_counter()))
for k, i, t in stamps: print(k, i, round((t - t0) * 1e3, 2))
# time embed_u8 host duration inside the pipeline
import types
orig = eng.embed_u8
durs = []
def timed_embed(*a, **k):
    t = time.perf_counter(); r = orig(*a, **k); durs.append(round((time.perf_counter() - t) * 1e3, 2)); return r
eng.embed_u8 = timed_embed
list(be.embed_stream(feed(8)))
print("embed_u8 host ms:", durs)
