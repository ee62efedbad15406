"""A/B of an environment switch on one box: python ab_models.py VAR model batch  (runs VAR=1 / VAR=0 alternately in subprocesses)."""
import json, os, subprocess, sys
var, model, batch = sys.argv[1], sys.argv[2], sys.argv[3]
code = f"""
import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from lumen_b200 import VIT_CONFIGS, synthetic_weights
from lumen_b200.engine import Engine, pack_images
cfg = VIT_CONFIGS['{model}']; B = {batch}
eng = Engine(cfg, device=0, max_batch=B); eng.load_weights(synthetic_weights(cfg, 0))
S = cfg['image_size']
imgs = list(np.random.default_rng(1).integers(0, 256, (B, S, S, 3), dtype=np.uint8))
buf, off, hw, total = pack_images(imgs); src = torch.from_numpy(buf[:total]).cuda()
out = torch.empty((B, cfg['embed_dim']), dtype=torch.float32, device='cuda')
for _ in range(3): eng.embed_u8(src, off, hw, 0, out=out)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(8): eng.embed_u8(src, off, hw, 0, out=out)
b.record(); torch.cuda.synchronize()
print(a.elapsed_time(b) / 8)
"""
for rep in range(3):
    for v in ("1", "0"):
        r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **{var: v}), capture_output=True, text=True)
        print(var, v, model, r.stdout.strip() or r.stderr[-300:], flush=True)
