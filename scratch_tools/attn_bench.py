"""Times b200clip_attention alone (env B200_ATTN_TC=0/1 picks the kernel) on the three model shapes."""
import sys, torch
sys.path.insert(0, ".")
from lumen_b200.engine import Engine
eng = Engine({}, device=0)
g = torch.Generator(device="cuda").manual_seed(0)
for name, B, S, heads in (("B/32", 512, 50, 12), ("B/16", 256, 197, 12), ("L/14@336", 128, 577, 16)):
    d = heads * 64
    qkv = torch.randn(B * S, 3 * d, device="cuda", generator=g).to(torch.bfloat16)
    for _ in range(5): eng.attention(qkv, B, S, d, heads)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50): eng.attention(qkv, B, S, d, heads)
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / 50
    fl = 4.0 * B * heads * S * S * 64
    print(f"{name}: {us:.1f} us  {fl / us / 1e6:.1f} TFLOP/s (useful)")
