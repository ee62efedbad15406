"""One attention shape, a few launches (for ncu): python attn_one.py B S heads"""
import sys, torch
sys.path.insert(0, ".")
from lumen_b200.engine import Engine
B, S, heads = map(int, sys.argv[1:4])
eng = Engine({}, device=0)
d = heads * 64
qkv = torch.randn(B * S, 3 * d, device="cuda").to(torch.bfloat16)
for _ in range(3): eng.attention(qkv, B, S, d, heads)
eng.sync()
