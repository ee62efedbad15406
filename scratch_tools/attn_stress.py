import sys, torch
sys.path.insert(0, ".")
from lumen_b200.engine import Engine
eng = Engine({}, device=0)
g = torch.Generator(device="cuda").manual_seed(4)
bad = 0
for S in (65, 66, 96, 127, 128, 191, 192, 193, 255, 256, 257, 320, 511, 513, 1025):
    for heads in (1, 2, 3):
        for B in (1, 3):
            for causal in (False, True):
                d = heads * 64
                qkv = torch.randn(B * S, 3 * d, device="cuda", generator=g).to(torch.bfloat16)
                out = eng.attention(qkv, B, S, d, heads, causal=causal); eng.sync()
                q, k, v = [t.view(B, S, heads, 64).transpose(1, 2) for t in qkv.float().split(d, dim=1)]
                sc = q @ k.transpose(-1, -2) * 0.125
                if causal: sc = sc + torch.full((S, S), float("-inf"), device="cuda").triu(1)
                ref = (torch.softmax(sc, -1) @ v).transpose(1, 2).reshape(B * S, d)
                err = (out.float() - ref).abs().max().item()
                if not err < 2e-2: bad += 1; print("BAD", S, heads, B, causal, err)
print("stress done, bad =", bad)
