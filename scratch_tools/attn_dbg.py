import sys, torch
sys.path.insert(0, ".")
from lumen_b200.engine import Engine
eng = Engine({}, device=0)
g = torch.Generator(device="cuda").manual_seed(2)
B, S, heads = 1, 64, 2
d = heads * 64
base = torch.randn(B * S, 3 * d, device="cuda", generator=g)
def run(qkv, tag):
    qkv = qkv.to(torch.bfloat16)
    out = eng.attention(qkv, B, S, d, heads); eng.sync()
    q, k, v = [t.view(B, S, heads, 64).transpose(1, 2) for t in qkv.float().split(d, dim=1)]
    ref = (torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1) @ v)
    o = out.float().view(B, S, heads, 64).transpose(1, 2)
    print(tag, "err", (o - ref).abs().max().item(), "o[0,0,:3,:4]", o[0, 0, :3, :4].tolist(), "ref", ref[0, 0, :3, :4].tolist())
x = base.clone(); x[:, 2 * d:] = 1.0; run(x, "V=1")
x = base.clone(); x[:, :d] = 0.0; run(x, "Q=0")
x = base.clone(); x[:, d:2 * d] = 0.0; run(x, "K=0")
x = base.clone(); x[:, 2 * d:] = torch.arange(S, device="cuda").float()[:, None] / 64; run(x, "V=key index")
x = base.clone(); x[:, :d] = 0; x[:, 2 * d:] = torch.arange(S, device="cuda").float()[:, None] / 64; run(x, "Q=0,V=key index (expect 0.492)")
