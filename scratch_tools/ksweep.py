import sys, json, torch
sys.path.insert(0, '/root/repo')
from lumen_b200 import _lib
from lumen_b200.engine import Engine
from bench_paths import timed
eng = Engine({}, device=0)
M, N = 25600, 2304
for epi in (_lib.EPI_BIAS_BF16, _lib.EPI_F32):
    for K in (128, 256, 512, 768, 1536, 3072):
        A = torch.randn(M, K, device="cuda").to(torch.bfloat16); W = torch.randn(N, K, device="cuda").to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        out = torch.zeros((M, N), dtype=torch.float32 if epi == _lib.EPI_F32 else torch.bfloat16, device="cuda")
        ms = timed(lambda: eng.gemm(A, W, bias, epi, 0, out=out), 20)
        print(f"epi={epi} K={K} us={ms*1e3:.1f} TF={2.0*M*N*K/ms/1e9:.0f}")
