"""GPU: differential test of the fused GEMM + top-k search (K10 + K12) against a full sort on exact data.

Small-integer embeddings make every dot product exact in fp32, and a span of one or two produces thousands of equal scores per query,
so indices - not only scores - must match the (score descending, index ascending) order bit for bit.  The shapes walk through every
regime of the kernel: one or several query tiles, fewer rows than a tile, ragged last tiles, galleries large enough for the seeding
pass (more than 16 x 16384 rows), k above the number of rows, lists that overflow their share of a query's slice."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _cases(seed, n):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        out.append((int(rng.choice([1, 3, 17, 64, 127, 128, 129, 200, 256, 300, 513, 700])),
                    int(rng.choice([37, 255, 256, 257, 1000, 5000, 33000, 70000, 262144, 263000, 300001, 524288, 600000])),
                    int(rng.choice([64, 128, 256, 512])), int(rng.choice([1, 5, 10, 50, 100, 128, 200, 256])),
                    int(rng.choice([1, 2, 8])), bool(rng.random() < 0.5)))
    return out


@pytest.mark.parametrize("seed", [1, 2])
def test_search_matches_full_sort_on_exact_data(seed):
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    gen = torch.Generator(device="cuda").manual_seed(100 + seed)
    for Q, N, D, k, span, dup in _cases(seed, 20):
        g = torch.randint(-span, span + 1, (N, D), device="cuda", generator=gen, dtype=torch.int32).to(torch.float16)
        q = torch.randint(-span, span + 1, (Q, D), device="cuda", generator=gen, dtype=torch.int32).to(torch.float16)
        if dup and N > 100:
            g[N // 2:N // 2 + N // 10] = g[:N // 10]                    # exact duplicates far apart
        s, i = eng.topk(q, g, k, index_base=7)
        eng.sync()
        full = q.float() @ g.float().T
        kk = min(k, N)
        order = torch.sort(-full, dim=1, stable=True).indices[:, :kk]   # stable: equal scores keep ascending index
        tag = dict(Q=Q, N=N, D=D, k=k, span=span, dup=dup)
        assert torch.equal(i[:, :kk], order + 7), tag
        assert torch.equal(s[:, :kk], torch.gather(full, 1, order)), tag
        if k > N:
            assert bool((i[:, N:] == -1).all()), tag
        del g, q, full, order
    eng.close()
