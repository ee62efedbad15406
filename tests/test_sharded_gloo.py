"""CPU, world_size 2 over gloo: the host-side logic of the sharded gallery search (shard bounds, index_base, the
all-gather layout and the merge call).  The device kernels are replaced by the oracle behind the same Engine methods;
the real kernels are covered by the -m gpu tests."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


class OracleEngine:
    """Same method surface as lumen_b200.engine.Engine for topk / merge_topk, computed by the oracle on CPU tensors."""

    def topk(self, q, g, k, index_base=0):
        from oracle import clip_oracle as O
        s, i = O.topk_cosine(q.numpy(), g.numpy(), k, index_base=index_base)
        return torch.from_numpy(s), torch.from_numpy(i)

    def merge_topk(self, ps, pi):
        from oracle import clip_oracle as O
        s, i = O.merge_topk(list(ps.numpy()), list(pi.numpy()), ps.shape[2])
        return torch.from_numpy(s), torch.from_numpy(i)


def _worker(rank, world, port, out_dir):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
    from lumen_b200.managers import ShardedGallery
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    N, D, Q, k = 1003, 64, 9, 20
    g = rng.integers(-8, 9, (N, D)).astype(np.float16)
    g[500:600] = g[:100]                                  # ties across the shard boundary (shard 0 | shard 1 at row 502)
    q = rng.integers(-8, 9, (Q, D)).astype(np.float16)
    lo, hi = ShardedGallery.shard_bounds(N, world, rank)
    gal = ShardedGallery(OracleEngine(), torch.from_numpy(g[lo:hi].copy()), index_base=lo)
    s, i = gal.search(torch.from_numpy(q), k)
    # a gallery of ONE row over two ranks: rank 1's shard is empty and contributes only empty slots
    lo1, hi1 = ShardedGallery.shard_bounds(1, world, rank)
    tiny = ShardedGallery(OracleEngine(), torch.from_numpy(g[lo1:hi1].copy()), index_base=lo1)
    s1, i1 = tiny.search(torch.from_numpy(q), 3)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), s=s.numpy(), i=i.numpy(), g=g, q=q, s1=s1.numpy(), i1=i1.numpy(), empty=np.array(hi1 - lo1))
    dist.destroy_process_group()


def test_sharded_gallery_world2_gloo(tmp_path):
    from oracle import clip_oracle as O
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    want_s, want_i = O.topk_cosine(r0["q"], r0["g"], 20)
    for r in (r0, r1):                                     # every rank holds the merged global answer
        assert np.array_equal(r["i"], want_i) and np.array_equal(r["s"], want_s)
    assert int(r0["empty"]) == 1 and int(r1["empty"]) == 0
    w1s, w1i = O.topk_cosine(r0["q"], r0["g"][:1], 3)
    for r in (r0, r1):
        assert np.array_equal(r["i1"], w1i) and np.array_equal(r["s1"], w1s)
