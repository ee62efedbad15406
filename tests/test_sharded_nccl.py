"""GPU, world_size 2 over NCCL: the sharded gallery search of BASELINE config C5 on real devices — every rank runs the fused
GEMM + top-k kernel on its row shard, the candidates cross NVLink in one NCCL all-gather and every rank merges them.  Same
fixture as tests/test_sharded_gloo.py (exact arithmetic, duplicated rows straddling the shard boundary), indices exact against the
oracle.  Needs two GPUs: skipped on a single-GPU box (run it with `gpurun --gpus 2`)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _worker(rank, world, port, out_dir):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
    from lumen_b200.engine import Engine
    from lumen_b200.gallery import DeviceGallery
    from lumen_b200.managers import ShardedGallery
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    eng = Engine({}, device=rank)
    rng = np.random.default_rng(5)
    out = {}
    # (a) the gloo fixture; (b) a batch of more than one query tile with k = 100 and a ragged last shard
    for tag, (N, D, Q, k) in {"a": (1003, 64, 9, 20), "b": (40001, 512, 300, 100)}.items():
        g = rng.integers(-8, 9, (N, D)).astype(np.float16)
        g[N // 2 - 2:N // 2 + 98] = g[:100]               # exact duplicates across the shard boundary -> index-ordered ties
        q = rng.integers(-8, 9, (Q, D)).astype(np.float16)
        lo, hi = ShardedGallery.shard_bounds(N, world, rank)
        gal = ShardedGallery(eng, torch.from_numpy(g[lo:hi].copy()).cuda(), index_base=lo)
        s, i = gal.search(torch.from_numpy(q).cuda(), k)
        eng.sync()
        out[tag + "_s"], out[tag + "_i"], out[tag + "_g"], out[tag + "_q"] = s.cpu().numpy(), i.cpu().numpy(), g, q
    # (c) the same through the gallery store: fp32 rows appended per rank, normalised + packed on the device
    gf = np.random.default_rng(6).standard_normal((5000, 128)).astype(np.float32)
    qf = np.random.default_rng(7).standard_normal((33, 128)).astype(np.float32)
    qf /= np.linalg.norm(qf, axis=1, keepdims=True)
    lo, hi = ShardedGallery.shard_bounds(5000, world, rank)
    dg = DeviceGallery(eng, 128, index_base=lo)
    dg.append(gf[lo:hi], normalize=True)
    s, i = dg.search(torch.from_numpy(qf.astype(np.float16)).cuda(), 10)
    eng.sync()
    out["c_s"], out["c_i"], out["c_rows"], out["c_q"] = s.cpu().numpy(), i.cpu().numpy(), dg.view().cpu().numpy(), qf
    # (d) the one-call C-ABI form on a RAW ncclComm_t (what a caller without torch.distributed would hold): the communicator is made
    # with ncclCommInitRank through ctypes on the libnccl torch has loaded; its unique id travels over the process group above
    import ctypes as C
    nccl = None
    for cand in (os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "nccl", "lib", "libnccl.so.2"), "libnccl.so.2"):
        try:
            nccl = C.CDLL(cand); break
        except OSError:
            continue
    if nccl is not None:
        class UniqueId(C.Structure):
            _fields_ = [("internal", C.c_byte * 128)]
        uid = UniqueId()
        if rank == 0:
            assert nccl.ncclGetUniqueId(C.byref(uid)) == 0
        t = torch.frombuffer(bytearray(bytes(uid)), dtype=torch.uint8).cuda()
        dist.broadcast(t, 0)
        C.memmove(C.byref(uid), bytes(t.cpu().numpy().tobytes()), 128)
        comm = C.c_void_p()
        nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
        assert nccl.ncclCommInitRank(C.byref(comm), world, uid, rank) == 0
        g, q = out["b_g"], out["b_q"]
        lo, hi = ShardedGallery.shard_bounds(g.shape[0], world, rank)
        s, i = eng.search_sharded(torch.from_numpy(q).cuda(), torch.from_numpy(g[lo:hi].copy()).cuda(), 100, lo, comm.value, world)
        eng.sync()
        out["d_s"], out["d_i"] = s.cpu().numpy(), i.cpu().numpy()
        # an empty shard on rank 1 (all rows on rank 0) must still give the global answer everywhere
        shard = torch.from_numpy(g.copy()).cuda() if rank == 0 else None
        s, i = eng.search_sharded(torch.from_numpy(q).cuda(), shard, 100, 0 if rank == 0 else g.shape[0], comm.value, world)
        eng.sync()
        out["e_s"], out["e_i"] = s.cpu().numpy(), i.cpu().numpy()
        nccl.ncclCommDestroy.argtypes = [C.c_void_p]
        nccl.ncclCommDestroy(comm)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), **out)
    eng.close()
    dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_gallery_world2_nccl(tmp_path):
    from oracle import clip_oracle as O
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    for tag, k in (("a", 20), ("b", 100)):
        want_s, want_i = O.topk_cosine(r0[tag + "_q"], r0[tag + "_g"], k)
        for r in (r0, r1):                                     # every rank holds the merged global answer
            assert np.array_equal(r[tag + "_i"], want_i), tag
            assert np.array_equal(r[tag + "_s"], want_s), tag
    assert "d_i" in r0.files and "d_i" in r1.files, "libnccl not loadable through ctypes"
    want_s, want_i = O.topk_cosine(r0["b_q"], r0["b_g"], 100)
    for r in (r0, r1):                                         # b200clip_search_sharded on a raw communicator; with an empty shard
        for tag in ("d", "e"):
            assert np.array_equal(r[tag + "_i"], want_i) and np.array_equal(r[tag + "_s"], want_s), tag
    g16 = np.concatenate([r0["c_rows"], r1["c_rows"]])        # the rows as the device packed them (<= 1 fp16 ulp from the oracle's)
    assert g16.shape == (5000, 128)
    want_s, want_i = O.topk_cosine(r0["c_q"].astype(np.float16), g16, 10)
    for r in (r0, r1):
        np.testing.assert_allclose(r["c_s"], want_s, atol=2e-6)
        assert np.mean(r["c_i"] == want_i) > 0.97             # swaps among near-ties only
    assert np.array_equal(r0["c_i"], r1["c_i"]) and np.array_equal(r0["c_s"], r1["c_s"])
