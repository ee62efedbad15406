"""GPU: b200clip_topk_static — the search entry point that allocates nothing.

Some callers cannot allocate device memory when they search: a thread whose stream is under CUDA-graph capture, a
latency-critical serving loop that must stay clear of the allocator, a plug-in confined to another framework's pool.
For them the library carries its scratch as a fixed array inside the .so.  This file runs last on purpose (pytest
collects it after every other test file), checks the entry point through the C ABI against the handle-based
b200clip_topk on two consecutive calls (the array keeps the first call's contents, the second must not see them),
and checks that a request beyond the array's size is refused with a clear message."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _static_topk(lib, q, g, k, index_base=0):
    Q, D = q.shape
    scores = torch.empty((Q, k), dtype=torch.float32, device=q.device)
    idx = torch.empty((Q, k), dtype=torch.int64, device=q.device)
    rc = lib.b200clip_topk_static(C.c_void_p(q.data_ptr()), Q, C.c_void_p(g.data_ptr()), g.shape[0], D, k, index_base,
                                  C.c_void_p(scores.data_ptr()), C.c_void_p(idx.data_ptr()),
                                  C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, scores, idx


def test_static_scratch_topk_matches_handle_path_twice():
    from lumen_b200 import _lib
    from lumen_b200.engine import Engine
    lib = _lib.load()
    torch.cuda.set_device(0)
    eng = Engine({}, device=0)
    rng = np.random.default_rng(77)
    # call 1: exact-arithmetic gallery with duplicated rows (index-ordered ties) -> bit-exact agreement required
    g1 = rng.integers(-8, 9, (40000, 128)).astype(np.float16); g1[20000:24000] = g1[:4000]
    q1 = rng.integers(-8, 9, (300, 128)).astype(np.float16)
    # call 2: different shapes and Gaussian unit vectors, straight after call 1 (stale scratch must not leak)
    g2 = rng.standard_normal((70001, 512)).astype(np.float32); g2 /= np.linalg.norm(g2, axis=1, keepdims=True)
    q2 = rng.standard_normal((37, 512)).astype(np.float32); q2 /= np.linalg.norm(q2, axis=1, keepdims=True)
    cases = [(q1, g1, 100, 5), (q2.astype(np.float16), g2.astype(np.float16), 64, 0)]
    from oracle import clip_oracle as O
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.b200clip_topk_static_status(stream) == 0            # nothing has run yet on this device, or nothing tripped
    for n, (q, g, k, base) in enumerate(cases):
        qd, gd = torch.from_numpy(q).cuda(), torch.from_numpy(g).cuda()
        rc, s, i = _static_topk(lib, qd, gd, k, base)
        assert rc == 0, lib.b200clip_last_error()
        assert lib.b200clip_topk_static_status(stream) == 0, lib.b200clip_last_error()    # the watchdog words of the array are clean
        rs, ri = eng.topk(qd, gd, k, index_base=base)
        torch.cuda.synchronize()
        assert torch.equal(i, ri)
        assert torch.equal(s, rs)           # same kernels, same per-element arithmetic: identical bits
        if n == 0:                          # exact arithmetic: the oracle's answer, bit for bit, ties in index order
            so, io = O.topk_cosine(q, g, k, index_base=base)
            assert np.array_equal(i.cpu().numpy(), io) and np.array_equal(s.cpu().numpy(), so)
    eng.close()


def test_static_scratch_rejects_oversize():
    from lumen_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    q = torch.zeros((20000, 64), dtype=torch.float16, device="cuda")
    g = torch.zeros((512, 64), dtype=torch.float16, device="cuda")
    rc, _, _ = _static_topk(lib, q, g, 100)
    assert rc == _lib.EINVAL
    msg = lib.b200clip_last_error().decode()
    assert "built-in array" in msg and "b200clip_topk" in msg
    rc, _, _ = _static_topk(lib, q[:8], g, 300)
    assert rc == _lib.EINVAL and "256" in lib.b200clip_last_error().decode()
