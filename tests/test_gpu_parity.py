"""GPU parity tests proper: the CUDA path, called through the C ABI, against the oracle and the golden fixtures."""
import numpy as np
import pytest

from conftest import synth_image

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def eng224():
    from lumen_b200.engine import Engine
    e = Engine({"patch": 32}, device=0, image_size=224)
    yield e
    e.close()


def _run_pre(eng, imgs, mode, layout=0, bf16=False):
    from lumen_b200.engine import pack_images
    buf, off, hw, total = pack_images(imgs)
    src = torch.from_numpy(buf[:total]).cuda()
    out = eng.preprocess(src, off, hw, mode=mode, layout=layout, bf16=bf16)
    eng.sync()
    return out


# ------------------------------------------------------------------------------------------ K1 preprocess
@pytest.mark.parametrize("mode", [0, 1])
def test_preprocess_golden_bit_exact(golden_pre, mode):
    """Integer part bit-exact against Pillow/torchvision goldens; fp32 normalise bit-exact against the oracle."""
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    by_size = {}
    for n, (seed, h, w, S) in enumerate(golden_pre["cases"]):
        by_size.setdefault(int(S), []).append((n, int(seed), int(h), int(w)))
    for S, cases in by_size.items():
        eng = Engine({"patch": 16 if S == 64 else 32 if S == 224 else 14}, device=0, image_size=S)
        imgs = [synth_image(seed, h, w) for _, seed, h, w in cases]
        out = _run_pre(eng, imgs, mode).cpu().numpy()
        mean = np.array(O.OPENAI_MEAN, np.float32)[:, None, None]
        std = np.array(O.OPENAI_STD, np.float32)[:, None, None]
        for j, (n, seed, h, w) in enumerate(cases):
            key = ("squash_u8_%d" if mode == 0 else "crop_u8_%d") % n
            want_u8 = golden_pre[key].transpose(2, 0, 1)
            want = (want_u8.astype(np.float32) / np.float32(255.0) - mean) / std
            got_u8 = np.rint((out[j] * std + mean) * 255.0)
            assert np.array_equal(got_u8, want_u8), f"S={S} case {n}: max |d|={np.abs(got_u8 - want_u8).max()}"
            assert np.array_equal(out[j], want), f"S={S} case {n}: float normalise differs"
            assert np.array_equal(out[j], O.preprocess(imgs[j], S, mode))
        eng.close()


def test_preprocess_ragged_batch_and_layouts(eng224):
    from oracle import clip_oracle as O
    rng = np.random.default_rng(3)
    # (the last three are more than 100 times taller than wide: PIL resizes those vertically first, and so must the device)
    sizes = [(224, 224), (31, 57), (640, 480), (480, 640), (1000, 200), (225, 224), (224, 223), (1080, 1920), (1, 1), (2, 3), (1, 400),
             (2300, 19), (701, 7), (700, 7), (450, 1)]
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in sizes]
    for mode in (0, 1):
        want = O.preprocess_batch(imgs, 224, mode)
        got = _run_pre(eng224, imgs, mode).cpu().numpy()
        assert np.array_equal(got, want)
        # bf16 NCHW = round-to-nearest of the fp32 result
        got16 = _run_pre(eng224, imgs, mode, bf16=True)
        assert torch.equal(got16.cpu(), torch.from_numpy(want).to(torch.bfloat16))
        # patch-major layout: [B*49, 3072] with K index (c, i, j)
        pm = _run_pre(eng224, imgs, mode, layout=1).cpu().numpy()
        ref = want.reshape(len(imgs), 3, 7, 32, 7, 32).transpose(0, 2, 4, 1, 3, 5).reshape(len(imgs) * 49, 3072)
        assert np.array_equal(pm, ref)


def test_preprocess_patch14_padding():
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({"patch": 14}, device=0, image_size=336)
    imgs = [synth_image(21, 500, 400), synth_image(22, 336, 336)]
    want = O.preprocess_batch(imgs, 336, 1)
    pm = _run_pre(eng, imgs, 1, layout=1, bf16=True).float().cpu().numpy()
    assert pm.shape == (2 * 576, 640)
    ref = want.reshape(2, 3, 24, 14, 24, 14).transpose(0, 2, 4, 1, 3, 5).reshape(2 * 576, 588)
    assert np.array_equal(pm[:, :588], torch.from_numpy(ref).to(torch.bfloat16).float().numpy())
    assert (pm[:, 588:] == 0).all()
    eng.close()


def test_preprocess_oversize_sources_and_many_sizes(eng224):
    """Sources no shared-memory tile can hold (a 23,000-pixel-wide strip, a 12,000-pixel-tall one: 53x vertical reduction) take the
    global-memory kernels and are still bit-exact, mixed in one batch with ordinary images; and a stream of ever-new source sizes
    (more than the coefficient-table cache holds) keeps working - the cache evicts, it never fails."""
    from oracle import clip_oracle as O
    rng = np.random.default_rng(8)
    sizes = [(30, 23000), (300, 400), (12000, 40), (224, 224), (26, 22500)]
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in sizes]
    for mode in (0, 1):
        want = O.preprocess_batch(imgs, 224, mode)
        got = _run_pre(eng224, imgs, mode).cpu().numpy()
        assert np.array_equal(got, want), mode
        pm = _run_pre(eng224, imgs, mode, layout=1, bf16=True).float().cpu().numpy()
        ref = want.reshape(len(imgs), 3, 7, 32, 7, 32).transpose(0, 2, 4, 1, 3, 5).reshape(len(imgs) * 49, 3072)
        assert np.array_equal(pm, torch.from_numpy(ref).to(torch.bfloat16).float().numpy())
    # 4500 distinct widths -> 4500 horizontal tables pass through a 4096-entry cache
    base = rng.integers(0, 256, (40, 5000, 3), dtype=np.uint8)
    for lo in range(0, 4500, 500):
        batch = [np.ascontiguousarray(base[:, :300 + lo + j]) for j in range(500)]
        out = _run_pre(eng224, batch, 0)
        if lo in (0, 4000):
            pick = [0, 250, 499]
            assert np.array_equal(out[pick].cpu().numpy(), O.preprocess_batch([batch[j] for j in pick], 224, 0))
    # the first sizes were evicted long ago: asking for them again rebuilds their tables
    again = [np.ascontiguousarray(base[:, :300 + j]) for j in (0, 1, 2)]
    assert np.array_equal(_run_pre(eng224, again, 0).cpu().numpy(), O.preprocess_batch(again, 224, 0))


def test_preprocess_empty_and_errors(eng224):
    from lumen_b200 import _lib
    src = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = eng224.preprocess(src, np.zeros(0, np.int64), np.zeros((0, 2), np.int32))
    assert out.shape == (0, 3, 224, 224)
    with pytest.raises(_lib.B200ClipError):
        eng224.preprocess(src, np.array([8], np.int64), np.array([[2, 2]], np.int32))     # misaligned offset
    with pytest.raises(_lib.B200ClipError):
        eng224.preprocess(src, np.array([0], np.int64), np.array([[0, 2]], np.int32))     # empty image


# ------------------------------------------------------------------------------------------ GEMM / LN / attention
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 128), (300, 384, 128), (1000, 768, 768), (3200, 2304, 768),
                                   (77, 64, 256), (25600, 768, 3072)])
def test_gemm_epilogues(M, N, K):
    from lumen_b200 import _lib
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g)).to(torch.bfloat16)
    W = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda", generator=g)
    ref = A.float() @ W.float().T
    out = eng.gemm(A, W)
    eng.sync()
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-3)
    out = eng.gemm(A, W, bias, _lib.EPI_BIAS_BF16)
    torch.testing.assert_close(out.float(), (ref + bias).to(torch.bfloat16).float(), rtol=2e-2, atol=2e-2)
    out = eng.gemm(A, W, bias, _lib.EPI_BIAS_ACT_BF16, 0)
    y = ref + bias
    torch.testing.assert_close(out.float(), (y * torch.sigmoid(1.702 * y)), rtol=2e-2, atol=2e-2)
    out = eng.gemm(A, W, bias, _lib.EPI_BIAS_ACT_BF16, 1)
    torch.testing.assert_close(out.float(), torch.nn.functional.gelu(y), rtol=2e-2, atol=2e-2)
    res = torch.randn(M, N, device="cuda", generator=g)
    out = eng.gemm(A, W, bias, _lib.EPI_BIAS_RES_F32, out=res.clone())
    torch.testing.assert_close(out, res + ref + bias, rtol=1e-4, atol=2e-3)
    eng.sync()
    eng.close()


def test_gemm_exact_small_integers():
    """Integer-valued operands: every partial sum is exact, so the tensor-core result must equal fp32 exactly
    (catches any descriptor / swizzle / tile-mapping error, not just large ones)."""
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    g = torch.Generator(device="cuda").manual_seed(1)
    for M, N, K in ((128, 256, 64), (512, 512, 256), (131, 200, 192)):
        A = torch.randint(-4, 5, (M, K), device="cuda", generator=g).to(torch.bfloat16)
        W = torch.randint(-4, 5, (N, K), device="cuda", generator=g).to(torch.bfloat16)
        out = eng.gemm(A, W)
        eng.sync()
        assert torch.equal(out, A.float() @ W.float().T), (M, N, K)
    eng.close()


def test_layernorm_and_attention():
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    g = torch.Generator(device="cuda").manual_seed(2)
    for d in (128, 768, 1024):
        x = torch.randn(333, d, device="cuda", generator=g) * 3 + 1
        gam, bet = torch.randn(d, device="cuda", generator=g), torch.randn(d, device="cuda", generator=g)
        out = eng.layernorm(x, gam, bet)
        ref = torch.nn.functional.layer_norm(x, (d,), gam, bet, 1e-5)
        torch.testing.assert_close(out.float(), ref, rtol=1e-2, atol=2e-2)
    # S > 64: attention_tc_kernel, 64-key steps; S <= 64: attention_short_kernel (two problems per tile; odd problem counts leave
    # the last tile half empty); 300 x 50, 513 x 50 and 40 x 197 give each persistent CTA several tiles
    for B, S, heads in ((3, 50, 12), (2, 197, 12), (1, 577, 16), (5, 17, 2), (4, 64, 4), (2, 128, 2), (2, 129, 2), (3, 257, 16),
                        (300, 50, 12), (40, 197, 12), (2, 50, 3), (1, 1, 2), (1, 50, 3), (3, 33, 5), (1, 64, 1), (513, 50, 12)):
        d = heads * 64
        qkv = torch.randn(B * S, 3 * d, device="cuda", generator=g)
        if B == 40:   # keys grow along the sequence: later key steps raise the row maximum by far more than 2^8 (in-place O rescale)
            qkv[:, d:2 * d] *= torch.linspace(0.2, 3.0, S, device="cuda").repeat(B)[:, None]
        qkv = qkv.to(torch.bfloat16)
        out = eng.attention(qkv, B, S, d, heads)
        q, k, v = [t.view(B, S, heads, 64).transpose(1, 2) for t in qkv.float().split(d, dim=1)]
        ref = torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1) @ v
        ref = ref.transpose(1, 2).reshape(B * S, d)
        torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)
    eng.sync()
    eng.close()


# ------------------------------------------------------------------------------------------ encoder end to end
@pytest.mark.parametrize("name,B,max_batch", [("ViT-tiny-test", 5, 3), ("ViT-B-32", 4, 3), ("ViT-B-16", 6, 4), ("ViT-L-14-336", 2, 2)])
def test_encoder_cosine_vs_oracle_and_golden(golden_enc, name, B, max_batch):
    """Every tower BASELINE.json's configs name (B/32: C1-C2, B/16: C3, L/14@336: C4), composed end to end in bf16 against the fp32
    oracle and against transformers' own output (golden)."""
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    cfg = VIT_CONFIGS[name]
    w = synthetic_weights(cfg, seed=0)
    eng = Engine(cfg, device=0, max_batch=max_batch)             # max_batch < B exercises chunking
    eng.load_weights(w)
    rng = np.random.default_rng(12)
    imgs = [rng.integers(0, 256, (cfg["image_size"] + 13 * i, cfg["image_size"] + 7 * i, 3), dtype=np.uint8) for i in range(B)]
    emb = eng.embed_host(imgs, mode=0)
    ref = O.vit_encode(w, O.preprocess_batch(imgs, cfg["image_size"], 0), cfg)
    cos = (emb * ref).sum(1)
    assert cos.min() >= 0.999, cos                                  # BASELINE.json tolerance
    np.testing.assert_allclose(np.linalg.norm(emb, axis=1), 1.0, atol=1e-5)
    # golden (transformers' CLIPVisionModelWithProjection on seeded Gaussian pixels): feed the same pixels as patches
    gold = golden_enc[f"{name}/image_embeds"]
    G = gold.shape[0]
    px = np.random.default_rng(11).standard_normal((G, 3, cfg["image_size"], cfg["image_size"])).astype(np.float32)
    P, n_side = cfg["patch"], cfg["image_size"] // cfg["patch"]
    pm = px.reshape(G, 3, n_side, P, n_side, P).transpose(0, 2, 4, 1, 3, 5).reshape(G * n_side * n_side, 3 * P * P)
    kpad = (3 * P * P + 63) // 64 * 64
    pmp = np.zeros((pm.shape[0], kpad), np.float32); pmp[:, :pm.shape[1]] = pm
    e2 = eng.encode(torch.from_numpy(pmp).to(torch.bfloat16).cuda(), G).cpu().numpy()
    gold = gold / np.linalg.norm(gold, axis=1, keepdims=True)
    assert (e2 * gold).sum(1).min() >= 0.999
    eng.sync()
    eng.close()


def test_backend_api_bytes_roundtrip():
    """image_to_vector / image_batch_to_vectors on encoded PNG bytes, error contract of onnxrt_backend.py:465-551."""
    import io
    from PIL import Image
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    from lumen_b200.interface import InferenceError, InvalidInputError
    from oracle import clip_oracle as O
    name = "ViT-tiny-test"; cfg = VIT_CONFIGS[name]; w = synthetic_weights(cfg, 0)
    be = B200ClipBackend(SyntheticResources(model_name=name), device_preference="cuda:0", max_batch_size=4, weights=w)
    be.initialize(); be.initialize()
    imgs = [synth_image(40 + i, 80 + 9 * i, 100 - 5 * i) for i in range(6)]
    blobs = []
    for im in imgs:
        bio = io.BytesIO(); Image.fromarray(im, "RGB").save(bio, format="PNG"); blobs.append(bio.getvalue())
    batch = be.image_batch_to_vectors(blobs)
    one = be.image_to_vector(blobs[2])
    ref = O.vit_encode(w, O.preprocess_batch(imgs, cfg["image_size"], 0), cfg)
    assert batch.dtype == np.float32 and batch.shape == ref.shape and batch.flags["C_CONTIGUOUS"]
    assert (batch * ref).sum(1).min() >= 0.999 and float(one @ ref[2]) >= 0.999
    with pytest.raises(InvalidInputError):
        be.image_to_vector(b"")
    with pytest.raises(InferenceError):
        be.image_to_vector(b"not an image")
    with pytest.raises(InferenceError):
        be.image_batch_to_vectors([blobs[0], b""])
    be.close()


# ------------------------------------------------------------------------------------------ classify / search
def test_classify_matches_reference_golden(golden_cls):
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    text = np.random.default_rng(7).standard_normal((1000, 512)).astype(np.float32)
    scene = np.random.default_rng(8).standard_normal((8, 512)).astype(np.float32)
    eng.set_labels(text, 0); eng.set_labels(scene, 1)
    q = torch.from_numpy(golden_cls["queries"]).cuda()
    vals, idx = eng.classify(q, 5)
    assert np.array_equal(idx.cpu().numpy(), golden_cls["cls_idx"])
    np.testing.assert_allclose(vals.cpu().numpy(), golden_cls["cls_p"], rtol=2e-3, atol=1e-6)
    sv, si = eng.classify(q, 1, slot=1, scale=1.0)
    assert np.array_equal(si.cpu().numpy()[:, 0], golden_cls["scene_idx"])
    np.testing.assert_allclose(sv.cpu().numpy()[:, 0], golden_cls["scene_p"], rtol=1e-4)
    eng.close()


def test_topk_exact_integer_gallery_with_ties():
    """Small-integer embeddings: all dot products exact in fp32, duplicated rows force ties -> indices must be exact."""
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({}, device=0)
    rng = np.random.default_rng(20)
    # (the last case is large enough for the seeding pass: two query tiles, more than 16 x 16384 rows)
    for Q, N, D, k in ((5, 3000, 64, 10), (130, 20000, 512, 100), (1, 700, 128, 256), (33, 257, 64, 300), (4, 50, 64, 100), (200, 300000, 64, 100)):
        g = rng.integers(-8, 9, (N, D)).astype(np.float16)
        g[N // 2:N // 2 + N // 10] = g[:N // 10]                   # exact duplicates -> index-ordered ties
        q = rng.integers(-8, 9, (Q, D)).astype(np.float16)
        s, i = eng.topk(torch.from_numpy(q).cuda(), torch.from_numpy(g).cuda(), k, index_base=1000)
        eng.sync()
        so, io = O.topk_cosine(q, g, k, index_base=1000)
        io = np.where(io >= 0, io, -1)
        assert np.array_equal(i.cpu().numpy(), io), (Q, N, D, k)
        assert np.array_equal(s.cpu().numpy(), so)
    eng.close()


def test_topk_gaussian_matches_reference_golden(golden_cls):
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({}, device=0)
    g = np.random.default_rng(10).standard_normal((5000, 512)).astype(np.float32)
    g /= np.linalg.norm(g, axis=1, keepdims=True)
    g16 = g.astype(np.float16); q16 = golden_cls["queries"].astype(np.float16)
    s, i = eng.topk(torch.from_numpy(q16).cuda(), torch.from_numpy(g16).cuda(), 10)
    so, io = O.topk_cosine(q16, g16, 10)            # oracle on the same fp16-rounded values, fp32 math
    s, i = s.cpu().numpy(), i.cpu().numpy()
    np.testing.assert_allclose(s, so, atol=2e-6)
    for r in range(len(q16)):                       # sets equal up to swaps among near-ties (< 1e-5)
        if not np.array_equal(i[r], io[r]):
            diff = set(i[r]) ^ set(io[r])
            assert all(abs(float(q16[r].astype(np.float32) @ g16[j].astype(np.float32)) - so[r, -1]) < 1e-5 for j in diff)
    # and against the reference's own fp32 run (golden): same neighbours, scores within fp16 rounding
    assert np.mean(i == golden_cls["bio_idx"]) > 0.9
    np.testing.assert_allclose(s, golden_cls["bio_s"], atol=1e-3)
    eng.close()


def test_topk_adversarial_ascending_and_merge():
    """Every new row beats the threshold (worst case for the candidate lists) + merge of shard results."""
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({}, device=0)
    N, D, k = 6000, 64, 50
    g = np.zeros((N, D), np.float16); g[:, 0] = (np.arange(N) // 3)       # ascending scores with 3-way ties
    q = np.zeros((3, D), np.float16); q[:, 0] = [1, -1, 0]
    gt = torch.from_numpy(g).cuda(); qt = torch.from_numpy(q).cuda()
    s, i = eng.topk(qt, gt, k)
    so, io = O.topk_cosine(q, g, k)
    assert np.array_equal(i.cpu().numpy(), io) and np.array_equal(s.cpu().numpy(), so)
    parts = [eng.topk(qt, gt[lo:lo + 1500].contiguous(), k, index_base=lo) for lo in range(0, N, 1500)]
    ms, mi = eng.merge_topk(torch.stack([p[0] for p in parts]), torch.stack([p[1] for p in parts]))
    assert np.array_equal(mi.cpu().numpy(), io) and np.array_equal(ms.cpu().numpy(), so)
    # the same with two query tiles: there a scanner hands its whole list to the merge kernel, or - when the list ended the scan
    # longer than its share of the query's slice, as ascending scores make it - selects its k best first
    N2 = 40000
    g2 = np.zeros((N2, D), np.float16); g2[:, 0] = (np.arange(N2) // 3) % 2048; g2[:, 1] = np.arange(N2) // (3 * 2048)
    q2 = np.zeros((200, D), np.float16); q2[:, 0] = np.resize([1, -1, 0, 2], 200); q2[:, 1] = np.resize([8, 0, -8], 200)
    for kk in (50, 100):
        s2, i2 = eng.topk(torch.from_numpy(q2).cuda(), torch.from_numpy(g2).cuda(), kk)
        so2, io2 = O.topk_cosine(q2, g2, kk)
        assert np.array_equal(i2.cpu().numpy(), io2) and np.array_equal(s2.cpu().numpy(), so2), kk
    eng.sync()
    eng.close()


def test_topk_exchange_keys_merge_like_a_sharded_search():
    """The 64-bit candidate form the sharded search moves between GPUs, emulated on one GPU: per-shard keys -> merge with the shards'
    index bases = the unsharded answer, ties across shard boundaries in global index order, ragged last shard, k > rows of a shard."""
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({}, device=0)
    rng = np.random.default_rng(21)
    for N, D, Q, k, shards in ((10007, 64, 37, 100, 4), (1003, 64, 9, 20, 2), (300, 128, 130, 120, 3)):
        g = rng.integers(-8, 9, (N, D)).astype(np.float16)
        g[N // 2 - 2:N // 2 + 48] = g[:50]                       # duplicates straddling a shard boundary
        q = rng.integers(-8, 9, (Q, D)).astype(np.float16)
        qt = torch.from_numpy(q).cuda()
        per = (N + shards - 1) // shards
        bases, keys = [], []
        for r in range(shards):
            lo, hi = r * per, min(N, (r + 1) * per)
            keys.append(eng.topk_keys(qt, torch.from_numpy(g[lo:hi].copy()).cuda(), k))
            bases.append(lo + 5000)                              # a global offset on top, as DeviceGallery.index_base allows
        s, i = eng.merge_topk_keys(torch.stack(keys), torch.tensor(bases, dtype=torch.int64, device="cuda"))
        eng.sync()
        so, io = O.topk_cosine(q, g, k, index_base=5000)
        assert np.array_equal(i.cpu().numpy(), io) and np.array_equal(s.cpu().numpy(), so), (N, D, Q, k)
    eng.close()


def test_topk_empty_shard_contributes_nothing():
    """More ranks than gallery rows leave some shards empty: their part is all empty slots and the merge ignores it."""
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({}, device=0)
    rng = np.random.default_rng(4)
    g = rng.integers(-8, 9, (300, 64)).astype(np.float16); q = rng.integers(-8, 9, (7, 64)).astype(np.float16)
    gt, qt = torch.from_numpy(g).cuda(), torch.from_numpy(q).cuda()
    s0, i0 = eng.topk(qt, gt[:0].contiguous(), 10)
    assert bool((i0 == -1).all()) and bool(torch.isinf(s0).all())
    keys = torch.stack([eng.topk_keys(qt, gt[:0].contiguous(), 10), eng.topk_keys(qt, gt, 10)])
    ms, mi = eng.merge_topk_keys(keys, torch.tensor([0, 5], dtype=torch.int64, device="cuda"))
    so, io = O.topk_cosine(q, g, 10, index_base=5)
    eng.sync()
    assert np.array_equal(mi.cpu().numpy(), io) and np.array_equal(ms.cpu().numpy(), so)
    ps, pi = eng.merge_topk(torch.stack([s0, eng.topk(qt, gt, 10, 5)[0]]), torch.stack([i0, eng.topk(qt, gt, 10, 5)[1]]))
    eng.sync()
    assert np.array_equal(pi.cpu().numpy(), io) and np.array_equal(ps.cpu().numpy(), so)
    eng.close()


def test_topk_exact_zero_scores_across_splits():
    """Scores of exactly +0.0 (zero / padding rows, a zero query) tie across the gallery splits that share a threshold: the
    k smallest indices must win, every run (the shared bound must not exclude scores equal to it)."""
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    eng = Engine({}, device=0)
    N, D, k = 400_000, 64, 100
    g = np.zeros((N, D), np.float16)
    g[::7, 1] = -1.0                                     # some rows score below zero for query 0, above for query 2
    q = np.zeros((3, D), np.float16); q[0, 1] = 1.0; q[2, 1] = -1.0          # query 1 is all zero: every score is +0.0
    gt, qt = torch.from_numpy(g).cuda(), torch.from_numpy(q).cuda()
    so, io = O.topk_cosine(q, g, k)
    for _ in range(5):                                   # which split publishes its bound first is a race: repeat
        s, i = eng.topk(qt, gt, k)
        eng.sync()
        assert np.array_equal(i.cpu().numpy(), io) and np.array_equal(s.cpu().numpy(), so)
    assert io[1].tolist() == list(range(k))
    eng.close()


def test_full_size_properties_c2_and_search():
    """BASELINE sizes: batch 512 of 224x224 (C2) — determinism and batch-independence; 1M x 512 gallery, Q=256,
    k=100 — every returned score is the true dot product, sorted, and no better row was missed (spot rows)."""
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.engine import Engine
    cfg = VIT_CONFIGS["ViT-B-32"]
    eng = Engine(cfg, device=0, max_batch=512)
    eng.load_weights(synthetic_weights(cfg, 0))
    imgs = list(np.random.default_rng(1234).integers(0, 256, (512, 224, 224, 3), dtype=np.uint8))
    e1 = eng.embed_host(imgs); e2 = eng.embed_host(imgs)
    assert np.array_equal(e1, e2)
    e3 = eng.embed_host(imgs[100:108])
    assert ((e1[100:108] * e3).sum(1) > 0.99999).all()
    np.testing.assert_allclose(np.linalg.norm(e1, axis=1), 1.0, atol=1e-5)
    # the batch-512 pass is the one bench.py times (CTA-pair GEMMs, M = 25,600): 16 of its rows against the oracle
    from oracle import clip_oracle as O
    pick = list(range(0, 512, 32))
    w = synthetic_weights(cfg, 0)
    ref = O.vit_encode(w, O.preprocess_batch([imgs[j] for j in pick], 224, 0), cfg)
    cos = (e1[pick] * ref).sum(1)
    assert cos.min() >= 0.999, cos                                  # BASELINE.json tolerance
    gen = torch.Generator(device="cuda").manual_seed(42)
    N, Q, k = 1_000_000, 256, 100
    g = torch.nn.functional.normalize(torch.randn(N, 512, device="cuda", generator=gen), dim=1).half()
    q = torch.nn.functional.normalize(torch.randn(Q, 512, device="cuda", generator=gen), dim=1).half()
    s, i = eng.topk(q, g, k)
    eng.sync()
    assert (s[:, :-1] >= s[:, 1:]).all() and (i >= 0).all() and (i < N).all()
    rows = [0, 17, 255]
    full = q[rows].float() @ g.float().T
    ts, ti = torch.topk(full, k, dim=1)
    torch.testing.assert_close(s[rows], ts, rtol=0, atol=2e-6)
    for r, row in enumerate(rows):
        assert len(set(i[row].tolist()) ^ set(ti[r].tolist())) <= 4
    eng.close()


_GEMM_VARIANT = r"""
import sys, numpy as np
sys.path.insert(0, ".")
from lumen_b200 import VIT_CONFIGS, synthetic_weights
from lumen_b200.engine import Engine
from oracle import clip_oracle as O
cfg = VIT_CONFIGS["ViT-B-32"]; w = synthetic_weights(cfg, 0)
eng = Engine(cfg, device=0, max_batch=64); eng.load_weights(w)
imgs = list(np.random.default_rng(77).integers(0, 256, (64, 224, 224, 3), dtype=np.uint8))
e = eng.embed_host(imgs)
ref = O.vit_encode(w, O.preprocess_batch(imgs[:8], 224, 0), cfg)
cos = (e[:8] * ref).sum(1)
assert cos.min() >= 0.999, cos
np.save(sys.argv[1], e)
print("ok")
"""


def test_encoder_both_gemm_variants_vs_oracle(tmp_path):
    """The single-CTA and the CTA-pair (cta_group::2) GEMM kernels each carry a whole tower to cosine >= 0.999 against the oracle
    (B200_GEMM_PAIR is read once per process, hence the subprocesses), and agree with each other."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for pair in ("0", "1"):
        f = str(tmp_path / f"emb_pair{pair}.npy")
        r = subprocess.run([sys.executable, "-c", _GEMM_VARIANT, f], cwd=root, env=dict(os.environ, B200_GEMM_PAIR=pair),
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-2000:] + r.stderr[-2000:]
        outs.append(np.load(f))
    assert (outs[0] * outs[1]).sum(1).min() > 0.9999


def test_encoder_exact_gelu_variant():
    """LAION open_clip checkpoints use exact GELU instead of QuickGELU: same gate, other epilogue."""
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    cfg = dict(VIT_CONFIGS["ViT-tiny-test"], activation="gelu")
    w = synthetic_weights(cfg, seed=5)
    eng = Engine(cfg, device=0, max_batch=8)
    eng.load_weights(w)
    imgs = [synth_image(60 + i, 70 + i, 90 - i) for i in range(6)]
    emb = eng.embed_host(imgs, mode=1)
    ref = O.vit_encode(w, O.preprocess_batch(imgs, cfg["image_size"], 1), cfg)
    assert (emb * ref).sum(1).min() >= 0.999
    quick = O.vit_encode(w, O.preprocess_batch(imgs, cfg["image_size"], 1), dict(cfg, activation="quick_gelu"))
    assert (emb * quick).sum(1).min() < (emb * ref).sum(1).min()     # the two activations are distinguishable
    eng.close()


def test_embed_stream_matches_blocking_calls():
    """Pipelined host-buffer path (upload of batch i+1 overlapping the encode of batch i): same vectors, same order."""
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    from lumen_b200.engine import pack_images
    name = "ViT-tiny-test"; cfg = VIT_CONFIGS[name]
    be = B200ClipBackend(SyntheticResources(model_name=name), device_preference="cuda:0", max_batch_size=16, weights=synthetic_weights(cfg, 0))
    be.initialize()
    rng = np.random.default_rng(31)
    batches, want = [], []
    for n in (5, 16, 1, 9, 12):
        imgs = [rng.integers(0, 256, (rng.integers(40, 120), rng.integers(40, 120), 3), dtype=np.uint8) for _ in range(n)]
        want.append(be.embed_arrays(imgs))
        buf, off, hw, total = pack_images(imgs)
        pinned = torch.from_numpy(buf[:total].copy()).pin_memory()
        batches.append((pinned, total, off, hw))
    got = list(be.embed_stream(iter(batches)))
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g.shape == w.shape and np.array_equal(g, w)
    assert list(be.embed_stream(iter([]))) == []
    be.close()
