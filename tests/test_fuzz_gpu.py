"""GPU: differential fuzz of the other kernels against the oracle and, where it is installed, against the dependency itself (Pillow).
Seeded, so every run sees the same cases; the search has its own file (test_search_fuzz_gpu.py).  The K1 fuzz is what found
PIL.Image.resize's vertical-first branch for sources more than 100 times taller than wide."""
import io

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_towers_of_random_shape_vs_oracle():
    """Width, depth, heads, MLP size, patch, image size, activation and batch drawn at random: every GEMM tile variant (64 / 128 / 192 /
    256 columns, single CTA and CTA pair), both attention kernels, ragged row / column tails, the CLS-only last layer."""
    from lumen_b200 import synthetic_weights
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    rng = np.random.default_rng(1)
    for case in range(10):
        heads = int(rng.choice([2, 4, 6, 8, 12, 16])); width = heads * 64
        patch = int(rng.choice([14, 16, 32]))
        grid = int(rng.choice([4, 8, 12, 16])) if patch == 14 else int(rng.choice([1, 2, 3, 5, 7, 9, 14]))
        S = patch * grid
        cfg = dict(image_size=S, patch=patch, width=width, layers=int(rng.choice([1, 2, 3])), heads=heads,
                   mlp_dim=int(rng.choice([64, 192, 256, 1024, 1536, 3072])), embed_dim=int(rng.choice([4, 64, 100, 512, 768])),
                   activation=str(rng.choice(["quick_gelu", "gelu"])), ln_eps=1e-5)
        B = int(rng.choice([1, 2, 5, 17, 64, 130, 300]))
        B = max(1, min(B, 30000 // (grid * grid + 1)))
        w = synthetic_weights(cfg, seed=case)
        imgs = [rng.integers(0, 256, (S, S, 3), dtype=np.uint8) for _ in range(B)]
        eng = Engine(cfg, device=0, max_batch=B)
        eng.load_weights(w)
        e = eng.embed_host(imgs)
        pick = list(range(0, B, max(1, B // 6)))[:6]
        ref = O.vit_encode(w, O.preprocess_batch([imgs[j] for j in pick], S, 0), cfg)
        cos = (e[pick] * ref).sum(1)
        assert cos.min() >= 0.999, (cfg, B, cos)                       # BASELINE.json tolerance
        assert np.isfinite(e).all() and np.allclose(np.linalg.norm(e, axis=1), 1.0, atol=1e-4), (cfg, B)
        eng.close()


def test_preprocess_random_sizes_vs_oracle_and_pillow():
    from lumen_b200.engine import Engine, pack_images
    from oracle import clip_oracle as O
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(3)
    for S, patch in ((224, 32), (64, 16)):
        eng = Engine({"patch": patch}, device=0, image_size=S)
        sizes = [(1, 1), (1, 2), (2, 1), (1, 500), (500, 1), (2, 2), (3, 3), (S - 1, S + 1), (S + 1, S - 1), (7, 1000), (1000, 7), (2300, 19)]
        sizes += [(int(rng.integers(1, 900)), int(rng.integers(1, 900))) for _ in range(30)]
        imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in sizes]
        buf, offsets, hw, total = pack_images(imgs)
        src = torch.from_numpy(buf[:total + 16].copy()).cuda()
        for mode in (0, 1):
            got = eng.preprocess(src, offsets, hw, mode).cpu().numpy()
            eng.sync()
            want = O.preprocess_batch(imgs, S, mode)
            for j, hw_j in enumerate(sizes):
                assert np.array_equal(got[j], want[j]), (S, mode, hw_j)
        for j in (0, 3, 4, 9, 10, 11, 20, 30):                          # the oracle itself against Pillow on the odd ones
            pil = np.asarray(Image.fromarray(imgs[j]).resize((S, S), Image.Resampling.BICUBIC))
            assert np.array_equal(pil, O.resize_bicubic(imgs[j], S, S)), sizes[j]
        eng.close()


def test_device_jpeg_decoder_vs_live_pillow():
    """Random sizes (down to 1 x 1), qualities 1-100, all three chroma layouts, grey-scale, optimised tables, restart intervals,
    progressive scans: an image may be handed back to the host decoder (status 1), a decoded one must equal Pillow's pixels."""
    from lumen_b200.engine import Engine
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(1)
    datas, refs, tags = [], [], []
    small = [1, 2, 7, 8, 9, 15, 16, 17, 31, 33, 64, 100, 223, 224, 500]
    for n in range(150):
        h = int(rng.choice(small)) if rng.random() < 0.5 else int(rng.integers(1, 700))
        w = int(rng.choice(small)) if rng.random() < 0.5 else int(rng.integers(1, 700))
        kind = int(rng.integers(0, 3))
        if kind == 0:
            a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        elif kind == 1:
            a = np.full((h, w, 3), int(rng.integers(0, 256)), np.uint8)
        else:
            yy, xx = np.mgrid[0:h, 0:w]
            a = np.stack([128 + 100 * np.sin(xx / 9.0), 128 + 100 * np.cos(yy / 7.0), (xx * 3 + yy * 5) % 256], -1).clip(0, 255).astype(np.uint8)
        gray = rng.random() < 0.15
        im = Image.fromarray(a).convert("L") if gray else Image.fromarray(a)
        kw = dict(quality=int(rng.integers(1, 101)))
        if not gray: kw["subsampling"] = int(rng.choice([0, 1, 2]))
        if rng.random() < 0.3: kw["optimize"] = True
        if rng.random() < 0.15: kw["progressive"] = True
        if rng.random() < 0.2: kw["restart_marker_blocks"] = int(rng.integers(1, 20))
        buf = io.BytesIO()
        try:
            im.save(buf, "JPEG", **kw)
        except OSError:                                                 # a few option combinations Pillow's encoder itself refuses
            continue
        datas.append(buf.getvalue())
        refs.append(np.asarray(Image.open(io.BytesIO(datas[-1])).convert("RGB")))
        tags.append((h, w, gray, kw))
    eng = Engine({}, device=0)
    dev, off, hw, status, used = eng.decode_jpeg(datas)
    eng.sync()
    host = dev[:used].cpu().numpy()
    decoded = 0
    for i, ref in enumerate(refs):
        if status[i] != 0:
            continue
        decoded += 1
        assert tuple(hw[i]) == ref.shape[:2], tags[i]
        assert np.array_equal(host[off[i]:off[i] + ref.size].reshape(ref.shape), ref), tags[i]
    assert decoded >= 100                                              # (progressive and a few exotic ones go back to the host)
    eng.close()


def test_classify_random_shapes_vs_oracle():
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    rng = np.random.default_rng(1)
    eng = Engine({}, device=0)
    for case in range(24):
        N = int(rng.choice([1, 2, 5, 8, 100, 999, 1000, 1001, 3000, 20000])); D = int(rng.choice([64, 100, 512, 768])); B = int(rng.choice([1, 3, 64, 257]))
        k = int(min(N, rng.choice([1, 3, 5, 10, 50]))); soft = bool(rng.random() < 0.7); scale = float(rng.choice([1.0, 100.0]))
        T = rng.standard_normal((N, D)).astype(np.float32) * rng.uniform(0.1, 5.0, (N, 1)).astype(np.float32)
        if N > 10 and rng.random() < 0.5:
            T[N // 2:N // 2 + 3] = T[:3]                               # exact duplicate labels -> index-ordered ties
        X = rng.standard_normal((B, D)).astype(np.float32)
        eng.set_labels(T)
        vals, idx = eng.classify(torch.from_numpy(X).cuda(), k, scale=scale, softmax=soft)
        eng.sync()
        vals, idx = vals.cpu().numpy(), idx.cpu().numpy()
        tag = dict(N=N, D=D, B=B, k=k, soft=soft, scale=scale)
        for r in range(min(B, 6)):
            wi, wv = O.classify(X[r], T, k, scale, soft)
            np.testing.assert_allclose(vals[r], wv, rtol=2e-3, atol=1e-6, err_msg=str(tag))
            if not np.array_equal(idx[r], wi):                         # swaps among near-equal values only
                fi, fv = O.classify(X[r], T, N, scale, soft)
                vmap = dict(zip(fi.tolist(), fv.tolist()))
                assert all(abs(vmap[int(a)] - vmap[int(b)]) <= 2e-3 * abs(vmap[int(b)]) + 1e-7 for a, b in zip(idx[r], wi)), tag
    eng.close()
