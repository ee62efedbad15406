"""GPU: b200clip_clone - a second handle sharing the weights - and the two-batches-in-flight stream built on it."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_clone_shares_weights_and_runs_concurrently():
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200._lib import B200ClipError
    from lumen_b200.engine import Engine, pack_images
    from oracle import clip_oracle as O
    cfg = VIT_CONFIGS["ViT-tiny-test"]
    w = synthetic_weights(cfg, 3)
    eng = Engine(cfg, device=0, max_batch=16)
    with pytest.raises(B200ClipError):
        eng.clone()                                                    # weights first
    eng.load_weights(w)
    twin = eng.clone()
    rng = np.random.default_rng(0)
    imgs = [rng.integers(0, 256, (70 + i, 90, 3), dtype=np.uint8) for i in range(12)]
    a, b = eng.embed_host(imgs), twin.embed_host(imgs)
    assert np.array_equal(a, b)                                         # same weights, same arithmetic
    ref = O.vit_encode(w, O.preprocess_batch(imgs, cfg["image_size"], 0), cfg)
    assert ((a * ref).sum(1) >= 0.999).all()
    with pytest.raises(B200ClipError):                                  # weights go in through the source only
        twin.load_weights(w)
    # two streams, two handles, interleaved batches: every result still equals the sequential one
    buf, offsets, hw, total = pack_images(imgs)
    src = torch.from_numpy(buf[:total + 16].copy()).cuda()
    outs = [torch.empty((12, cfg["embed_dim"]), device="cuda") for _ in range(2)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for s in streams:
        s.wait_stream(torch.cuda.current_stream())
    for i in range(20):
        k = i & 1
        with torch.cuda.stream(streams[k]):
            (eng, twin)[k].embed_u8(src, offsets, hw, 0, out=outs[k])
    torch.cuda.synchronize()
    assert np.array_equal(outs[0].cpu().numpy(), a) and np.array_equal(outs[1].cpu().numpy(), a)
    # the source can go first: the weights live until the last handle using them is closed
    eng.close()
    assert np.array_equal(twin.embed_host(imgs), a)
    twin.close()


def test_embed_stream_with_twin_equals_single_engine(monkeypatch):
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.engine import Engine, pack_images
    cfg = VIT_CONFIGS["ViT-tiny-test"]
    eng = Engine(cfg, device=0, max_batch=8)
    eng.load_weights(synthetic_weights(cfg, 4))
    rng = np.random.default_rng(1)
    items, want = [], []
    for n in range(7):                                                  # ragged batches, odd count
        imgs = [rng.integers(0, 256, (40 + 3 * n, 50 + i, 3), dtype=np.uint8) for i in range(1 + n % 5)]
        buf, offsets, hw, total = pack_images(imgs)
        pinned = torch.from_numpy(buf[:total + 16].copy()).pin_memory()
        items.append((pinned, total, offsets, hw))
        want.append(eng.embed_host(imgs))
    got = list(eng.embed_stream(iter(items)))
    assert len(got) == len(want) and all(np.array_equal(g, w) for g, w in zip(got, want))
    # a consumer that walks away after two results leaves nothing running behind it
    gen = eng.embed_stream(iter(items))
    first = [next(gen), next(gen)]
    gen.close()
    assert np.array_equal(first[0], want[0]) and np.array_equal(first[1], want[1])
    assert all(np.array_equal(g, w) for g, w in zip(eng.embed_stream(iter(items)), want))
    monkeypatch.setenv("B200_STREAM_TWIN", "0")
    got1 = list(eng.embed_stream(iter(items)))
    assert all(np.array_equal(g, w) for g, w in zip(got1, want))
    eng.close()
