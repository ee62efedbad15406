"""P0 / N2 — JPEG decode.  CPU: the C restatement (oracle/jpeg_decode.c) against pixels Pillow itself produced
(tests/golden/jpeg.npz), and the host-only header probe of the product.  GPU: the device decoder, through the C ABI, bit for bit
against the goldens, the oracle and (when importable) Pillow on larger images; and the bytes -> embedding path."""
import io
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).parent / "golden" / "jpeg.npz"


@pytest.fixture(scope="module")
def golden_jpeg():
    g = np.load(GOLD)
    return [(str(n), g[f"bytes_{i}"].tobytes(), g[f"rgb_{i}"]) for i, n in enumerate(g["names"])]


def test_oracle_matches_pillow_goldens(golden_jpeg):
    from oracle import clip_oracle as O
    for name, data, rgb in golden_jpeg:
        out = O.decode_jpeg(data)
        if name.startswith("refuse"):
            assert out is None, name
        else:
            assert np.array_equal(out, rgb), name


def test_oracle_rejects_corrupt_stream(golden_jpeg):
    from oracle import clip_oracle as O
    name, data, _ = next(c for c in golden_jpeg if c[0].startswith("64x48") and "420" in c[0])
    with pytest.raises(ValueError):
        O.decode_jpeg(data[:len(data) // 8])              # headers cut in the middle of a segment
    with pytest.raises(ValueError):
        O.decode_jpeg(data[:len(data) - 40])              # entropy-coded data cut: no EOI (Pillow: "image file is truncated")


def test_oracle_matches_live_pillow_on_random_cases():
    """Beyond the committed fixtures: when Pillow is importable (it is in this image), fresh random sizes / qualities / samplings."""
    Image = pytest.importorskip("PIL.Image")
    from oracle import clip_oracle as O
    rng = np.random.default_rng(123)
    for case in range(40):
        h, w = int(rng.integers(1, 200)), int(rng.integers(1, 200))
        yy, xx = np.mgrid[0:h, 0:w]
        a = np.stack([(xx * int(rng.integers(1, 9)) + yy) % 256, rng.integers(0, 256, (h, w)), 128 + 100 * np.sin((xx + yy) / float(rng.uniform(2, 30)))], -1)
        kw = dict(quality=int(rng.integers(5, 101)), subsampling=int(rng.integers(0, 3)))
        if case % 5 == 0:
            kw["optimize"] = True
        if case % 7 == 0:
            kw["restart_marker_blocks"] = int(rng.integers(1, 9))
        buf = io.BytesIO()
        Image.fromarray(a.clip(0, 255).astype(np.uint8)).save(buf, "JPEG", **kw)
        want = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"))
        assert np.array_equal(O.decode_jpeg(buf.getvalue()), want), (case, h, w, kw)


def test_probe_is_host_only_and_matches_headers(golden_jpeg):
    """b200clip_jpeg_probe parses headers on the host: callable without a GPU."""
    from lumen_b200.engine import Engine
    for name, data, rgb in golden_jpeg:
        info = Engine.jpeg_probe(data)
        if name.startswith("refuse"):
            assert info is None, name
        else:
            assert info[:2] == rgb.shape[:2], name
            assert info[2] == (1 if "gray" in name else 3), name


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_device_decode_bit_exact_vs_goldens(golden_jpeg):
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    for threads in (1, 0):
        datas = [c[1] for c in golden_jpeg]
        dev, off, hw, status, used = eng.decode_jpeg(datas, host_threads=threads)
        eng.sync()
        host = dev[:used].cpu().numpy()
        for i, (name, _, rgb) in enumerate(golden_jpeg):
            if name.startswith("refuse"):
                assert status[i] == 1 and tuple(hw[i]) == (0, 0), name
                continue
            assert status[i] == 0 and tuple(hw[i]) == rgb.shape[:2] and off[i] % 16 == 0, name
            got = host[off[i]:off[i] + rgb.size].reshape(rgb.shape)
            assert np.array_equal(got, rgb), f"{name}: {np.count_nonzero(got != rgb)} bytes differ"
    eng.close()


@pytest.mark.gpu
def test_device_decode_large_images_vs_oracle_and_pillow():
    from lumen_b200.engine import Engine
    from oracle import clip_oracle as O
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(3)
    datas, refs = [], []
    for (h, w, sub, q) in ((480, 640, 2, 85), (1080, 1920, 2, 90), (777, 1023, 1, 70), (600, 401, 0, 95), (1200, 1600, 2, 50)):
        yy, xx = np.mgrid[0:h, 0:w]
        a = np.stack([128 + 90 * np.sin(xx / 31.0) + rng.normal(0, 12, (h, w)), 128 + 90 * np.cos(yy / 17.0) + rng.normal(0, 12, (h, w)),
                      128 + 60 * np.sin((xx + yy) / 23.0) + rng.normal(0, 12, (h, w))], -1).clip(0, 255).astype(np.uint8)
        buf = io.BytesIO()
        Image.fromarray(a).save(buf, "JPEG", quality=q, subsampling=sub)
        datas.append(buf.getvalue())
        refs.append(np.asarray(Image.open(io.BytesIO(datas[-1])).convert("RGB")))
        assert np.array_equal(O.decode_jpeg(datas[-1]), refs[-1])
    eng = Engine({}, device=0)
    dev, off, hw, status, used = eng.decode_jpeg(datas)
    eng.sync()
    host = dev[:used].cpu().numpy()
    for i, ref in enumerate(refs):
        assert status[i] == 0
        assert np.array_equal(host[off[i]:off[i] + ref.size].reshape(ref.shape), ref), i
    # a second, smaller batch reuses the arenas
    dev, off, hw, status, used = eng.decode_jpeg(datas[:1])
    eng.sync()
    assert np.array_equal(dev[off[0]:off[0] + refs[0].size].cpu().numpy().reshape(refs[0].shape), refs[0])
    # corrupt baseline stream: the call fails loudly
    from lumen_b200 import _lib
    with pytest.raises(_lib.B200ClipError):
        eng.decode_jpeg([datas[0][:len(datas[0]) // 2]])       # cut in the middle of the entropy-coded data: no EOI
    eng.close()


@pytest.mark.gpu
def test_bytes_to_embeddings_equals_decoded_path(golden_jpeg):
    """image_batch_to_vectors on bytes: device-decoded JPEGs, a PNG through the host decoder, same embeddings as feeding the
    reference's own decoded arrays."""
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.backend import B200ClipBackend
    from lumen_b200.engine import Engine
    cfg = VIT_CONFIGS["ViT-tiny-test"]
    eng = Engine(cfg, device=0, max_batch=8)
    eng.load_weights(synthetic_weights(cfg, seed=0))
    pick = [c for c in golden_jpeg if c[0].startswith(("100x150", "120x160", "17x23"))][:11] + [c for c in golden_jpeg if c[0] == "refuse_png"]
    assert pick[-1][0] == "refuse_png" and len(pick) == 12
    want = eng.embed_host([c[2] for c in pick], mode=0)
    got = eng.embed_bytes([c[1] for c in pick], mode=0, host_decode=B200ClipBackend._decode)
    assert np.array_equal(got, want)
    from lumen_b200 import _lib
    with pytest.raises(_lib.B200ClipError):
        eng.embed_bytes([pick[-1][1]], mode=0)               # not a JPEG and no host decoder given
    eng.close()
