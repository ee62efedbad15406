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


def _dht_segment(cls_id: int, counts, symbols=b"") -> bytes:
    body = bytes([cls_id]) + bytes(counts) + bytes(symbols)
    return b"\xff\xc4" + (len(body) + 2).to_bytes(2, "big") + body


def test_probe_survives_malformed_headers(golden_jpeg):
    """Client bytes reach the header parser unfiltered (image_to_vector -> embed_bytes -> jpeg_probe): crafted or damaged segments
    must be refused, never crash.  The first case is a 280-byte file whose DHT over-subscribes length 1 (255 codes of one bit)."""
    from lumen_b200.engine import Engine
    soi = b"\xff\xd8"
    counts = [0] * 16; counts[0] = 255
    assert Engine.jpeg_probe(soi + _dht_segment(0x00, counts, bytes(255)) + b"\xff\xd9") is None
    for bad_len in range(1, 17):                                 # every single over-subscribed length, DC and AC tables
        counts = [0] * 16; counts[bad_len - 1] = min(255, (1 << bad_len) + 1)
        for cls_id in (0x00, 0x10, 0x03, 0x13):
            assert Engine.jpeg_probe(soi + _dht_segment(cls_id, counts, bytes(counts[bad_len - 1]))) is None
    counts = [16] * 16                                           # 256 symbols but not a prefix code
    assert Engine.jpeg_probe(soi + _dht_segment(0x00, counts, bytes(256))) is None
    name, data, rgb = next(c for c in golden_jpeg if c[0].startswith("64x48") and "420" in c[0])
    assert Engine.jpeg_probe(data)[:2] == rgb.shape[:2]
    sos = data.index(b"\xff\xda")
    assert Engine.jpeg_probe(data[:sos] + b"\xff\xda\x00\x02") is None          # SOS with no payload at the end of the buffer
    assert Engine.jpeg_probe(data[:sos] + b"\xff\xda\x00\x03\x07") is None      # ... and with a component count only
    sof = data.index(b"\xff\xc0")
    huge = bytearray(data); huge[sof + 5:sof + 9] = b"\xff\xff\xff\xff"        # 65535 x 65535: beyond PIL's pixel limit
    assert Engine.jpeg_probe(bytes(huge)) is None
    rng = np.random.default_rng(2024)
    hdr = sos + 14
    for it in range(3000):                                       # random damage confined to the header segments
        b = bytearray(data)
        for _ in range(int(rng.integers(1, 6))):
            b[int(rng.integers(2, hdr))] = int(rng.integers(0, 256))
        if it % 3 == 0:
            b = b[:int(rng.integers(2, hdr))]
        info = Engine.jpeg_probe(bytes(b))
        assert info is None or (0 < info[0] <= 65535 and 0 < info[1] <= 65535 and info[2] in (1, 3))


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
    # a damaged baseline stream (cut in the middle of the entropy-coded data: no EOI) is handed back to the caller's decoder,
    # per image; its neighbours in the batch are decoded as usual
    dev, off, hw, status, used = eng.decode_jpeg([datas[0], datas[0][:len(datas[0]) // 2], datas[3]])
    eng.sync()
    assert status.tolist() == [0, 1, 0] and tuple(hw[1]) == (0, 0)
    host = dev[:used].cpu().numpy()
    for i, ref in ((0, refs[0]), (2, refs[3])):
        assert np.array_equal(host[off[i]:off[i] + ref.size].reshape(ref.shape), ref), i
    eng.close()


@pytest.mark.gpu
def test_device_decode_survives_damaged_streams(golden_jpeg):
    """Random damage anywhere in the file (headers, tables, entropy-coded data): every image comes back as decoded or as
    'left to the host decoder', the batch call itself never fails and nothing crashes."""
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    rng = np.random.default_rng(99)
    base = [c[1] for c in golden_jpeg if not c[0].startswith("refuse")]
    datas = []
    for it in range(400):
        b = bytearray(base[it % len(base)])
        for _ in range(int(rng.integers(1, 8))):
            b[int(rng.integers(2, len(b)))] = int(rng.integers(0, 256))
        if it % 5 == 0:
            b = b[:int(rng.integers(2, len(b)))]
        datas.append(bytes(b))
    datas.append(base[0])
    dev, off, hw, status, used = eng.decode_jpeg(datas)
    eng.sync()
    assert set(status.tolist()) <= {0, 1} and status[-1] == 0 and (status[:-1] == 1).sum() > 20
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
    # a JPEG cut short: the device decoder hands it back, PIL refuses it ("image file is truncated") -> the call raises, as the
    # reference's Image.open(...).convert("RGB") does
    cut = pick[0][1][:len(pick[0][1]) - 40]
    with pytest.raises(Exception):
        eng.embed_bytes([pick[1][1], cut], mode=0, host_decode=B200ClipBackend._decode)
    with pytest.raises(_lib.B200ClipError):
        eng.embed_bytes([cut], mode=0)
    eng.close()
