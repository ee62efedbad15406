"""N4 — device gallery: fp32 .npy -> fp16 rows on the device, incremental append, on-disk round trip, search."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")


def test_oracle_gallery_pack_is_the_reference_normalisation():
    from oracle import clip_oracle as O
    t = np.random.default_rng(0).standard_normal((7, 16)).astype(np.float32) * 3
    g = O.gallery_pack(t)
    np.testing.assert_allclose(np.linalg.norm(g.astype(np.float32), axis=1), 1.0, atol=2e-3)
    assert np.array_equal(O.gallery_pack(t, normalize=False), t.astype(np.float16))


def _ulp_diff_f16(a: np.ndarray, b: np.ndarray) -> int:
    ia, ib = a.view(np.int16).astype(np.int32), b.view(np.int16).astype(np.int32)
    ia = np.where(ia < 0, -32768 - ia, ia)
    ib = np.where(ib < 0, -32768 - ib, ib)
    return int(np.abs(ia - ib).max())


@pytest.mark.gpu
def test_gallery_from_npy_append_save_load_search(tmp_path):
    from lumen_b200.engine import Engine
    from lumen_b200.gallery import DeviceGallery
    from oracle import clip_oracle as O
    rng = np.random.default_rng(5)
    N, D = 30011, 512
    t = (rng.standard_normal((N, D)) * rng.uniform(0.1, 9.0, (N, 1))).astype(np.float32)
    np.save(tmp_path / "gallery_f32.npy", t)
    eng = Engine({}, device=0)
    DeviceGallery.CHUNK_BYTES = 8 << 20                     # several chunks
    g = DeviceGallery.from_npy(eng, tmp_path / "gallery_f32.npy")
    eng.sync()
    want = O.gallery_pack(t)
    got = g.view().cpu().numpy()
    assert got.shape == want.shape and _ulp_diff_f16(got, want) <= 1     # fp32 norm summed in a different order: at most one fp16 ulp
    assert (got != want).mean() < 0.02
    # un-normalised conversion is exact (round to nearest even, like numpy)
    raw = DeviceGallery(eng, D)
    raw.append(t[:1000], normalize=False)
    assert np.array_equal(raw.view().cpu().numpy(), t[:1000].astype(np.float16))
    # incremental append: indices are stable, earlier rows survive the reallocation, device rows are taken as they are
    inc = DeviceGallery(eng, D, capacity=16, index_base=100)
    first, n = inc.append(t[:5000])
    assert (first, n) == (100, 5000)
    first, n = inc.append(torch.from_numpy(t[5000:12000]).cuda())
    assert (first, n) == (5100, 7000)
    first, n = inc.append(g.view()[12000:12010])                          # fp16 device rows
    assert (first, n, inc.rows) == (12100, 10, 12010)
    eng.sync()
    assert torch.equal(inc.view(), g.view()[:12010])
    # on-disk round trip: plain fp16 .npy + json sidecar
    g.save(tmp_path / "gallery_f16.npy")
    assert np.array_equal(np.load(tmp_path / "gallery_f16.npy", mmap_mode="r"), got)
    g2 = DeviceGallery.load(eng, tmp_path / "gallery_f16.npy")
    assert g2.rows == N and g2.index_base == 0 and torch.equal(g2.view(), g.view())
    # search on it = the oracle's cosine top-k over the same fp16 rows
    q = torch.from_numpy(O.gallery_pack(rng.standard_normal((33, D)).astype(np.float32))).cuda()
    s, i = g2.search(q, 10)
    eng.sync()
    ws, wi = O.topk_cosine(q.cpu().numpy().astype(np.float32), got.astype(np.float32), 10)
    assert np.array_equal(i.cpu().numpy(), wi)
    np.testing.assert_allclose(s.cpu().numpy(), ws, rtol=0, atol=2e-3)
    # row shards of two ranks = the single-GPU gallery, with global indices
    a = DeviceGallery.from_npy(eng, tmp_path / "gallery_f32.npy", rank=0, world=2)
    b = DeviceGallery.from_npy(eng, tmp_path / "gallery_f32.npy", rank=1, world=2)
    assert a.index_base == 0 and b.index_base == a.rows and a.rows + b.rows == N
    assert torch.equal(torch.cat([a.view(), b.view()]), g.view())
    eng.close()
