"""CPU: the C-ABI library builds, loads and exports every symbol include/b200clip.h declares; host logic."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def lib():
    from lumen_b200 import build, _lib
    build.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    header = (ROOT / "include" / "b200clip.h").read_text()
    declared = set(re.findall(r"\b(b200clip_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 18
    from lumen_b200 import _lib
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert getattr(lib, name) is not None


def test_create_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from lumen_b200 import _lib
    from lumen_b200.engine import Engine
    with pytest.raises(_lib.B200ClipError) as e:
        Engine({}, device=0)
    assert e.value.code == _lib.EDEVICE and "no CPU fallback" in e.value.message


def test_missing_library_is_an_error(monkeypatch, tmp_path):
    from lumen_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_pack_images_alignment():
    from lumen_b200.engine import pack_images
    rng = np.random.default_rng(0)
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in ((3, 5), (7, 7), (1, 1))]
    buf, off, hw, total = pack_images(imgs)
    assert (off % 16 == 0).all() and total == off[-1] + 16 + 16
    for im, o in zip(imgs, off):
        assert np.array_equal(buf[o:o + im.size], im.reshape(-1))
    assert hw.tolist() == [[3, 5], [7, 7], [1, 1]]
    with pytest.raises(ValueError):
        pack_images([np.zeros((4, 4), np.uint8)])


def test_backend_error_contract_without_init():
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    from lumen_b200.interface import InferenceError
    be = B200ClipBackend(SyntheticResources(model_name="ViT-tiny-test"))
    assert not be.is_initialized
    with pytest.raises(RuntimeError, match="not initialized"):
        be.image_to_vector(b"abc")
    assert be.image_batch_to_vectors([]).shape == (0, 0)          # onnxrt_backend.py:499-500
    with pytest.raises(InferenceError):
        be.text_to_vector("a cat")
    info = be.get_info()
    assert info.runtime == "b200" and info.supports_image_batch and info.image_embedding_dim == 64


def test_shard_bounds_cover_gallery():
    from lumen_b200.managers import ShardedGallery
    for n, w in ((10_000_000, 8), (1001, 4), (3, 8)):
        spans = [ShardedGallery.shard_bounds(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
