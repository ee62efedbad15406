"""CPU: the reference's own `create_backend` (backends/factory.py:75-141) builds a B200ClipBackend once
`install_as_reference_backend` has run - SURVEY §8b "B-factory".  Needs the reference tree (build container only)."""
import sys
from pathlib import Path

import pytest


@pytest.fixture()
def reference_factory(monkeypatch):
    ref = Path("/root/reference/packages")
    if not ref.exists():
        pytest.skip("reference tree not mounted")
    for p in (ref / "lumen-clip" / "src", ref / "lumen-resources" / "src"):
        monkeypatch.syspath_prepend(str(p))
    try:
        from lumen_clip.backends import factory
        from lumen_resources.lumen_config import BackendSettings
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference factory not importable here: {e}")
    monkeypatch.setattr(factory, "_BACKEND_REGISTRY", {}, raising=False)
    saved = {k: sys.modules.get(k) for k in ("lumen_clip.backends.torch_backend", "lumen_clip.backends.onnxrt_backend")}
    yield factory, BackendSettings
    for k, v in saved.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v


@pytest.mark.parametrize("runtime", ["torch", "onnx"])
def test_reference_create_backend_builds_the_b200_backend(reference_factory, runtime):
    factory, BackendSettings = reference_factory
    from lumen_b200.backend import B200ClipBackend, SyntheticResources, install_as_reference_backend
    from lumen_b200.interface import BaseClipBackend
    if runtime == "onnx":
        # the factory only offers `onnx` when onnxruntime can be found; the deployment this is for has it, this container does not
        import importlib.util
        real = importlib.util.find_spec
        import unittest.mock as mock
        patcher = mock.patch.object(importlib.util, "find_spec", lambda name, *a, **k: object() if name == "onnxruntime" else real(name, *a, **k))
        patcher.start()
    try:
        cls = install_as_reference_backend(runtime, preprocess_mode="squash")
        resources = SyntheticResources(model_name="ViT-B-32")
        settings = BackendSettings(device="cuda", batch_size=64)
        backend = factory.create_backend(backend_config=settings, resources=resources, runtime=runtime, precision="fp16")
    finally:
        if runtime == "onnx":
            patcher.stop()
    assert isinstance(backend, B200ClipBackend) and isinstance(backend, cls) and isinstance(backend, BaseClipBackend)
    assert backend.resources is resources
    assert backend._max_batch_size == 64 and backend._device_index == 0
    assert backend.requested_precision == ("fp16" if runtime == "onnx" else None)   # factory.py:118-132: only the ONNX branch forwards it
    assert not backend.is_initialized                         # the factory constructs, the service initialises
    info = backend.get_info()
    assert info.runtime == "b200" and info.image_embedding_dim == 512 and info.supports_image_batch
    with pytest.raises(RuntimeError):
        backend.image_to_vector(b"x")                         # used before initialize(): the reference's error contract
    with pytest.raises(ValueError):
        factory.create_backend(backend_config=settings, resources=resources, runtime="b200")   # the runtime enum stays closed


def test_install_rejects_unknown_runtime():
    from lumen_b200.backend import install_as_reference_backend
    with pytest.raises(ValueError):
        install_as_reference_backend("rknn")
