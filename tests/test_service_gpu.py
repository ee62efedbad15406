"""SURVEY §8f N3 with a REAL backend: the request coalescer in front of `B200ClipBackend`, driven the way the reference's server
drives a backend - ten worker threads (packages/lumen-clip/src/lumen_clip/server.py:294-297), one image per request
(general_clip/clip_service.py:352-366).  The second test goes through the reference's own `GeneralCLIPService.Infer` and needs both a
GPU and the reference tree: it is skipped wherever one of them is missing (the GPU box has no reference tree, the build container
has no GPU) and is what a deployment would run."""
import io
import threading

import numpy as np
import pytest

from conftest import synth_image

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _blob(img, fmt):
    from PIL import Image
    bio = io.BytesIO()
    Image.fromarray(img, "RGB").save(bio, format=fmt, **({"quality": 90} if fmt == "JPEG" else {}))
    return bio.getvalue()


@pytest.fixture(scope="module")
def real_backend():
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    cfg = VIT_CONFIGS["ViT-tiny-test"]
    be = B200ClipBackend(SyntheticResources(model_name="ViT-tiny-test"), device_preference="cuda:0", max_batch_size=64, weights=synthetic_weights(cfg, 0))
    be.initialize()
    yield be
    be.close()


def test_ten_threads_through_the_coalescer(real_backend):
    from lumen_b200.interface import InferenceError, InvalidInputError
    from lumen_b200.service import BatchedBackendView
    view = BatchedBackendView(real_backend, max_batch=64, max_wait_ms=20)
    blobs = [_blob(synth_image(300 + i, 60 + 7 * i, 90 - 3 * i), "JPEG" if i % 2 else "PNG") for i in range(10)]
    want = real_backend.image_batch_to_vectors(blobs)
    out, errs = {}, {}

    def work(i, payload):
        try:
            out[i] = view.image_to_vector(payload)
        except Exception as e:  # noqa: BLE001
            errs[i] = e

    th = [threading.Thread(target=work, args=(i, b)) for i, b in enumerate(blobs)]
    [t.start() for t in th]; [t.join(timeout=60) for t in th]
    assert not errs and len(out) == 10
    for i in range(10):
        assert out[i].dtype == np.float32 and out[i].shape == (64,) and out[i].flags["C_CONTIGUOUS"]
        assert float(out[i] @ want[i]) > 0.99999                     # the same vector the direct batched call gives
    assert view._batcher.batches_run < 10                            # ten requests, fewer GPU passes
    # a damaged image among good ones: only its caller sees an error
    n0 = view._batcher.batches_run
    payloads = blobs[:4] + [blobs[1][: len(blobs[1]) // 2]] + blobs[4:8]
    out.clear(); errs.clear()
    th = [threading.Thread(target=work, args=(i, b)) for i, b in enumerate(payloads)]
    [t.start() for t in th]; [t.join(timeout=60) for t in th]
    assert list(errs) == [4] and isinstance(errs[4], InferenceError) and len(out) == 8
    assert view._batcher.batches_run > n0
    with pytest.raises(InvalidInputError):
        view.image_to_vector(b"")


def test_infer_rpc_through_the_reference_service(real_backend):
    pytest.importorskip("grpc")
    import json
    import sys
    from pathlib import Path
    ref = Path("/root/reference/packages")
    if not ref.exists():
        pytest.skip("reference tree not on this machine")
    for p in (ref / "lumen-clip" / "src", ref / "lumen-resources" / "src"):
        if str(p) not in sys.path:
            sys.path.insert(0, str(p))
    try:
        from lumen_b200.service import b200_clip_service_class
        from lumen_clip.proto import ml_service_pb2 as pb
        cls = b200_clip_service_class()
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference service not importable here: {e}")
    svc = cls(real_backend, real_backend.resources)
    svc.initialize()
    blobs = [_blob(synth_image(400 + i, 64, 64), "JPEG") for i in range(10)]
    results = {}

    def call(i):
        req = pb.InferRequest(correlation_id=f"c{i}", task="clip_image_embed", payload=blobs[i], payload_mime="image/jpeg")
        results[i] = list(svc.Infer(iter([req]), context=None))

    th = [threading.Thread(target=call, args=(i,)) for i in range(10)]
    [t.start() for t in th]; [t.join(timeout=60) for t in th]
    for i in range(10):
        (resp,) = results[i]
        assert resp.is_final and not resp.error.message and "lat_ms" in resp.meta
        body = json.loads(resp.result)
        assert body["dim"] == 64 and len(body["vector"]) == 64 and abs(sum(v * v for v in body["vector"]) - 1.0) < 1e-4
    assert svc.model.backend._batcher.batches_run < 10
