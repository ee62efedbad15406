"""CPU, build container only: the oracle against the reference's OWN code run live on random inputs (the committed goldens pin a
handful of seeded cases; this walks many more).  Skipped where /root/reference does not exist (the GPU box).

  * backends/onnxrt_backend.py:378-433    the preprocess closure (PIL resize + normalise), incl. sources PIL resizes vertically first
  * general_clip/clip_model.py:232-359    CLIPModelManager.classify_image / classify_scene
  * expert_bioclip/bioclip_model.py:219-335  BioCLIPModelManager.classify_image (raw rows, no normalisation at search time)
"""
import sys
from pathlib import Path

import numpy as np
import pytest

from lumen_b200.vit_config import OPENAI_MEAN, OPENAI_STD
from oracle import clip_oracle as O

REF = Path("/root/reference/packages")
pytestmark = pytest.mark.skipif(not REF.exists(), reason="reference tree not mounted (build container only)")


@pytest.fixture(scope="module")
def ref():
    for p in (REF / "lumen-clip" / "src", REF / "lumen-resources" / "src"):
        if str(p) not in sys.path:
            sys.path.insert(0, str(p))
    try:
        from lumen_clip.backends.base import BackendInfo, BaseClipBackend
        from lumen_clip.backends.onnxrt_backend import ONNXRTBackend
        from lumen_clip.expert_bioclip.bioclip_model import BioCLIPModelManager
        from lumen_clip.general_clip.clip_model import CLIPModelManager
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference not importable here: {e}")

    class FakeBackend(BaseClipBackend):
        def __init__(self): self._initialized = True; self.vec = None; self.scene = None
        def initialize(self): pass
        def text_to_vector(self, text): raise RuntimeError
        def text_batch_to_vectors(self, texts): return self.scene
        def image_to_vector(self, b): return self.vec
        def get_info(self): return BackendInfo(runtime="fake")

    class Res:
        model_name = "fake"; runtime = "onnx"
        def __init__(self, labels, emb): self.labels = labels; self.label_embeddings = emb
        def has_classification_support(self): return True

    return dict(ONNXRTBackend=ONNXRTBackend, CLIP=CLIPModelManager, Bio=BioCLIPModelManager, FakeBackend=FakeBackend, Res=Res)


def test_preprocess_closure_live(ref):
    Image = pytest.importorskip("PIL.Image")

    class _Res:
        def __init__(self, s): self.s = s
        def get_image_size(self): return (self.s, self.s)
        def get_normalization_stats(self): return {"mean": list(OPENAI_MEAN), "std": list(OPENAI_STD)}

    rng = np.random.default_rng(21)
    tall = 0
    for n in range(60):
        kind = n % 3
        if kind == 0: h, w = int(rng.integers(1, 30)), int(rng.integers(40, 2500))
        elif kind == 1: h, w = int(rng.integers(40, 2500)), int(rng.integers(1, 30))
        else: h, w = int(rng.integers(1, 900)), int(rng.integers(1, 900))
        S = int(rng.choice([64, 224, 336]))
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        be = object.__new__(ref["ONNXRTBackend"])
        be.resources = _Res(S)
        be._vision_input_dtype = np.dtype(np.float32)
        want = be._create_image_preprocessor()(Image.fromarray(img, "RGB"))[0]
        assert np.array_equal(O.preprocess(img, S, mode=0), want), (h, w, S)
        tall += h > 100 * w and S < h
    assert tall >= 3


def test_clip_manager_live(ref):
    rng = np.random.default_rng(22)
    be = ref["FakeBackend"]()
    for case in range(12):
        N = int(rng.choice([5, 8, 100, 1000, 3001])); D = int(rng.choice([64, 512, 768]))
        text = (rng.standard_normal((N, D)) * rng.uniform(0.2, 4.0, (N, 1))).astype(np.float32)
        be.scene = rng.standard_normal((8, D)).astype(np.float32)
        labels = np.array([f"label_{i}" for i in range(N)], dtype=object)
        mgr = ref["CLIP"](be, ref["Res"](labels, text))
        mgr.initialize()
        for _ in range(4):
            v = rng.standard_normal(D).astype(np.float32)
            if rng.random() < 0.5:                                      # sit near a label so that softmax(100 s) is not flat
                v = 0.6 * v / np.linalg.norm(v) + 0.4 * text[int(rng.integers(N))] / np.linalg.norm(text[0])
            be.vec = v.astype(np.float32)
            k = int(min(N, rng.choice([1, 3, 5])))
            got = mgr.classify_image(b"x", top_k=k)
            idx, p = O.classify(be.vec, text, k)
            assert [l for l, _ in got] == [f"label_{i}" for i in idx], (N, D, k)
            np.testing.assert_allclose([q for _, q in got], p, rtol=1e-5, atol=1e-9)
            lab, conf = mgr.classify_scene(b"x")
            si, sp = O.classify_scene(be.vec, be.scene)
            cand = [s.replace("a photo of ", "").replace("an ", "") for s in mgr.scene_prompts]
            assert cand.index(lab) == si and abs(conf - sp) < 1e-6


def test_bioclip_manager_live(ref):
    rng = np.random.default_rng(23)
    be = ref["FakeBackend"]()
    for case in range(8):
        N = int(rng.choice([50, 1000, 5000])); D = int(rng.choice([64, 512]))
        g = rng.standard_normal((N, D)).astype(np.float32)
        if case % 2:
            g *= rng.uniform(0.5, 2.0, (N, 1)).astype(np.float32)       # NOT unit length: the reference multiplies the rows as given
        else:
            g /= np.linalg.norm(g, axis=1, keepdims=True)
        labels = np.array([f"sp_{i}" for i in range(N)], dtype=object)
        mgr = ref["Bio"](be, ref["Res"](labels, g))
        mgr.initialize()
        for _ in range(4):
            be.vec = rng.standard_normal(D).astype(np.float32)
            k = int(rng.choice([1, 3, 10]))
            got = mgr.classify_image(b"x", top_k=k)
            idx, sc = O.bioclip_classify_image(be.vec, g, k)
            assert [int(l.split("_")[1]) for l, _ in got] == list(idx), (N, D, k)
            np.testing.assert_allclose([s for _, s in got], sc, rtol=1e-5, atol=1e-7)


def test_extract_name_matches_the_reference(ref):
    """`B200BioCLIPModelManager.extract_name` against the reference's static method on every label shape a TreeOfLife file can hold
    (and a few it should not): host-only, so it runs wherever the reference tree is."""
    from lumen_b200.managers import B200BioCLIPModelManager as M
    R = ref["Bio"]
    cases = ["sp_3", "", ["a", "b", "c"], [["Animalia", "Chordata", "Panthera", "leo"], "Lion"], [["Animalia", "Chordata", "Panthera", "leo"], ""],
             [["Animalia", "Chordata", "Panthera", "leo"], "   "], [["leo"], ""], [[], ""], [["x", "y"], None], [["x", "y"], 7], ["tax", "Common"],
             [("a", "b"), ""], [["a", "b"], "\t\n"], [["a", "b", "c"], " Tiger "], [None, None], [[1, 2], ""]]
    for c in cases:
        assert M.extract_name(c) == R.extract_name(c), c


def test_scene_prompts_match_the_reference(ref):
    from lumen_b200.managers import B200CLIPModelManager as M
    be = ref["FakeBackend"]()
    be.scene = np.zeros((8, 4), np.float32)
    mgr = ref["CLIP"](be, ref["Res"](np.array(["a"], dtype=object), np.ones((1, 4), np.float32)))
    assert list(M.scene_prompts) == list(mgr.scene_prompts)
