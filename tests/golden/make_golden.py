"""Generate the committed golden fixtures from the REAL reference code (run in the build container only).

    PYTHONPATH=/root/reference/packages/lumen-clip/src:/root/reference/packages/lumen-resources/src \
        python tests/golden/make_golden.py

Sources of truth executed here (all relative to /root/reference/packages/lumen-clip/src/lumen_clip):
  * backends/onnxrt_backend.py:378-433   the reference's own preprocess closure (PIL resize + normalise)
  * PIL.Image.resize(BICUBIC) + torchvision Resize/CenterCrop (the open_clip transform of
    backends/torch_backend.py:201-204; open_clip itself is not installed)
  * general_clip/clip_model.py:232-359   CLIPModelManager.classify_image / classify_scene (fake backend)
  * expert_bioclip/bioclip_model.py:219-335  BioCLIPModelManager.classify_image (fake backend)
  * transformers CLIPVisionModelWithProjection (the class family torch_backend.py:290-295 loads)
Inputs are seeded, so only seeds + outputs are stored.  Nothing here is read at test time from
/root/reference; the .npz files are.
"""
from __future__ import annotations

import io
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
OUT = Path(__file__).resolve().parent

from lumen_b200.vit_config import OPENAI_MEAN, OPENAI_STD, VIT_CONFIGS, synthetic_weights  # noqa: E402


def synth_image(seed: int, h: int, w: int) -> np.ndarray:
    """Noise + smooth gradient + saturated blocks (exercises the clip-to-uint8 between the two passes)."""
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    grad = ((yy * 255 // max(h - 1, 1)) ^ (xx * 255 // max(w - 1, 1))).astype(np.uint8)
    img[: h // 3, :, 1] = grad[: h // 3]
    img[h // 2: h // 2 + max(h // 8, 1), : w // 2] = 255
    img[h // 2: h // 2 + max(h // 8, 1), w // 2:] = 0
    return img


PRE_CASES = [  # (seed, h_in, w_in, S)
    (1, 224, 224, 224), (2, 97, 131, 64), (3, 480, 640, 224), (4, 500, 333, 224),
    (5, 768, 1024, 64), (6, 50, 40, 64), (7, 1080, 1920, 224), (8, 336, 336, 336), (9, 700, 300, 336),
    # more than 100 times taller than wide and shrinking vertically: PIL.Image.resize runs the VERTICAL resize first (two complete
    # resizes); the last one is tall enough to take that branch under the shortest-side transform as well
    (10, 1000, 7, 224), (11, 500, 3, 64), (12, 26000, 240, 224),
]


def gen_preprocess():
    from PIL import Image
    import torchvision.transforms as T
    from lumen_clip.backends.onnxrt_backend import ONNXRTBackend

    class _Res:
        def __init__(self, s): self.s = s
        def get_image_size(self): return (self.s, self.s)
        def get_normalization_stats(self): return {"mean": list(OPENAI_MEAN), "std": list(OPENAI_STD)}

    out = {"cases": np.array(PRE_CASES, np.int32)}
    for n, (seed, h, w, S) in enumerate(PRE_CASES):
        img = synth_image(seed, h, w)
        pil = Image.fromarray(img, "RGB")
        be = object.__new__(ONNXRTBackend)
        be.resources = _Res(S)
        be._vision_input_dtype = np.dtype(np.float32)
        ref = be._create_image_preprocessor()(pil)[0]                     # the reference closure, fp32 [3,S,S]
        squash_u8 = np.asarray(pil.resize((S, S), Image.Resampling.BICUBIC))
        # the closure must be exactly normalise(PIL resize): recover u8 and check before storing
        chk = ((squash_u8.astype(np.float32) / 255.0 - np.array(OPENAI_MEAN, np.float32)) / np.array(OPENAI_STD, np.float32)).transpose(2, 0, 1)
        assert np.array_equal(chk, ref)
        crop = T.Compose([T.Resize(S, interpolation=T.InterpolationMode.BICUBIC), T.CenterCrop(S)])(pil)
        out[f"squash_u8_{n}"] = squash_u8
        out[f"crop_u8_{n}"] = np.asarray(crop)
        if S == 64:
            out[f"closure_f32_{n}"] = ref
    np.savez_compressed(OUT / "preprocess.npz", **out)
    print("preprocess.npz", sum(v.nbytes for v in out.values()))


def gen_classify():
    from lumen_clip.backends.base import BackendInfo, BaseClipBackend
    from lumen_clip.expert_bioclip.bioclip_model import BioCLIPModelManager
    from lumen_clip.general_clip.clip_model import CLIPModelManager

    class FakeBackend(BaseClipBackend):
        def __init__(self): self._initialized = True; self.vec = None
        def initialize(self): pass
        def text_to_vector(self, text): raise RuntimeError
        def text_batch_to_vectors(self, texts): return self.scene
        def image_to_vector(self, b): return self.vec
        def get_info(self): return BackendInfo(runtime="fake")

    class Res:
        model_name = "fake"; runtime = "onnx"
        def __init__(self, labels, emb): self.labels = labels; self.label_embeddings = emb
        def has_classification_support(self): return True

    D, N, NQ = 512, 1000, 8
    labels = np.array([f"label_{i}" for i in range(N)], dtype=object)
    text = np.random.default_rng(7).standard_normal((N, D)).astype(np.float32)          # NOT normalised on purpose
    scene = np.random.default_rng(8).standard_normal((8, D)).astype(np.float32)
    q = np.random.default_rng(9).standard_normal((NQ, D)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    # make query r sit close to label 10*r so the softmax(100*s) is not flat
    for r in range(NQ):
        q[r] = 0.7 * q[r] + 0.3 * text[10 * r] / np.linalg.norm(text[10 * r])
        q[r] /= np.linalg.norm(q[r])
    be = FakeBackend(); be.scene = scene
    mgr = CLIPModelManager(be, Res(labels, text)); mgr.initialize()
    cls_idx = np.zeros((NQ, 5), np.int64); cls_p = np.zeros((NQ, 5), np.float32)
    sc_idx = np.zeros(NQ, np.int64); sc_p = np.zeros(NQ, np.float32)
    for r in range(NQ):
        be.vec = q[r].copy()
        res = mgr.classify_image(b"x", top_k=5)
        cls_idx[r] = [int(l.split("_")[1]) for l, _ in res]; cls_p[r] = [p for _, p in res]
        lab, conf = mgr.classify_scene(b"x")
        cand = [p.replace("a photo of ", "").replace("an ", "") for p in mgr.scene_prompts]
        sc_idx[r] = cand.index(lab); sc_p[r] = conf
    # BioCLIP: unit-norm fp32 gallery, raw cosine top-k
    NG = 5000
    g = np.random.default_rng(10).standard_normal((NG, D)).astype(np.float32)
    g /= np.linalg.norm(g, axis=1, keepdims=True)
    glabels = np.array([f"sp_{i}" for i in range(NG)], dtype=object)
    bio = BioCLIPModelManager(be, Res(glabels, g)); bio.initialize()
    bio_idx = np.zeros((NQ, 10), np.int64); bio_s = np.zeros((NQ, 10), np.float32)
    for r in range(NQ):
        be.vec = q[r].copy()
        res = bio.classify_image(b"x", top_k=10)
        bio_idx[r] = [int(l.split("_")[1]) for l, _ in res]; bio_s[r] = [s for _, s in res]
    # BioCLIP on a gallery that is NOT unit length: the reference computes a normalised copy (bioclip_model.py:251-256) but the dot
    # product at :313 runs on the raw rows (:293), so ranking and scores follow the raw matrix.  Rows scaled by 0.5 .. 2.
    g_raw = (g * np.random.default_rng(12).uniform(0.5, 2.0, (NG, 1))).astype(np.float32)
    bio2 = BioCLIPModelManager(be, Res(glabels, g_raw)); bio2.initialize()
    braw_idx = np.zeros((NQ, 10), np.int64); braw_s = np.zeros((NQ, 10), np.float32)
    for r in range(NQ):
        be.vec = q[r].copy()
        res = bio2.classify_image(b"x", top_k=10)
        braw_idx[r] = [int(l.split("_")[1]) for l, _ in res]; braw_s[r] = [s for _, s in res]
    assert not np.array_equal(braw_idx, bio_idx)
    np.savez_compressed(OUT / "classify.npz", seeds=np.array([7, 8, 9, 10, 12]), queries=q, cls_idx=cls_idx, cls_p=cls_p,
                        scene_idx=sc_idx, scene_p=sc_p, bio_idx=bio_idx, bio_s=bio_s, bio_raw_idx=braw_idx, bio_raw_s=braw_s)
    print("classify.npz ok")


def gen_encoder():
    import torch
    from transformers import CLIPVisionConfig, CLIPVisionModelWithProjection

    out = {}
    # ViT-B/16 (BASELINE config C3) and ViT-L/14@336 (C4): two images each keep the fixture small; 24 layers at d = 1024 is where a
    # bf16 tower drifts furthest from fp32, so that is the depth the 0.999 gate has to be shown at
    for name, B in (("ViT-tiny-test", 4), ("ViT-B-32", 4), ("ViT-B-16", 2), ("ViT-L-14-336", 2)):
        cfg = VIT_CONFIGS[name]
        hf = CLIPVisionConfig(hidden_size=cfg["width"], intermediate_size=cfg["mlp_dim"], projection_dim=cfg["embed_dim"],
                              num_hidden_layers=cfg["layers"], num_attention_heads=cfg["heads"], image_size=cfg["image_size"],
                              patch_size=cfg["patch"], hidden_act="quick_gelu", layer_norm_eps=cfg["ln_eps"])
        model = CLIPVisionModelWithProjection(hf).eval()
        w = synthetic_weights(cfg, seed=0)
        sd = {k: torch.from_numpy(v) for k, v in w.items()}
        missing, unexpected = model.load_state_dict(sd, strict=False)
        assert not unexpected and all("position_ids" in m for m in missing), (missing, unexpected)
        px = np.random.default_rng(11).standard_normal((B, 3, cfg["image_size"], cfg["image_size"])).astype(np.float32)
        with torch.no_grad():
            emb = model(pixel_values=torch.from_numpy(px)).image_embeds.numpy()
        out[f"{name}/image_embeds"] = emb.astype(np.float32)
    np.savez_compressed(OUT / "encoder.npz", **out)
    print("encoder.npz ok")


if __name__ == "__main__":
    gen_preprocess()
    gen_classify()
    gen_encoder()
