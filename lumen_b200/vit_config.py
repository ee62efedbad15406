"""Shapes of the CLIP image towers the reference ships (packages/lumen-clip/README.md:119-125) and
synthetic weights with the transformers key names the reference's torch backend loads
(backends/torch_backend.py:290-295).  Host-side only; no device code here."""
from __future__ import annotations

import numpy as np

VIT_CONFIGS: dict[str, dict] = {
    "ViT-B-32": dict(image_size=224, patch=32, width=768, layers=12, heads=12, mlp_dim=3072, embed_dim=512),
    "ViT-B-16": dict(image_size=224, patch=16, width=768, layers=12, heads=12, mlp_dim=3072, embed_dim=512),
    "ViT-L-14": dict(image_size=224, patch=14, width=1024, layers=24, heads=16, mlp_dim=4096, embed_dim=768),       # CN-CLIP_ViT-L-14, bioclip-2
    "ViT-L-14-336": dict(image_size=336, patch=14, width=1024, layers=24, heads=16, mlp_dim=4096, embed_dim=768),
    # small shapes for fast parity tests (same code path, every dimension a legal tile multiple)
    "ViT-tiny-test": dict(image_size=64, patch=16, width=128, layers=2, heads=2, mlp_dim=256, embed_dim=64),
}
for _c in VIT_CONFIGS.values():
    _c.setdefault("activation", "quick_gelu")
    _c.setdefault("ln_eps", 1e-5)

OPENAI_MEAN = (0.48145466, 0.4578275, 0.40821073)
OPENAI_STD = (0.26862954, 0.26130258, 0.27577711)


def weight_shapes(cfg: dict) -> dict[str, tuple[int, ...]]:
    d, L, P, mlp, D = cfg["width"], cfg["layers"], cfg["patch"], cfg["mlp_dim"], cfg["embed_dim"]
    n = (cfg["image_size"] // P) ** 2
    s: dict[str, tuple[int, ...]] = {
        "vision_model.embeddings.class_embedding": (d,),
        "vision_model.embeddings.patch_embedding.weight": (d, 3, P, P),
        "vision_model.embeddings.position_embedding.weight": (n + 1, d),
        "vision_model.pre_layrnorm.weight": (d,),
        "vision_model.pre_layrnorm.bias": (d,),
        "vision_model.post_layernorm.weight": (d,),
        "vision_model.post_layernorm.bias": (d,),
        "visual_projection.weight": (D, d),
    }
    for i in range(L):
        p = f"vision_model.encoder.layers.{i}."
        for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
            s[p + f"self_attn.{nm}.weight"] = (d, d)
            s[p + f"self_attn.{nm}.bias"] = (d,)
        for nm in ("layer_norm1", "layer_norm2"):
            s[p + nm + ".weight"] = (d,)
            s[p + nm + ".bias"] = (d,)
        s[p + "mlp.fc1.weight"] = (mlp, d)
        s[p + "mlp.fc1.bias"] = (mlp,)
        s[p + "mlp.fc2.weight"] = (d, mlp)
        s[p + "mlp.fc2.bias"] = (d,)
    return s


def synthetic_weights(cfg: dict, seed: int = 0) -> dict[str, np.ndarray]:
    """Seeded random weights of the named architecture (no checkpoints exist offline).  Linear and conv
    weights ~ N(0, 1/fan_in) so activations keep unit scale; biases, LayerNorm gains and offsets are
    perturbed so every epilogue term is exercised."""
    rng = np.random.default_rng(seed)
    out: dict[str, np.ndarray] = {}
    for name, shape in weight_shapes(cfg).items():
        if "norm" in name and name.endswith(".weight"):
            w = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith(".bias"):
            w = 0.1 * rng.standard_normal(shape)
        elif name.endswith("class_embedding") or "position_embedding" in name:
            w = 0.5 * rng.standard_normal(shape)
        else:
            fan_in = int(np.prod(shape[1:]))
            w = rng.standard_normal(shape) / np.sqrt(fan_in)
        out[name] = w.astype(np.float32)
    return out


# ------------------------------------------------------------------------------------------------ text towers (N4)
TEXT_CONFIGS: dict[str, dict] = {
    "ViT-B-32": dict(context_length=77, vocab_size=49408, width=512, layers=12, heads=8, mlp_dim=2048, embed_dim=512),
    "ViT-B-16": dict(context_length=77, vocab_size=49408, width=512, layers=12, heads=8, mlp_dim=2048, embed_dim=512),
    "ViT-L-14": dict(context_length=77, vocab_size=49408, width=768, layers=12, heads=12, mlp_dim=3072, embed_dim=768),
    "ViT-L-14-336": dict(context_length=77, vocab_size=49408, width=768, layers=12, heads=12, mlp_dim=3072, embed_dim=768),
    "ViT-tiny-test": dict(context_length=77, vocab_size=1000, width=128, layers=2, heads=2, mlp_dim=256, embed_dim=64),
}
for _c in TEXT_CONFIGS.values():
    _c.setdefault("activation", "quick_gelu")
    _c.setdefault("ln_eps", 1e-5)


def text_weight_shapes(cfg: dict) -> dict[str, tuple[int, ...]]:
    d, L, mlp, D = cfg["width"], cfg["layers"], cfg["mlp_dim"], cfg["embed_dim"]
    s: dict[str, tuple[int, ...]] = {
        "text_model.embeddings.token_embedding.weight": (cfg["vocab_size"], d),
        "text_model.embeddings.position_embedding.weight": (cfg["context_length"], d),
        "text_model.final_layer_norm.weight": (d,),
        "text_model.final_layer_norm.bias": (d,),
        "text_projection.weight": (D, d),
    }
    for i in range(L):
        p = f"text_model.encoder.layers.{i}."
        for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
            s[p + f"self_attn.{nm}.weight"] = (d, d)
            s[p + f"self_attn.{nm}.bias"] = (d,)
        for nm in ("layer_norm1", "layer_norm2"):
            s[p + nm + ".weight"] = (d,)
            s[p + nm + ".bias"] = (d,)
        s[p + "mlp.fc1.weight"] = (mlp, d)
        s[p + "mlp.fc1.bias"] = (mlp,)
        s[p + "mlp.fc2.weight"] = (d, mlp)
        s[p + "mlp.fc2.bias"] = (d,)
    return s


def synthetic_text_weights(cfg: dict, seed: int = 0) -> dict[str, np.ndarray]:
    """Seeded random text-tower weights, same recipe as synthetic_weights."""
    rng = np.random.default_rng(seed + 1000)
    out: dict[str, np.ndarray] = {}
    for name, shape in text_weight_shapes(cfg).items():
        if "norm" in name and name.endswith(".weight"):
            w = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith(".bias"):
            w = 0.1 * rng.standard_normal(shape)
        elif "embedding" in name:
            w = 0.5 * rng.standard_normal(shape)
        else:
            w = rng.standard_normal(shape) / np.sqrt(int(np.prod(shape[1:])))
        out[name] = w.astype(np.float32)
    return out


def synthetic_token_ids(cfg: dict, B: int, seed: int = 0) -> np.ndarray:
    """[B, context_length] int32 in CLIP's layout: BOS, a few tokens, EOT (= the largest id), zero padding."""
    rng = np.random.default_rng(seed)
    V, ctx = cfg["vocab_size"], cfg["context_length"]
    ids = np.zeros((B, ctx), np.int32)
    for b in range(B):
        n = int(rng.integers(1, ctx - 2)) if b else ctx - 2          # row 0: the longest legal prompt (EOT in the last slot)
        ids[b, 0] = V - 2
        ids[b, 1:1 + n] = rng.integers(1, V - 2, n)
        ids[b, 1 + n] = V - 1
    return ids
