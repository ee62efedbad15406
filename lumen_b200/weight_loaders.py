"""Vision-tower weight loading: the checkpoint formats the reference's torch backend reads
(packages/lumen-clip/src/lumen_clip/backends/torch_backend.py:228-241 open_clip `open_clip_pytorch_model.bin`,
:290-295 HuggingFace `model.safetensors` / `pytorch_model.bin`, :280-288 Chinese-CLIP) mapped onto the transformers
key names that `b200clip_load_weight` takes.  Host-side only (numpy); nothing here touches the GPU."""
from __future__ import annotations

import json
import re
from pathlib import Path

import numpy as np

from .interface import ModelLoadingError

_OC_BLOCK = re.compile(r"^visual\.transformer\.resblocks\.(\d+)\.(.+)$")
_OC_MAP = {
    "ln_1.weight": "layer_norm1.weight", "ln_1.bias": "layer_norm1.bias",
    "ln_2.weight": "layer_norm2.weight", "ln_2.bias": "layer_norm2.bias",
    "attn.out_proj.weight": "self_attn.out_proj.weight", "attn.out_proj.bias": "self_attn.out_proj.bias",
    "mlp.c_fc.weight": "mlp.fc1.weight", "mlp.c_fc.bias": "mlp.fc1.bias",
    "mlp.c_proj.weight": "mlp.fc2.weight", "mlp.c_proj.bias": "mlp.fc2.bias",
}


def _np(t) -> np.ndarray:
    if hasattr(t, "detach"):
        t = t.detach().to("cpu").float().numpy()
    return np.ascontiguousarray(t, dtype=np.float32)


def from_openclip_state_dict(sd: dict) -> dict[str, np.ndarray]:
    """open_clip `VisionTransformer` names -> transformers names.  The fused `attn.in_proj_*` is split into q/k/v,
    `visual.proj` (d x D, applied as x @ proj) becomes `visual_projection.weight` (D x d)."""
    if "model" in sd and isinstance(sd["model"], dict):
        sd = sd["model"]
    if "state_dict" in sd and isinstance(sd["state_dict"], dict):
        sd = sd["state_dict"]
    sd = {k[len("module."):] if k.startswith("module.") else k: v for k, v in sd.items()}
    if "visual.conv1.weight" not in sd:
        raise ModelLoadingError("not an open_clip ViT checkpoint (visual.conv1.weight missing; timm / ResNet towers are unsupported)")
    out = {
        "vision_model.embeddings.class_embedding": _np(sd["visual.class_embedding"]),
        "vision_model.embeddings.patch_embedding.weight": _np(sd["visual.conv1.weight"]),
        "vision_model.embeddings.position_embedding.weight": _np(sd["visual.positional_embedding"]),
        "vision_model.pre_layrnorm.weight": _np(sd["visual.ln_pre.weight"]),
        "vision_model.pre_layrnorm.bias": _np(sd["visual.ln_pre.bias"]),
        "vision_model.post_layernorm.weight": _np(sd["visual.ln_post.weight"]),
        "vision_model.post_layernorm.bias": _np(sd["visual.ln_post.bias"]),
        "visual_projection.weight": np.ascontiguousarray(_np(sd["visual.proj"]).T),
    }
    for key, val in sd.items():
        m = _OC_BLOCK.match(key)
        if not m:
            continue
        pre = f"vision_model.encoder.layers.{int(m.group(1))}."
        rest = m.group(2)
        if rest in ("attn.in_proj_weight", "attn.in_proj_bias"):
            kind = "weight" if rest.endswith("weight") else "bias"
            q, k, v = np.split(_np(val), 3, axis=0)
            for name, part in (("q_proj", q), ("k_proj", k), ("v_proj", v)):
                out[pre + f"self_attn.{name}.{kind}"] = np.ascontiguousarray(part)
        elif rest in _OC_MAP:
            out[pre + _OC_MAP[rest]] = _np(val)
    return out


def from_hf_state_dict(sd: dict) -> dict[str, np.ndarray]:
    """CLIPModel / ChineseCLIPModel / CLIPVisionModelWithProjection state dict -> the vision tower's tensors."""
    out = {}
    for k, v in sd.items():
        k = k[len("model."):] if k.startswith("model.") else k
        if k.endswith("position_ids"):
            continue
        if k.startswith("vision_model.") or k == "visual_projection.weight":
            out[k] = _np(v)
    if "vision_model.embeddings.class_embedding" not in out:
        raise ModelLoadingError("state dict holds no transformers CLIP vision tower")
    return out


def infer_vit_config(weights: dict[str, np.ndarray], activation: str | None = None) -> dict:
    """Shape of the tower from the tensors themselves."""
    d = int(weights["vision_model.embeddings.class_embedding"].shape[0])
    P = int(weights["vision_model.embeddings.patch_embedding.weight"].shape[-1])
    S = int(weights["vision_model.embeddings.position_embedding.weight"].shape[0])
    side = int(round((S - 1) ** 0.5))
    if side * side != S - 1:
        raise ModelLoadingError(f"position table with {S} rows is not 1 + a square grid")
    layers = 1 + max(int(k.split(".")[3]) for k in weights if k.startswith("vision_model.encoder.layers."))
    return dict(image_size=side * P, patch=P, width=d, layers=layers, heads=d // 64,
                mlp_dim=int(weights["vision_model.encoder.layers.0.mlp.fc1.weight"].shape[0]),
                embed_dim=int(weights["visual_projection.weight"].shape[0]), activation=activation or "quick_gelu", ln_eps=1e-5)


def load_vision_weights(model_dir: str | Path) -> tuple[dict[str, np.ndarray], dict]:
    """Finds a checkpoint under `model_dir` (or its `torch/` runtime sub-directory) and returns (weights, vit_config).
    QuickGELU vs exact GELU follows `open_clip_config.json` (`model_cfg.quick_gelu`) or `config.json`
    (`vision_config.hidden_act`); OpenAI / HF CLIP default to QuickGELU, LAION open_clip checkpoints to GELU."""
    root = Path(model_dir)
    dirs = [root, root / "torch"]
    activation = None
    for dname in dirs:
        oc = dname / "open_clip_config.json"
        if oc.exists():
            cfg = json.loads(oc.read_text())
            activation = "quick_gelu" if cfg.get("model_cfg", {}).get("quick_gelu") else "gelu"
        hc = dname / "config.json"
        if hc.exists() and activation is None:
            act = (json.loads(hc.read_text()).get("vision_config") or {}).get("hidden_act")
            if act:
                activation = "quick_gelu" if act == "quick_gelu" else "gelu"
    for dname in dirs:
        st = dname / "model.safetensors"
        if st.exists():
            from safetensors.numpy import load_file
            w = from_hf_state_dict(load_file(str(st)))
            return w, infer_vit_config(w, activation)
        for fname, conv in (("open_clip_pytorch_model.bin", from_openclip_state_dict), ("pytorch_model.bin", from_hf_state_dict)):
            f = dname / fname
            if f.exists():
                import torch
                sd = torch.load(str(f), map_location="cpu", weights_only=True)
                w = conv(sd)
                return w, infer_vit_config(w, activation if activation else ("gelu" if conv is from_openclip_state_dict else "quick_gelu"))
    raise ModelLoadingError(f"no model.safetensors / open_clip_pytorch_model.bin / pytorch_model.bin under {root}")
