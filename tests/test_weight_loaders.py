"""CPU: checkpoint-format converters round-trip against the synthetic tower."""
import numpy as np
import pytest

from lumen_b200.vit_config import VIT_CONFIGS, synthetic_weights
from lumen_b200 import weight_loaders as WL


def _to_openclip(w, cfg):
    sd = {"visual.class_embedding": w["vision_model.embeddings.class_embedding"],
          "visual.conv1.weight": w["vision_model.embeddings.patch_embedding.weight"],
          "visual.positional_embedding": w["vision_model.embeddings.position_embedding.weight"],
          "visual.ln_pre.weight": w["vision_model.pre_layrnorm.weight"], "visual.ln_pre.bias": w["vision_model.pre_layrnorm.bias"],
          "visual.ln_post.weight": w["vision_model.post_layernorm.weight"], "visual.ln_post.bias": w["vision_model.post_layernorm.bias"],
          "visual.proj": w["visual_projection.weight"].T.copy(), "logit_scale": np.float32(4.6), "token_embedding.weight": np.zeros((3, 3), np.float32)}
    for i in range(cfg["layers"]):
        p, o = f"vision_model.encoder.layers.{i}.", f"visual.transformer.resblocks.{i}."
        sd[o + "attn.in_proj_weight"] = np.concatenate([w[p + f"self_attn.{n}.weight"] for n in ("q_proj", "k_proj", "v_proj")])
        sd[o + "attn.in_proj_bias"] = np.concatenate([w[p + f"self_attn.{n}.bias"] for n in ("q_proj", "k_proj", "v_proj")])
        for a, b in (("attn.out_proj", "self_attn.out_proj"), ("ln_1", "layer_norm1"), ("ln_2", "layer_norm2"), ("mlp.c_fc", "mlp.fc1"), ("mlp.c_proj", "mlp.fc2")):
            sd[o + a + ".weight"] = w[p + b + ".weight"]; sd[o + a + ".bias"] = w[p + b + ".bias"]
    return sd


def test_openclip_names_roundtrip_and_config():
    cfg = VIT_CONFIGS["ViT-tiny-test"]
    w = synthetic_weights(cfg, 3)
    back = WL.from_openclip_state_dict({"state_dict": {"module." + k: v for k, v in _to_openclip(w, cfg).items()}})
    assert set(back) == set(w)
    for k in w:
        assert np.array_equal(back[k], w[k]), k
    got = WL.infer_vit_config(back, "gelu")
    for key in ("image_size", "patch", "width", "layers", "heads", "mlp_dim", "embed_dim"):
        assert got[key] == cfg[key]
    assert got["activation"] == "gelu"
    with pytest.raises(Exception):
        WL.from_openclip_state_dict({"visual.trunk.stem.weight": np.zeros(3)})


def test_checkpoint_files(tmp_path):
    torch = pytest.importorskip("torch")
    from safetensors.numpy import save_file
    cfg = VIT_CONFIGS["ViT-tiny-test"]
    w = synthetic_weights(cfg, 4)
    hf_dir = tmp_path / "hf"; hf_dir.mkdir()
    extra = {"text_model.embeddings.token_embedding.weight": np.zeros((4, 4), np.float32), "logit_scale": np.array(4.6, np.float32),
             "vision_model.embeddings.position_ids": np.arange(17, dtype=np.int64)[None]}
    save_file({**w, **extra}, str(hf_dir / "model.safetensors"))
    (hf_dir / "config.json").write_text('{"vision_config": {"hidden_act": "quick_gelu"}}')
    got, c = WL.load_vision_weights(hf_dir)
    assert set(got) == set(w) and all(np.array_equal(got[k], w[k]) for k in w) and c["activation"] == "quick_gelu" and c["layers"] == cfg["layers"]
    oc_dir = tmp_path / "oc" / "torch"; oc_dir.mkdir(parents=True)
    torch.save({k: torch.from_numpy(np.asarray(v)) for k, v in _to_openclip(w, cfg).items()}, str(oc_dir / "open_clip_pytorch_model.bin"))
    (oc_dir / "open_clip_config.json").write_text('{"model_cfg": {"embed_dim": 64, "quick_gelu": false}}')
    got, c = WL.load_vision_weights(tmp_path / "oc")
    assert all(np.array_equal(got[k], w[k]) for k in w) and c["activation"] == "gelu" and c["embed_dim"] == cfg["embed_dim"]
    with pytest.raises(Exception):
        WL.load_vision_weights(tmp_path / "nothing")
