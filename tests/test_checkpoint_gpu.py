"""SURVEY §8f N1 end to end: a checkpoint FILE in each format the reference's torch backend reads
(packages/lumen-clip/src/lumen_clip/backends/torch_backend.py:228-241 open_clip `open_clip_pytorch_model.bin`, :290-295 transformers
`model.safetensors`) -> `B200ClipBackend(resources)` with no `weights=` -> `initialize()` -> embeddings, against the module that
wrote the file, run in fp32 on the CPU.

The files come from real module definitions, not from an inverse of our own name mapping:
  * transformers' `CLIPModel.save_pretrained` (the class `torch_backend.py:290-295` loads), full model with text tower and logit scale;
  * an open_clip-layout vision tower assembled from torch's own `nn.MultiheadAttention` (fused `in_proj_weight` / `in_proj_bias`,
    `out_proj`), which is the layer open_clip's `ResidualAttentionBlock` uses; `visual.proj` is a d x D matrix applied as x @ proj.
    open_clip itself is not installed in this image, so its class is restated here (names from open_clip's state-dict keys)."""
import json

import numpy as np
import pytest

from conftest import synth_image

torch = pytest.importorskip("torch")
nn = torch.nn


class _OCBlock(nn.Module):
    def __init__(self, d, heads, mlp, quick):
        super().__init__()
        self.ln_1 = nn.LayerNorm(d); self.attn = nn.MultiheadAttention(d, heads, batch_first=True); self.ln_2 = nn.LayerNorm(d)
        self.mlp = nn.Sequential()
        self.mlp.add_module("c_fc", nn.Linear(d, mlp)); self.mlp.add_module("gelu", nn.Identity()); self.mlp.add_module("c_proj", nn.Linear(mlp, d))
        self.quick = quick

    def forward(self, x):
        y = self.ln_1(x)
        x = x + self.attn(y, y, y, need_weights=False)[0]
        h = self.mlp.c_fc(self.ln_2(x))
        h = h * torch.sigmoid(1.702 * h) if self.quick else nn.functional.gelu(h)
        return x + self.mlp.c_proj(h)


class _OCVisual(nn.Module):
    """open_clip.transformer.VisionTransformer, reduced to what its state dict and forward pass define."""
    def __init__(self, size, patch, d, layers, heads, mlp, D, quick):
        super().__init__()
        n = (size // patch) ** 2
        self.conv1 = nn.Conv2d(3, d, patch, patch, bias=False)
        self.class_embedding = nn.Parameter(torch.randn(d) * 0.5)
        self.positional_embedding = nn.Parameter(torch.randn(n + 1, d) * 0.5)
        self.ln_pre = nn.LayerNorm(d)
        self.transformer = nn.Module()
        self.transformer.resblocks = nn.ModuleList([_OCBlock(d, heads, mlp, quick) for _ in range(layers)])
        self.ln_post = nn.LayerNorm(d)
        self.proj = nn.Parameter(torch.randn(d, D) / d ** 0.5)

    def forward(self, px):
        x = self.conv1(px).flatten(2).transpose(1, 2)
        x = torch.cat([self.class_embedding.expand(x.shape[0], 1, -1), x], 1) + self.positional_embedding
        x = self.ln_pre(x)
        for b in self.transformer.resblocks:
            x = b(x)
        return self.ln_post(x[:, 0]) @ self.proj


def _perturb(module, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for n, p in module.named_parameters():
            if p.ndim == 1:                                   # LayerNorm gains / offsets / biases: move them off their defaults
                p.add_(0.1 * torch.randn(p.shape, generator=g))


def _write_hf(dirpath):
    from transformers import CLIPConfig, CLIPModel
    torch.manual_seed(0)
    cfg = CLIPConfig(projection_dim=64,
                     vision_config=dict(hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=2, image_size=64,
                                        patch_size=16, hidden_act="quick_gelu", projection_dim=64),
                     text_config=dict(hidden_size=64, intermediate_size=128, num_hidden_layers=1, num_attention_heads=2, vocab_size=100,
                                      max_position_embeddings=16, projection_dim=64))
    model = CLIPModel(cfg).eval()
    _perturb(model, 1)
    model.save_pretrained(str(dirpath), safe_serialization=True)
    return model


def _write_openclip(dirpath, quick):
    torch.manual_seed(1)
    vis = _OCVisual(64, 16, 128, 2, 2, 256, 64, quick).eval()
    _perturb(vis, 2)
    sd = {"visual." + k: v for k, v in vis.state_dict().items()}
    sd.update({"logit_scale": torch.tensor(4.6), "positional_embedding": torch.zeros(16, 64), "token_embedding.weight": torch.zeros(100, 64)})
    (dirpath / "torch").mkdir(parents=True)
    torch.save(sd, str(dirpath / "torch" / "open_clip_pytorch_model.bin"))
    (dirpath / "torch" / "open_clip_config.json").write_text(json.dumps({"model_cfg": {"embed_dim": 64, "quick_gelu": quick,
                                                                                        "vision_cfg": {"image_size": 64, "layers": 2, "width": 128, "patch_size": 16}}}))
    return vis


def test_checkpoint_files_load_on_the_host(tmp_path):
    """CPU part: both real files parse, the tower shape and the activation are inferred from them."""
    from lumen_b200 import weight_loaders as WL
    _write_hf(tmp_path / "hf")
    w, c = WL.load_vision_weights(tmp_path / "hf")
    assert (c["image_size"], c["patch"], c["width"], c["layers"], c["heads"], c["mlp_dim"], c["embed_dim"], c["activation"]) == (64, 16, 128, 2, 2, 256, 64, "quick_gelu")
    assert not any(k.startswith("text_") or k == "logit_scale" for k in w)
    for quick in (True, False):
        d = tmp_path / f"oc{int(quick)}"
        _write_openclip(d, quick)
        w, c = WL.load_vision_weights(d)
        assert c["activation"] == ("quick_gelu" if quick else "gelu") and c["layers"] == 2 and c["embed_dim"] == 64
        assert w["visual_projection.weight"].shape == (64, 128) and w["vision_model.encoder.layers.1.self_attn.k_proj.weight"].shape == (128, 128)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["hf_safetensors", "openclip_quickgelu", "openclip_gelu"])
def test_checkpoint_file_to_embeddings(tmp_path, fmt):
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    from oracle import clip_oracle as O
    if fmt == "hf_safetensors":
        model = _write_hf(tmp_path)
        forward = lambda px: model.get_image_features(pixel_values=px)            # noqa: E731
    else:
        vis = _write_openclip(tmp_path, quick=fmt.endswith("quickgelu"))
        forward = vis
    res = SyntheticResources(model_name="from-file", model_root_path=tmp_path)    # no weights=, no shape: everything comes from the file
    be = B200ClipBackend(res, device_preference="cuda:0", max_batch_size=4)
    be.initialize()
    imgs = [synth_image(200 + i, 70 + 11 * i, 95 - 6 * i) for i in range(6)]
    emb = be.embed_arrays(imgs)
    px = torch.from_numpy(O.preprocess_batch(imgs, 64, 0))
    with torch.no_grad():
        ref = forward(px)
        ref = getattr(ref, "pooler_output", ref)
        ref = torch.nn.functional.normalize(ref.float(), dim=1).numpy()
    assert emb.shape == ref.shape == (6, 64)
    cos = (emb * ref).sum(1)
    assert cos.min() >= 0.999, cos
    assert be.get_info().image_embedding_dim == 64
    be.close()
    # a directory without a checkpoint: the reference's ModelLoadingError contract
    from lumen_b200.interface import ModelLoadingError
    empty = tmp_path / "empty"; empty.mkdir()
    with pytest.raises(ModelLoadingError):
        B200ClipBackend(SyntheticResources(model_name="from-file", model_root_path=empty), device_preference="cuda:0").initialize()
