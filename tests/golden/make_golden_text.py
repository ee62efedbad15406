"""tests/golden/text.npz: text_embeds of transformers' CLIPTextModelWithProjection (the class the reference's torch backend runs,
backends/torch_backend.py) on seeded synthetic weights and token ids.  Pins oracle.text_encode; run in this container."""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_text_weights, synthetic_token_ids  # noqa: E402
from transformers import CLIPTextConfig, CLIPTextModelWithProjection  # noqa: E402

out = {}
for name, B in (("ViT-tiny-test", 6), ("ViT-B-32", 3)):
    cfg = TEXT_CONFIGS[name]
    hf = CLIPTextConfig(vocab_size=cfg["vocab_size"], hidden_size=cfg["width"], intermediate_size=cfg["mlp_dim"], projection_dim=cfg["embed_dim"],
                        num_hidden_layers=cfg["layers"], num_attention_heads=cfg["heads"], max_position_embeddings=cfg["context_length"],
                        hidden_act="quick_gelu", layer_norm_eps=cfg["ln_eps"], eos_token_id=cfg["vocab_size"] - 1, bos_token_id=cfg["vocab_size"] - 2,
                        pad_token_id=0)
    model = CLIPTextModelWithProjection(hf).eval()
    w = synthetic_text_weights(cfg, seed=0)
    missing, unexpected = model.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}, strict=False)
    assert not unexpected and all("position_ids" in m for m in missing), (missing, unexpected)
    ids = synthetic_token_ids(cfg, B, seed=3)
    with torch.no_grad():
        emb = model(input_ids=torch.from_numpy(ids.astype(np.int64))).text_embeds.numpy()
    out[f"{name}/ids"] = ids
    out[f"{name}/text_embeds"] = emb.astype(np.float32)
np.savez_compressed(Path(__file__).with_name("text.npz"), **out)
print("text.npz ok", {k: v.shape for k, v in out.items()})
