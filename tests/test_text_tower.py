"""N4 — text tower.  CPU: the fp32 restatement against transformers' CLIPTextModelWithProjection (tests/golden/text.npz).
GPU: the device tower through the C ABI against the restatement (cosine >= 0.999, BASELINE.json's tolerance for embeddings)."""
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")
GOLD = Path(__file__).parent / "golden" / "text.npz"


@pytest.mark.parametrize("name", ["ViT-tiny-test", "ViT-B-32"])
def test_oracle_text_tower_matches_transformers(name):
    from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_text_weights
    from oracle import clip_oracle as O
    g = np.load(GOLD)
    cfg = TEXT_CONFIGS[name]
    e = O.text_encode(synthetic_text_weights(cfg, 0), g[f"{name}/ids"], cfg, normalize=False)
    np.testing.assert_allclose(e, g[f"{name}/text_embeds"], rtol=0, atol=2e-5)


def test_synthetic_ids_layout():
    from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_token_ids
    cfg = TEXT_CONFIGS["ViT-tiny-test"]
    ids = synthetic_token_ids(cfg, 5, seed=1)
    assert ids.shape == (5, 77) and ids.max() == cfg["vocab_size"] - 1 and (ids[:, 0] == cfg["vocab_size"] - 2).all()
    assert ids[0].argmax() == 76                                # the longest prompt puts EOT in the last slot


@pytest.mark.gpu
def test_causal_attention_vs_torch():
    from lumen_b200.engine import Engine
    eng = Engine({}, device=0)
    g = torch.Generator(device="cuda").manual_seed(9)
    for B, S, heads in ((3, 77, 2), (2, 77, 8), (2, 130, 2), (1, 257, 4), (40, 77, 8)):
        d = heads * 64
        qkv = torch.randn(B * S, 3 * d, device="cuda", generator=g).to(torch.bfloat16)
        out = eng.attention(qkv, B, S, d, heads, causal=True)
        q, k, v = [t.view(B, S, heads, 64).transpose(1, 2) for t in qkv.float().split(d, dim=1)]
        mask = torch.full((S, S), float("-inf"), device="cuda").triu(1)
        ref = (torch.softmax(q @ k.transpose(-1, -2) * 0.125 + mask, -1) @ v).transpose(1, 2).reshape(B * S, d)
        torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)
    from lumen_b200 import _lib
    with pytest.raises(_lib.B200ClipError):
        eng.attention(torch.zeros(50, 3 * 128, device="cuda", dtype=torch.bfloat16), 1, 50, 128, 2, causal=True)   # S <= 64
    eng.sync()
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,B", [("ViT-tiny-test", 9), ("ViT-B-32", 5)])
def test_text_tower_cosine_vs_oracle_and_golden(name, B):
    from lumen_b200.engine import Engine
    from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_text_weights, synthetic_token_ids
    from oracle import clip_oracle as O
    cfg = TEXT_CONFIGS[name]
    w = synthetic_text_weights(cfg, 0)
    eng = Engine({}, device=0)                                   # no image tower on this handle
    eng.init_text(cfg, max_batch=4)                              # max_batch < B exercises chunking
    eng.load_weights(w)
    ids = synthetic_token_ids(cfg, B, seed=5)
    emb = eng.encode_text(ids)
    ref = O.text_encode(w, ids, cfg)
    cos = (emb * ref).sum(1)
    assert cos.min() >= 0.999, cos
    np.testing.assert_allclose(np.linalg.norm(emb, axis=1), 1.0, atol=1e-5)
    g = np.load(GOLD)
    gold = g[f"{name}/text_embeds"]
    e2 = eng.encode_text(g[f"{name}/ids"])
    cos2 = (e2 * (gold / np.linalg.norm(gold, axis=1, keepdims=True))).sum(1)
    assert cos2.min() >= 0.999, cos2
    from lumen_b200 import _lib
    with pytest.raises(_lib.B200ClipError):
        eng.encode_text(np.full((1, 77), cfg["vocab_size"], np.int32))        # id outside the vocabulary
    with pytest.raises(_lib.B200ClipError):
        eng.encode_text(np.zeros((1, 10), np.int32))                          # wrong context length
    assert eng.encode_text(np.zeros((0, 77), np.int32)).shape == (0, cfg["embed_dim"])
    eng.close()


@pytest.mark.gpu
def test_both_towers_on_one_handle_and_label_bootstrap():
    """Image and text tower share a handle: text embeddings of label prompts become the classify label set (the reference's
    bootstrap, general_clip/clip_model.py:145-172), and classification equals the oracle's on the same embeddings."""
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.engine import Engine
    from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_text_weights, synthetic_token_ids
    from oracle import clip_oracle as O
    vc, tc = VIT_CONFIGS["ViT-tiny-test"], TEXT_CONFIGS["ViT-tiny-test"]
    eng = Engine(vc, device=0, max_batch=8)
    eng.init_text(tc, max_batch=16)
    w = dict(synthetic_weights(vc, 0)); w.update(synthetic_text_weights(tc, 0))
    eng.load_weights(w)
    labels = eng.encode_text(synthetic_token_ids(tc, 21, seed=8))
    eng.set_labels(labels, slot=0)
    rng = np.random.default_rng(2)
    imgs = [rng.integers(0, 256, (64, 64, 3), dtype=np.uint8) for _ in range(4)]
    emb = eng.embed_host(imgs, mode=0)
    vals, idx = eng.classify(torch.from_numpy(emb).cuda(), k=3, slot=0)
    eng.sync()
    for b in range(4):
        want_idx, _ = O.classify(emb[b], labels, top_k=3)
        assert [int(i) for i in idx[b].cpu()] == [int(j) for j in want_idx]
    eng.close()


@pytest.mark.gpu
def test_backend_text_methods_on_the_device():
    """BaseClipBackend.text_to_vector / text_batch_to_vectors served by the device tower when a tokenizer is supplied."""
    import zlib
    from lumen_b200 import VIT_CONFIGS, synthetic_weights
    from lumen_b200.backend import B200ClipBackend, SyntheticResources
    from lumen_b200.interface import InferenceError
    from lumen_b200.vit_config import TEXT_CONFIGS, synthetic_text_weights
    from oracle import clip_oracle as O
    name = "ViT-tiny-test"
    vc, tc = VIT_CONFIGS[name], TEXT_CONFIGS[name]
    w = dict(synthetic_weights(vc, 0)); w.update(synthetic_text_weights(tc, 0))

    def tokenizer(texts):                                         # stand-in for open_clip.tokenize: BOS, one id per word, EOT, padding
        ids = np.zeros((len(texts), tc["context_length"]), np.int64)
        for i, t in enumerate(texts):
            words = [1 + zlib.crc32(x.encode()) % (tc["vocab_size"] - 3) for x in t.split()][:tc["context_length"] - 2]
            ids[i, 0] = tc["vocab_size"] - 2
            ids[i, 1:1 + len(words)] = words
            ids[i, 1 + len(words)] = tc["vocab_size"] - 1
        return ids

    be = B200ClipBackend(SyntheticResources(model_name=name), device_preference="cuda:0", max_batch_size=8, weights=w, text_config=tc,
                         tokenizer=tokenizer)
    be.initialize()
    prompts = ["a photo of a cat", "a photo of a very small dog on a beach", "x"]
    got = be.text_batch_to_vectors(prompts)
    ref = O.text_encode(w, tokenizer(prompts), tc)
    assert got.dtype == np.float32 and got.shape == ref.shape and ((got * ref).sum(1) >= 0.999).all()
    one = be.text_to_vector(prompts[1])
    assert np.allclose(one, got[1], atol=1e-6)
    # without a tokenizer the backend keeps refusing, as before
    be2 = B200ClipBackend(SyntheticResources(model_name=name), device_preference="cuda:0", max_batch_size=8, weights=synthetic_weights(vc, 0))
    be2.initialize()
    with pytest.raises(InferenceError):
        be2.text_to_vector("a photo of a cat")
    be.close(); be2.close()
