"""lumen_b200 — B200-native (sm_100a) implementation of Lumen's lumen-clip image path:
batched preprocess -> CLIP ViT image embedding -> zero-shot classify / cosine top-k search.

Host side mirrors the reference's backend / model-manager interface; the arithmetic lives in
libb200clip.so (hand-written CUDA, C ABI in include/b200clip.h).  Importing the package does not touch the
GPU; constructing an Engine / initialising a backend without the built library or without a B200 fails."""
from .vit_config import VIT_CONFIGS, synthetic_weights  # noqa: F401

__all__ = ["VIT_CONFIGS", "synthetic_weights", "B200ClipBackend", "SyntheticResources", "Engine",
           "B200CLIPModelManager", "B200BioCLIPModelManager", "ShardedGallery"]


def __getattr__(name):
    if name in ("B200ClipBackend", "SyntheticResources"):
        from . import backend
        return getattr(backend, name)
    if name == "Engine":
        from .engine import Engine
        return Engine
    if name in ("B200CLIPModelManager", "B200BioCLIPModelManager", "ShardedGallery"):
        from . import managers
        return getattr(managers, name)
    raise AttributeError(name)
