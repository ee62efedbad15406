"""B200ClipBackend — the drop-in for lumen-clip's backend seam.

Mirrors `BaseClipBackend` (reference: packages/lumen-clip/src/lumen_clip/backends/base.py:91-292) with the
behaviour of the reference's ONNX backend (backends/onnxrt_backend.py:465-551): encoded image bytes in,
unit-normalised float32 NumPy out, `InvalidInputError` on empty bytes, every other failure wrapped in
`InferenceError`, `RuntimeError` when used before `initialize()`.  Decode (PIL) stays on the host; resize,
normalise, the ViT tower and the L2 normalisation run in libb200clip on the GPU.  There is no CPU path."""
from __future__ import annotations

import io
import threading
from collections.abc import Sequence
from dataclasses import dataclass, field
from typing import Any

import numpy as np

from .engine import MODE_SHORTEST_CROP, MODE_SQUASH, Engine
from .interface import (BackendInfo, BaseClipBackend, InferenceError, InvalidInputError, ModelLoadingError)
from .vit_config import OPENAI_MEAN, OPENAI_STD, VIT_CONFIGS

__all__ = ["B200ClipBackend", "SyntheticResources", "install_as_reference_backend"]


@dataclass
class SyntheticResources:
    """Stand-in for `ModelResources` (resources/loader.py:36-139) when no model directory exists: the same
    accessors the backend and the model managers use."""
    model_name: str = "ViT-B-32"
    runtime: str = "b200"
    config: dict[str, Any] = field(default_factory=dict)
    labels: np.ndarray | None = None
    label_embeddings: np.ndarray | None = None
    weights: dict[str, np.ndarray] | None = None
    model_root_path: Any = None          # a model directory with a checkpoint (weight_loaders.load_vision_weights), as ModelResources has

    def get_image_size(self):
        s = self.config.get("image_size") or (self.config.get("b200_vit") or VIT_CONFIGS.get(self.model_name) or {}).get("image_size")
        if s is None:
            return None                  # ModelResources.get_image_size (resources/loader.py:80-99) also returns None when the config is silent
        return (s, s) if isinstance(s, int) else (int(s[0]), int(s[1]))

    def get_normalization_stats(self):
        pc = self.config.get("preprocess_cfg") or {}
        return {"mean": list(pc.get("mean", OPENAI_MEAN)), "std": list(pc.get("std", OPENAI_STD))}

    def get_embedding_dim(self) -> int:
        return VIT_CONFIGS[self.model_name]["embed_dim"]

    def has_classification_support(self) -> bool:
        return self.labels is not None


def _vit_config_for(resources) -> dict:
    cfg = getattr(resources, "config", {}) or {}
    name = getattr(resources, "model_name", "")
    if "b200_vit" in cfg:
        return dict(cfg["b200_vit"])
    for key in (name, name.replace("_", "-")):
        if key in VIT_CONFIGS:
            return dict(VIT_CONFIGS[key])
    vc = cfg.get("vision_cfg")
    if isinstance(vc, dict) and {"width", "layers", "patch_size"} <= set(vc):
        d = int(vc["width"])
        return dict(image_size=int(vc.get("image_size", 224)), patch=int(vc["patch_size"]), width=d, layers=int(vc["layers"]),
                    heads=d // 64, mlp_dim=int(d * vc.get("mlp_ratio", 4.0)), embed_dim=int(cfg.get("embed_dim", 512)),
                    activation="quick_gelu" if cfg.get("quick_gelu") else "gelu", ln_eps=1e-5)
    raise ModelLoadingError(f"no ViT shape known for model '{name}'")


class B200ClipBackend(BaseClipBackend):
    def __init__(self, resources, device_preference: str | None = None, max_batch_size: int | None = None,
                 weights: dict[str, np.ndarray] | None = None, preprocess_mode: str = "squash",
                 text_backend: BaseClipBackend | None = None, text_config: dict | None = None, tokenizer=None) -> None:
        """text_config (a TEXT_CONFIGS entry) + tokenizer (callable: list[str] -> int ids [B, context_length], e.g. open_clip's
        tokenizer or a transformers CLIPTokenizer wrapper) put the text tower on the device too; its tensors are the text_model.* /
        text_projection.weight entries of `weights`."""
        super().__init__(resources, device_preference, max_batch_size)
        self._text_cfg, self._tokenizer = text_config, tokenizer
        self._weights = weights if weights is not None else getattr(resources, "weights", None)
        self._mode = MODE_SQUASH if preprocess_mode == "squash" else MODE_SHORTEST_CROP
        self._text = text_backend
        self._engine: Engine | None = None
        self._lock = threading.Lock()          # the reference server calls from 10 threads (server.py:294-297)
        self._device_index = 0
        if device_preference and ":" in device_preference:
            self._device_index = int(device_preference.split(":")[1])

    # ---------------------------------------------------------------- lifecycle
    def initialize(self) -> None:
        if self._initialized:
            return
        try:
            if self._weights is None:
                # real checkpoints: the formats the reference's torch backend reads (weight_loaders.py)
                from .weight_loaders import load_vision_weights
                root = getattr(self.resources, "runtime_files_path", None) or getattr(self.resources, "model_root_path", None)
                if root is not None:
                    self._weights, found = load_vision_weights(root)
                    cfgd = dict(getattr(self.resources, "config", {}) or {})
                    cfgd["b200_vit"] = found
                    try:
                        self.resources.config = cfgd
                    except Exception:  # noqa: BLE001 - frozen resources: fall back to name lookup
                        pass
            cfg = _vit_config_for(self.resources)
            size = self.resources.get_image_size()
            if size is not None:
                if size[0] != size[1]:
                    raise ModelLoadingError("non-square model input sizes are not supported")
                cfg["image_size"] = int(size[0])
            stats = self.resources.get_normalization_stats()
            if self._weights is None:
                raise ModelLoadingError("no vision-tower weights supplied (pass weights= or resources.weights)")
            self._engine = Engine(cfg, device=self._device_index, max_batch=self._max_batch_size or 512,
                                  mean=stats["mean"], std=stats["std"])
            has_text = self._text_cfg is not None and any(k.startswith("text_") for k in self._weights)
            if has_text:
                self._engine.init_text(self._text_cfg, max_batch=min(self._max_batch_size or 256, 256))
            self._engine.load_weights({k: v for k, v in self._weights.items() if has_text or not k.startswith("text_")})
            self._cfg = cfg
            self._initialized = True
        except ModelLoadingError:
            raise
        except Exception as e:  # noqa: BLE001
            raise ModelLoadingError(f"B200 model loading failed: {e}") from e

    def close(self) -> None:
        if self._engine is not None:
            self._engine.close()
            self._engine = None
        self._initialized = False

    def _ensure_initialized(self) -> None:
        if not self._initialized or self._engine is None:
            raise RuntimeError("B200ClipBackend is not initialized. Call initialize() first.")

    # ---------------------------------------------------------------- images
    @staticmethod
    def _decode(image_bytes: bytes) -> np.ndarray:
        from PIL import Image
        return np.asarray(Image.open(io.BytesIO(image_bytes)).convert("RGB"))

    def embed_arrays(self, images: Sequence[np.ndarray]) -> np.ndarray:
        """Batched entry on decoded uint8 HWC images (the reference's metric excludes decode)."""
        self._ensure_initialized()
        try:
            with self._lock:
                return self._engine.embed_host(images, self._mode)
        except Exception as e:  # noqa: BLE001
            raise InferenceError(f"Batch image encoding failed: {e}") from e

    def embed_packed(self, pinned_u8, total_bytes: int, offsets: np.ndarray, hw: np.ndarray) -> np.ndarray:
        """Same, for a batch the caller already packed into pinned host memory (engine.pack_images layout)."""
        self._ensure_initialized()
        try:
            with self._lock:
                return self._engine.embed_packed_host(pinned_u8, total_bytes, offsets, hw, self._mode)
        except Exception as e:  # noqa: BLE001
            raise InferenceError(f"Batch image encoding failed: {e}") from e

    def embed_stream(self, batches):
        """Pipelined embed_packed over an iterable of (pinned_u8, total_bytes, offsets, hw): the upload of the next batch
        overlaps the encoding of the current one; yields one float32 [B, D] array per batch, in order."""
        self._ensure_initialized()
        # the lock is taken per batch, never across a yield: other threads' requests interleave between batches, and a consumer
        # that abandons the generator leaves nothing locked
        gen = self._engine.embed_stream(batches, self._mode)
        try:
            while True:
                with self._lock:
                    try:
                        out = next(gen)
                    except StopIteration:
                        return
                yield out
        finally:
            gen.close()

    def image_to_vector(self, image_bytes: bytes) -> np.ndarray:
        self._ensure_initialized()
        if not image_bytes:
            raise InvalidInputError("image_bytes cannot be empty")
        try:
            with self._lock:
                return self._engine.embed_bytes([image_bytes], self._mode, host_decode=self._decode)[0]
        except Exception as e:  # noqa: BLE001
            raise InferenceError(f"Image encoding failed: {e}") from e

    def image_batch_to_vectors(self, images: Sequence[bytes]) -> np.ndarray:
        if not images:
            return np.empty((0, 0), dtype=np.float32)       # onnxrt_backend.py:499-500
        self._ensure_initialized()
        try:
            for b in images:
                if not b:
                    raise InvalidInputError("image bytes cannot be empty")
            # baseline JPEGs are decoded on the device (bit-exact with the reference's PIL call); other formats go through PIL
            with self._lock:
                return self._engine.embed_bytes(list(images), self._mode, host_decode=self._decode)
        except Exception as e:  # noqa: BLE001
            raise InferenceError(f"Batch image encoding failed: {e}") from e

    # ---------------------------------------------------------------- text (out of scope: delegated or refused)
    def _device_text(self, texts: Sequence[str]) -> np.ndarray:
        try:
            ids = np.asarray(self._tokenizer(list(texts)))
            with self._lock:
                return self._engine.encode_text(ids)
        except Exception as e:  # noqa: BLE001
            raise InferenceError(f"Text encoding failed: {e}") from e

    def _has_device_text(self) -> bool:
        return self._engine is not None and self._engine.text_cfg is not None and self._tokenizer is not None

    def text_to_vector(self, text: str) -> np.ndarray:
        if self._has_device_text():
            self._ensure_initialized()
            if not text or not text.strip():
                raise InvalidInputError("text cannot be empty")
            return self._device_text([text])[0]
        if self._text is not None:
            return self._text.text_to_vector(text)
        raise InferenceError("Text encoding failed: the B200 backend accelerates the image path only; "
                             "supply precomputed label embeddings or a text_backend")

    def text_batch_to_vectors(self, texts: Sequence[str]) -> np.ndarray:
        if self._has_device_text():
            self._ensure_initialized()
            if not texts:
                return np.empty((0, 0), dtype=np.float32)
            return self._device_text(texts)
        if self._text is not None:
            return self._text.text_batch_to_vectors(texts)
        raise InferenceError("Batch text encoding failed: the B200 backend accelerates the image path only")

    # ---------------------------------------------------------------- metadata
    def get_info(self) -> BackendInfo:
        size = self.resources.get_image_size()
        dim = _vit_config_for(self.resources)["embed_dim"]
        return BackendInfo(runtime="b200", device=f"cuda:{self._device_index}", model_id=self.resources.model_name,
                           model_name=getattr(self.resources, "config", {}).get("model_name", self.resources.model_name),
                           pretrained=None, version="libb200clip-0.1", image_embedding_dim=dim, text_embedding_dim=dim,
                           precisions=["bf16"], max_batch_size=self._max_batch_size, supports_image_batch=True,
                           extra={"image_size": f"{size[0]}x{size[1]}" if size else None,
                                  "preprocess": "squash" if self._mode == MODE_SQUASH else "shortest+center_crop"})

    @property
    def engine(self) -> Engine:
        self._ensure_initialized()
        return self._engine


def install_as_reference_backend(runtime: str = "torch", **backend_kwargs) -> type:
    """Makes the reference's own `create_backend(backend_config, resources, runtime=...)` (backends/factory.py:75-141) build a
    `B200ClipBackend`.  The factory hard-codes one class per runtime string and imports it from a fixed module path at call time
    (:112-139), and the `Runtime` enum is closed (lumen_resources/lumen_config.py:181-188), so the substitution happens at the
    module level - the technique of the reference's own test (packages/lumen-clip/tests/test_backend_factory.py:37-45): a module
    object exposing a class of the expected name, with the constructor keywords the factory passes, is placed in `sys.modules`.
    Call it once at start-up, before `create_backend`; YAML configs that say `runtime: torch` (or `onnx`) then get this backend with
    no change to any caller.  `backend_kwargs` (e.g. weights=, preprocess_mode=) are forwarded to every instance.  Returns the class."""
    import sys
    import types
    names = {"torch": ("lumen_clip.backends.torch_backend", "TorchBackend"), "onnx": ("lumen_clip.backends.onnxrt_backend", "ONNXRTBackend")}
    if runtime not in names:
        raise ValueError(f"runtime must be one of {sorted(names)}")
    mod_name, cls_name = names[runtime]

    class _B200AsReference(B200ClipBackend):
        def __init__(self, resources, providers=None, device_preference=None, max_batch_size=None, precision=None):
            dev = device_preference if device_preference and str(device_preference).startswith("cuda") else "cuda:0"
            super().__init__(resources, device_preference=dev, max_batch_size=max_batch_size, **backend_kwargs)
            self.requested_precision = precision            # the tower computes in bf16 whatever ONNX file precision was asked for

    _B200AsReference.__name__ = _B200AsReference.__qualname__ = cls_name
    mod = types.ModuleType(mod_name)
    setattr(mod, cls_name, _B200AsReference)
    sys.modules[mod_name] = mod
    try:                                                     # a registry filled before the substitution would keep the old class
        from lumen_clip.backends import factory
        factory._BACKEND_REGISTRY.pop(runtime, None)
    except Exception:  # noqa: BLE001 - reference package not importable: nothing to refresh
        pass
    return _B200AsReference
