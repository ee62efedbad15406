"""Model managers: the numeric half of the reference's L3 layer on the GPU.

`B200CLIPModelManager` mirrors `CLIPModelManager` (reference: packages/lumen-clip/src/lumen_clip/general_clip/
clip_model.py:48-403): `initialize`, `encode_image`, `classify_image(bytes, top_k)`, `classify_scene` with the same
argument meaning, return types and error behaviour — plus batched variants, which the reference lacks.
`B200BioCLIPModelManager` mirrors `BioCLIPModelManager.classify_image` (expert_bioclip/bioclip_model.py:219-335):
raw-cosine top-k over a large label matrix, here an fp16 device-resident gallery searched by the fused
GEMM + top-k kernel.  `ShardedGallery` row-shards a gallery over the ranks of a torch.distributed group and
merges the per-GPU candidates after one all-gather."""
from __future__ import annotations

import time
from collections.abc import Sequence

import numpy as np

from .backend import B200ClipBackend


def _torch():
    import torch
    return torch


class B200CLIPModelManager:
    scene_prompts = ["a photo of a person", "a photo of an animal", "a photo of a vehicle", "a photo of food",
                     "a photo of a building", "a photo of nature", "a photo of an object", "a photo of a landscape"]

    def __init__(self, backend: B200ClipBackend, resources, scene_embeddings: np.ndarray | None = None) -> None:
        self.backend = backend
        self.resources = resources
        self.labels = resources.labels.tolist() if resources.labels is not None else []
        self.text_embeddings = resources.label_embeddings
        self.model_id = f"{resources.model_name}_{resources.runtime}"
        self.supports_classification = resources.has_classification_support()
        self.scene_prompt_embeddings = scene_embeddings
        self._load_time = None
        self.is_initialized = False

    def initialize(self) -> None:
        if self.is_initialized:
            return
        t0 = time.time()
        try:
            self.backend.initialize()
            if self.supports_classification and self.text_embeddings is None:
                prompts = [f"a photo of a {label}" for label in self.labels]
                self.text_embeddings = self.backend.text_batch_to_vectors(prompts).astype(np.float32, copy=False)
            if self.scene_prompt_embeddings is None:
                try:
                    self.scene_prompt_embeddings = self.backend.text_batch_to_vectors(self.scene_prompts)
                except Exception:  # noqa: BLE001  (clip_model.py:195-197: scene classification becomes unavailable)
                    self.scene_prompt_embeddings = None
            eng = self.backend.engine
            if self.text_embeddings is not None:
                self._check_text(self.text_embeddings)
                eng.set_labels(self.text_embeddings, slot=0)
            if self.scene_prompt_embeddings is not None:
                eng.set_labels(self.scene_prompt_embeddings, slot=1)
            self.is_initialized = True
            self._load_time = time.time() - t0
        except Exception as e:  # noqa: BLE001
            raise RuntimeError(f"Model initialization failed: {e}") from e

    @staticmethod
    def _check_text(t: np.ndarray) -> None:
        if np.any(np.isnan(t)) or np.any(np.isinf(t)):
            raise RuntimeError("Text embeddings contain invalid values (NaN/Inf)")
        zero = int(np.sum(np.linalg.norm(t, axis=1) == 0))
        if zero:
            raise RuntimeError(f"{zero} text embeddings have zero norm - cannot normalize")

    def _ensure_initialized(self) -> None:
        if not self.is_initialized:
            raise RuntimeError("Model manager not initialized. Call initialize() first.")

    def encode_image(self, image_bytes: bytes) -> np.ndarray:
        self._ensure_initialized()
        return self.backend.image_to_vector(image_bytes)

    # -- classification on embeddings (device kernels), single and batched
    def classify_embeddings(self, emb: np.ndarray, top_k: int = 5, slot: int = 0, scale: float = 100.0, softmax: bool = True):
        self._ensure_initialized()
        torch = _torch()
        emb = np.ascontiguousarray(np.atleast_2d(emb), np.float32)
        if np.any(np.isnan(emb)) or np.any(np.isinf(emb)):
            raise RuntimeError("Image embedding contains invalid values (NaN/Inf)")
        if np.any(np.linalg.norm(emb, axis=1) == 0):
            raise RuntimeError("Image embedding has zero norm - cannot normalize")
        eng = self.backend.engine
        dev = torch.from_numpy(emb).to(f"cuda:{eng.device}")
        vals, idx = eng.classify(dev, top_k, slot=slot, scale=scale, softmax=softmax)
        return vals.cpu().numpy(), idx.cpu().numpy()

    def classify_image(self, image_bytes: bytes, top_k: int = 5) -> list[tuple[str, float]]:
        self._ensure_initialized()
        if not self.supports_classification:
            raise RuntimeError("Classification not supported: no dataset loaded. "
                               "Ensure ImageNet_1k.npz exists in the model directory.")
        if self.text_embeddings is None:
            raise RuntimeError("Text embeddings are not available")
        emb = self.encode_image(image_bytes)
        if top_k == 0:                                   # `sorted_idx[:0]` in the reference: an empty list, after the same guards
            self.classify_embeddings(emb, 1)
            return []
        vals, idx = self.classify_embeddings(emb, min(int(top_k), len(self.labels)))
        return [(self.labels[int(i)], float(v)) for v, i in zip(vals[0], idx[0]) if i >= 0]

    def classify_images(self, images: Sequence[bytes], top_k: int = 5) -> list[list[tuple[str, float]]]:
        self._ensure_initialized()
        vals, idx = self.classify_embeddings(self.backend.image_batch_to_vectors(images), top_k)
        return [[(self.labels[int(i)], float(v)) for v, i in zip(vr, ir) if i >= 0] for vr, ir in zip(vals, idx)]

    def classify_scene(self, image_bytes: bytes) -> tuple[str, float]:
        self._ensure_initialized()
        if self.scene_prompt_embeddings is None:
            raise RuntimeError("Scene embeddings are not available")
        vals, idx = self.classify_embeddings(self.encode_image(image_bytes), 1, slot=1, scale=1.0, softmax=True)
        label = self.scene_prompts[int(idx[0, 0])].replace("a photo of ", "").replace("an ", "")
        return label, float(vals[0, 0])


class ShardedGallery:
    """An fp16 embedding gallery row-sharded over the ranks of `group` (or a single GPU when torch.distributed is
    not initialised).  search(): each rank runs the fused GEMM + top-k kernel on its shard, ONE all-gather moves the candidates
    as 64-bit keys (score bits | ~row-in-shard: Q*k*8 bytes per GPU), every rank merges them with the (score desc, index asc) order."""

    def __init__(self, engine, shard_f16, index_base: int = 0, group=None) -> None:
        torch = _torch()
        assert shard_f16.dtype == torch.float16 and shard_f16.is_contiguous()
        self.engine, self.shard, self.index_base, self.group = engine, shard_f16, int(index_base), group
        self._bases = None                                  # index_base of every rank, gathered once

    @staticmethod
    def shard_bounds(n_total: int, world: int, rank: int) -> tuple[int, int]:
        per = (n_total + world - 1) // world
        lo = min(n_total, rank * per)
        return lo, min(n_total, lo + per)

    def _part_bases(self, world: int, device):
        torch = _torch()
        if self._bases is None:
            mine = torch.tensor([self.index_base], dtype=torch.int64, device=device)
            allb = torch.empty((world,), dtype=torch.int64, device=device)
            torch.distributed.all_gather_into_tensor(allb, mine, group=self.group)
            self._bases = allb
        return self._bases

    def search(self, queries_f16, k: int, events: list | None = None):
        """(scores [Q,k], global indices [Q,k]).  `events`, when given, receives four CUDA events recorded on the current stream:
        start | local top-k done | all-gather done | merge done (bench.py times the phases with them)."""
        torch = _torch()

        def mark():
            if events is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                events.append(e)

        dist = torch.distributed
        sharded = dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1
        if sharded and dist.get_backend(self.group) == "nccl" and hasattr(self.engine, "topk_keys"):
            world = dist.get_world_size(self.group)
            bases = self._part_bases(world, queries_f16.device)
            mark()
            keys = self.engine.topk_keys(queries_f16, self.shard, k)
            mark()
            all_k = torch.empty((world,) + tuple(keys.shape), dtype=keys.dtype, device=keys.device)
            dist.all_gather_into_tensor(all_k, keys, group=self.group)          # the one collective of the design, over NVLink
            mark()
            out = self.engine.merge_topk_keys(all_k, bases)
            mark()
            return out
        mark()
        scores, idx = self.engine.topk(queries_f16, self.shard, k, self.index_base)
        mark()
        if not sharded:
            mark(); mark()
            return scores, idx
        # any other backend (gloo in the CPU tests of the host logic): scores and indices travel as two tensors
        world = dist.get_world_size(self.group)
        all_s = torch.empty((world,) + tuple(scores.shape), dtype=scores.dtype, device=scores.device)
        all_i = torch.empty((world,) + tuple(idx.shape), dtype=idx.dtype, device=idx.device)
        dist.all_gather(list(all_s.unbind(0)), scores, group=self.group)
        dist.all_gather(list(all_i.unbind(0)), idx, group=self.group)
        mark()
        out = self.engine.merge_topk(all_s, all_i)
        mark()
        return out


class B200BioCLIPModelManager:
    """Raw-cosine top-k over a large label matrix (bioclip_model.py:219-335), stored fp16 on the device and searched with the
    fused kernel.

    Like the reference, the dot product runs on the label matrix AS GIVEN: `classify_image` there builds a unit-length copy
    (bioclip_model.py:251-256) but only logs it (:280-283); `similarities` is `np.dot(img, self.text_embeddings.T)` on the raw rows
    (:293, :313).  `normalize_gallery=True` is an opt-in extension (true cosine for label files that are not unit length), not the
    reference's behaviour.  Two stated deviations, both inside the 1e-3 score tolerance of the parity tests: the rows are kept as
    fp16 and the query is rounded to fp16 before the tensor-core dot product (fp32 accumulation)."""

    def __init__(self, backend: B200ClipBackend, resources, normalize_gallery: bool = False) -> None:
        self.backend = backend
        self.resources = resources
        self.labels = resources.labels.tolist() if resources.labels is not None else []
        self.text_embeddings = resources.label_embeddings
        self.model_id = f"{resources.model_name}_{resources.runtime}"
        self.supports_classification = resources.has_classification_support()
        self.normalize_gallery = bool(normalize_gallery)
        self._load_time = None
        self.is_initialized = False
        self._gallery = None

    @staticmethod
    def extract_name(label_data) -> str:
        if isinstance(label_data, list) and len(label_data) == 2:
            taxonomy, common = label_data
            if isinstance(common, str) and common.strip():
                return common
            if isinstance(taxonomy, list) and len(taxonomy) >= 2:
                return f"{taxonomy[-2]} {taxonomy[-1]}"
        return str(label_data)

    def initialize(self) -> None:
        if self.is_initialized:
            return
        t0 = time.time()
        try:
            self.backend.initialize()
            if self.supports_classification and self.text_embeddings is None and self.labels:
                prompts = [f"a photo of {name}" for name in self.labels]          # bioclip_model.py:135-143
                self.text_embeddings = self.backend.text_batch_to_vectors(prompts).astype(np.float32, copy=False)
            if self.text_embeddings is not None:
                t = np.asarray(self.text_embeddings, np.float32)
                dim = int(self.backend.engine.embed_dim)
                if t.ndim == 2 and t.shape[0] == len(self.labels) and t.shape[1] == dim:
                    pass
                elif t.ndim == 2 and t.shape[1] == len(self.labels) and t.shape[0] == dim:
                    t = t.T                                           # bioclip_model.py:294-302: wrong axis order, transposed
                else:
                    raise RuntimeError(f"Text embeddings shape incompatible: {t.shape}")      # :303-309
                from .gallery import DeviceGallery
                self._gallery = DeviceGallery(self.backend.engine, t.shape[1], capacity=t.shape[0], normalized=self.normalize_gallery)
                self._gallery.append(t, normalize=self.normalize_gallery)      # fp32 rows -> fp16 on the device
            self.is_initialized = True
            self._load_time = time.time() - t0
        except Exception as e:  # noqa: BLE001
            raise RuntimeError(f"Model initialization failed: {e}") from e

    def _ensure_initialized(self) -> None:
        if not self.is_initialized:
            raise RuntimeError("Model manager not initialized. Call initialize() first.")

    def encode_image(self, image_bytes: bytes) -> np.ndarray:
        self._ensure_initialized()
        return self.backend.image_to_vector(image_bytes)

    def search_embeddings(self, emb: np.ndarray, top_k: int):
        """Batched form of the search inside classify_image: float32 [Q, D] (or [D]) -> (scores [Q, k], indices [Q, k])."""
        self._ensure_initialized()
        if self._gallery is None:
            raise RuntimeError("Text embeddings are not available")
        torch = _torch()
        q = torch.from_numpy(np.ascontiguousarray(np.atleast_2d(emb).astype(np.float16))).to(self._gallery.view().device)
        s, i = self._gallery.search(q, top_k)
        return s.cpu().numpy(), i.cpu().numpy()

    def classify_image(self, image_bytes: bytes, top_k: int = 3) -> list[tuple[str, float]]:
        self._ensure_initialized()
        if not self.supports_classification:
            raise RuntimeError("Classification not supported: no dataset loaded. "
                               "Ensure TreeOfLife dataset exists in the model directory.")
        if self._gallery is None:
            raise RuntimeError("Text embeddings are not available")
        emb = self.encode_image(image_bytes)
        if top_k == 0:                                   # `[:0]` in the reference
            return []
        s, i = self.search_embeddings(emb, top_k)
        return [(self.extract_name(self.labels[int(j)]), float(v)) for v, j in zip(s[0], i[0]) if j >= 0]
