"""Service-level pieces (SURVEY.md §8b "B-service", §8f N3).

`MicroBatcher` coalesces the single-image requests that the reference's 10 gRPC worker threads issue concurrently
(packages/lumen-clip/src/lumen_clip/server.py:294-297; every `Infer` message carries exactly one image,
general_clip/clip_service.py:352-366) into one `image_batch_to_vectors` call, so the GPU path sees batches in
production instead of B = 1.

`b200_clip_service_class()` builds, when the reference package is importable, a `GeneralCLIPService` subclass whose
`from_config` creates a `B200ClipBackend` instead of going through `create_backend` (backends/factory.py:112-139 is
hard-wired per runtime string).  It is the class to name in `import_info.registry_class` of the YAML config."""
from __future__ import annotations

import threading
import time
from collections.abc import Sequence

import numpy as np

from .interface import BaseClipBackend


class MicroBatcher:
    """Thread-safe coalescer in front of any backend with `image_batch_to_vectors`.

    A caller blocks in `image_to_vector`.  Whoever finds no batch in flight leads ONE batch: it waits until the oldest queued
    request is `max_wait_ms` old (or `max_batch` requests are queued), runs one batched call, hands every caller of that batch its
    own row and gives the lead away; a caller whose request is still queued takes over.  So no caller serves other callers' batches
    after its own result is ready, however steady the traffic.  When a batched call fails, its requests are retried one by one, so
    only the caller whose image is at fault sees the error (the reference handles every request on its own)."""

    def __init__(self, backend, max_batch: int = 64, max_wait_ms: float = 2.0) -> None:
        self.backend, self.max_batch, self.max_wait = backend, int(max_batch), max_wait_ms / 1e3
        self._cv = threading.Condition()
        self._queue: list[dict] = []
        self._leader_active = False
        self.batches_run = 0
        self.retried_singly = 0

    def image_to_vector(self, image_bytes: bytes) -> np.ndarray:
        slot = {"bytes": image_bytes, "out": None, "err": None, "done": False, "t": time.monotonic()}
        with self._cv:
            self._queue.append(slot)
            self._cv.notify_all()
        while True:
            with self._cv:
                while not slot["done"] and self._leader_active:
                    self._cv.wait()
                if slot["done"]:
                    break
                self._leader_active = True                      # nobody is leading and this request is still queued
            try:
                self._lead_one_batch()
            finally:
                with self._cv:
                    self._leader_active = False
                    self._cv.notify_all()
        if slot["err"] is not None:
            raise slot["err"]
        return slot["out"]

    def _lead_one_batch(self) -> None:
        with self._cv:
            while len(self._queue) < self.max_batch:
                left = self._queue[0]["t"] + self.max_wait - time.monotonic() if self._queue else 0.0
                if left <= 0:
                    break
                self._cv.wait(left)
            batch, self._queue = self._queue[: self.max_batch], self._queue[self.max_batch:]
        if not batch:
            return
        try:
            out = self.backend.image_batch_to_vectors([s["bytes"] for s in batch])
            if len(out) != len(batch):
                raise RuntimeError(f"backend returned {len(out)} vectors for {len(batch)} images")
            for s, row in zip(batch, out):
                s["out"] = np.ascontiguousarray(row, dtype=np.float32)
        except Exception as e:  # noqa: BLE001
            if len(batch) == 1:
                batch[0]["err"] = e
            else:                                               # find the offender(s): everyone else still gets a result
                self.retried_singly += 1
                for s in batch:
                    try:
                        s["out"] = np.ascontiguousarray(self.backend.image_batch_to_vectors([s["bytes"]])[0], dtype=np.float32)
                    except Exception as e1:  # noqa: BLE001
                        s["err"] = e1
        with self._cv:
            self.batches_run += 1
            for s in batch:
                s["done"] = True
            self._cv.notify_all()


class BatchedBackendView(BaseClipBackend):
    """A backend whose single-image entry goes through a MicroBatcher; everything else is forwarded."""

    def __init__(self, inner, max_batch: int = 64, max_wait_ms: float = 2.0) -> None:
        self.inner = inner
        self.resources = inner.resources
        self._batcher = MicroBatcher(inner, max_batch, max_wait_ms)

    @property
    def is_initialized(self) -> bool:
        return self.inner.is_initialized

    def initialize(self) -> None:
        self.inner.initialize()

    def close(self) -> None:
        self.inner.close()

    def image_to_vector(self, image_bytes: bytes) -> np.ndarray:
        if not image_bytes:
            return self.inner.image_to_vector(image_bytes)       # raises InvalidInputError like the inner backend
        return self._batcher.image_to_vector(image_bytes)

    def image_batch_to_vectors(self, images: Sequence[bytes]) -> np.ndarray:
        return self.inner.image_batch_to_vectors(images)

    def text_to_vector(self, text: str) -> np.ndarray:
        return self.inner.text_to_vector(text)

    def text_batch_to_vectors(self, texts: Sequence[str]) -> np.ndarray:
        return self.inner.text_batch_to_vectors(texts)

    def get_info(self):
        return self.inner.get_info()

    def get_temperature(self):
        return self.inner.get_temperature()

    def __getattr__(self, name):            # engine, embed_arrays, ...
        return getattr(self.inner, name)


def b200_clip_service_class():
    """Returns `B200CLIPService` (a `GeneralCLIPService` subclass); needs the reference's lumen_clip, lumen_resources
    and grpc to be importable — it is meant to run inside a lumen-clip deployment."""
    from lumen_clip.general_clip.clip_service import GeneralCLIPService
    from lumen_clip.resources.loader import ResourceLoader

    from .backend import B200ClipBackend

    class B200CLIPService(GeneralCLIPService):
        def __init__(self, backend, resources, coalesce: bool = True) -> None:
            super().__init__(BatchedBackendView(backend) if coalesce else backend, resources)

        @classmethod
        def from_config(cls, service_config, cache_dir):
            model_config = next((service_config.models[k] for k in ("general", "clip", "general_clip") if k in service_config.models), None)
            if model_config is None:
                raise ValueError("No suitable model config found in service_config.models. "
                                 "Expected one of: 'general', 'clip', 'general_clip'")
            resources = ResourceLoader.load_model_resources(cache_dir, model_config)
            bs = service_config.backend_settings
            device = getattr(bs, "device", None) or "cuda:0"
            backend = B200ClipBackend(resources, device_preference=device if str(device).startswith("cuda") else "cuda:0",
                                      max_batch_size=max(64, int(getattr(bs, "batch_size", 1) or 1)))
            service = cls(backend, resources)
            service.initialize()                 # the hub router never calls initialize()
            return service

    return B200CLIPService
