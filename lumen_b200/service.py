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

    A caller blocks in `image_to_vector`; the first caller of a window becomes the leader, waits up to
    `max_wait_ms` for companions (or until `max_batch` requests are queued), runs ONE batched call and hands every
    caller its own row.  An exception of the batched call is re-raised in every caller of that batch."""

    def __init__(self, backend, max_batch: int = 64, max_wait_ms: float = 2.0) -> None:
        self.backend, self.max_batch, self.max_wait = backend, int(max_batch), max_wait_ms / 1e3
        self._cv = threading.Condition()
        self._queue: list[dict] = []
        self._leader_active = False
        self.batches_run = 0

    def image_to_vector(self, image_bytes: bytes) -> np.ndarray:
        slot = {"bytes": image_bytes, "out": None, "err": None, "done": False}
        with self._cv:
            self._queue.append(slot)
            self._cv.notify_all()
            lead = not self._leader_active
            if lead:
                self._leader_active = True
        if lead:
            self._lead()
        with self._cv:
            while not slot["done"]:
                self._cv.wait()
        if slot["err"] is not None:
            raise slot["err"]
        return slot["out"]

    def _lead(self) -> None:
        while True:
            deadline = time.monotonic() + self.max_wait
            with self._cv:
                while len(self._queue) < self.max_batch:
                    left = deadline - time.monotonic()
                    if left <= 0:
                        break
                    self._cv.wait(left)
                batch, self._queue = self._queue[: self.max_batch], self._queue[self.max_batch:]
            try:
                out = self.backend.image_batch_to_vectors([s["bytes"] for s in batch])
                for s, row in zip(batch, out):
                    s["out"] = np.ascontiguousarray(row, dtype=np.float32)
            except Exception as e:  # noqa: BLE001 - every waiter of this batch sees the failure
                for s in batch:
                    s["err"] = e
            with self._cv:
                self.batches_run += 1
                for s in batch:
                    s["done"] = True
                self._cv.notify_all()
                if not self._queue:
                    self._leader_active = False
                    return


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
