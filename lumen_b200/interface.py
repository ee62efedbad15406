"""The reference's backend interface for this path.

When the reference package is importable (drop-in deployment inside lumen-clip) its own classes are used, so
`B200ClipBackend` IS a `lumen_clip.backends.base.BaseClipBackend`.  Otherwise (bench box, tests) an
equivalent interface with the same names, arguments and error behaviour is defined here
(reference: packages/lumen-clip/src/lumen_clip/backends/base.py:47-292, backend_exceptions.py:7-62)."""
from __future__ import annotations

import abc
from collections.abc import Sequence
from dataclasses import dataclass, field

import numpy as np

try:  # pragma: no cover - exercised only inside a lumen-clip deployment
    from lumen_clip.backends.backend_exceptions import (BackendError, InferenceError, InvalidInputError,
                                                        ModelLoadingError)
    from lumen_clip.backends.base import BackendInfo, BaseClipBackend
    USING_REFERENCE_INTERFACE = True
except Exception:  # noqa: BLE001 - any import problem means "not deployed inside lumen-clip"
    USING_REFERENCE_INTERFACE = False

    class BackendError(Exception):
        """Base class for all backend errors (backend_exceptions.py:7)."""

    class InvalidInputError(BackendError):
        """Input data is invalid or malformed (backend_exceptions.py:19)."""

    class InferenceError(BackendError):
        """An inference operation failed (backend_exceptions.py:25)."""

    class ModelLoadingError(BackendError):
        """Model loading failed (backend_exceptions.py:30)."""

    @dataclass
    class BackendInfo:
        """base.py:47-88."""
        runtime: str
        device: str | None = None
        model_id: str | None = None
        model_name: str | None = None
        pretrained: str | None = None
        version: str | None = None
        image_embedding_dim: int | None = None
        text_embedding_dim: int | None = None
        precisions: list[str] = field(default_factory=list)
        max_batch_size: int | None = None
        supports_image_batch: bool = False
        extra: dict[str, str | None] = field(default_factory=dict)

        def as_dict(self) -> dict[str, object]:
            return {k: (list(v) if isinstance(v, list) else dict(v) if isinstance(v, dict) else v)
                    for k, v in self.__dict__.items()}

    class BaseClipBackend(abc.ABC):
        """base.py:91-292: lifecycle + unit-normalised float32 encoders + metadata."""

        def __init__(self, resources, device_preference: str | None = None, max_batch_size: int | None = None) -> None:
            self._initialized = False
            self.resources = resources
            self._device_pref = device_preference
            self._max_batch_size = max_batch_size

        @abc.abstractmethod
        def initialize(self) -> None: ...

        def close(self) -> None:
            return

        @property
        def is_initialized(self) -> bool:
            return self._initialized

        @abc.abstractmethod
        def text_to_vector(self, text: str) -> np.ndarray: ...

        @abc.abstractmethod
        def image_to_vector(self, image_bytes: bytes) -> np.ndarray: ...

        def image_batch_to_vectors(self, images: Sequence[bytes]) -> np.ndarray:
            if not images:
                return np.empty((0, 0), dtype=np.float32)
            vecs = []
            for img in images:
                try:
                    vecs.append(self.image_to_vector(img))
                except (BackendError, RuntimeError):
                    continue
            if not vecs:
                raise InferenceError("Batch image encoding failed: All images in batch failed to encode")
            return np.stack(vecs, axis=0).astype(np.float32, copy=False)

        def text_batch_to_vectors(self, texts: Sequence[str]) -> np.ndarray:
            if not texts:
                raise InferenceError("All texts in batch failed to encode, empty input.")
            return np.stack([self.text_to_vector(t) for t in texts], axis=0).astype(np.float32, copy=False)

        @abc.abstractmethod
        def get_info(self) -> BackendInfo: ...

        def get_temperature(self) -> float | None:
            return None

        @staticmethod
        def unit_normalize(vec: np.ndarray, axis: int = -1, eps: float = 1e-8) -> np.ndarray:
            v = vec.astype(np.float32, copy=False)
            return v / np.maximum(np.linalg.norm(v, ord=2, axis=axis, keepdims=True), eps)
