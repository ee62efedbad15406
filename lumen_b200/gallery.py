"""N4: a device-resident fp16 embedding gallery with incremental append and an on-disk format.

The reference keeps label / gallery embeddings as an fp32 `.npy` that `ModelResources` loads whole (lumen-resources
`loader.py:366`), normalises on the host when needed (`expert_bioclip/bioclip_model.py:251-256`) and multiplies on the CPU per
request.  Here the rows live in HBM as fp16 (what the fused GEMM + top-k kernel K10 reads):

* `DeviceGallery.from_npy(path)` memory-maps the fp32 file and streams it through pinned staging in chunks; the conversion
  (and the L2 normalisation, when asked for) runs on the device (`b200clip_gallery_pack`).  With `world > 1` each rank loads only its
  own row range (`ShardedGallery.shard_bounds`), so a gallery larger than one GPU's memory never exists in one place.
* `append(rows)` adds rows behind the existing ones (capacity grows geometrically; indices are stable), e.g. right after
  `embed_stream` produced them — device tensors are taken as they are, no host round trip.
* `save(path)` / `load(path)` use a plain fp16 `.npy` (so `np.load(..., mmap_mode="r")` and the reference's tooling still read it)
  plus `<path>.json` with `{"rows", "dim", "index_base", "normalized"}`; loading it is one H2D copy per chunk with no conversion.
* `search(queries, k)` is `ShardedGallery.search` on the filled prefix: K10 on this rank's rows, one all-gather, K12 merge.
"""
from __future__ import annotations

import ctypes as C
import json
from pathlib import Path

import numpy as np

from . import _lib
from .managers import ShardedGallery


def _torch():
    import torch
    return torch


class DeviceGallery:
    CHUNK_BYTES = 256 << 20          # pinned staging per chunk

    def __init__(self, engine, dim: int, capacity: int = 0, index_base: int = 0, group=None, normalized: bool = True) -> None:
        torch = _torch()
        self.engine, self.dim, self.index_base, self.group = engine, int(dim), int(index_base), group
        self.normalized = bool(normalized)
        self.rows = 0
        self._dev = f"cuda:{engine.device}"
        self._buf = torch.empty((max(int(capacity), 0), self.dim), dtype=torch.float16, device=self._dev)
        self._pin = None
        self._stage = None

    # ------------------------------------------------------------------ storage
    @property
    def capacity(self) -> int:
        return int(self._buf.shape[0])

    def view(self):
        """The filled rows, fp16 [rows, dim], no copy."""
        return self._buf[:self.rows]

    def reserve(self, rows: int) -> None:
        torch = _torch()
        if rows <= self.capacity:
            return
        new = torch.empty((max(rows, int(self.capacity * 1.5) + 1024), self.dim), dtype=torch.float16, device=self._dev)
        new[:self.rows].copy_(self._buf[:self.rows])
        self._buf = new

    def _pack(self, src_dev_f32, rows: int, normalize: bool) -> None:
        dst = self._buf[self.rows:self.rows + rows]
        _lib.check(self.engine.lib.b200clip_gallery_pack(self.engine._h, C.c_void_p(src_dev_f32.data_ptr()), rows, self.dim, int(normalize),
                                                         C.c_void_p(dst.data_ptr()), self.engine._stream()))
        self.rows += rows

    def append(self, rows, normalize: bool = True) -> tuple[int, int]:
        """rows: float32 [n, dim] as numpy (host, possibly memory-mapped) or a CUDA tensor; float16 device rows are copied as they
        are.  Returns (global index of the first new row, n)."""
        torch = _torch()
        first = self.index_base + self.rows
        n = int(rows.shape[0])
        if n == 0:
            return first, 0
        if rows.shape[1] != self.dim:
            raise ValueError(f"gallery rows have dim {rows.shape[1]}, expected {self.dim}")
        self.reserve(self.rows + n)
        if isinstance(rows, torch.Tensor):
            if rows.dtype == torch.float16:
                self._buf[self.rows:self.rows + n].copy_(rows)
                self.rows += n
            else:
                self._pack(rows.to(device=self._dev, dtype=torch.float32).contiguous(), n, normalize)
            return first, n
        per = max(1, self.CHUNK_BYTES // (4 * self.dim))
        if self._pin is None or self._pin.shape[0] < min(per, n):
            self._pin = torch.empty((min(per, max(n, 1024)), self.dim), dtype=torch.float32, pin_memory=True)
            self._stage = torch.empty_like(self._pin, device=self._dev)
        per = int(self._pin.shape[0])
        pin_np = self._pin.numpy()
        for lo in range(0, n, per):
            m = min(per, n - lo)
            np.copyto(pin_np[:m], rows[lo:lo + m], casting="same_kind")      # reads the mmap pages
            self._stage[:m].copy_(self._pin[:m], non_blocking=True)
            self._pack(self._stage, m, normalize)
            self.engine.sync()                                                # the staging buffers are reused by the next chunk
        return first, n

    # ------------------------------------------------------------------ files
    @classmethod
    def from_npy(cls, engine, path, normalize: bool = True, rank: int = 0, world: int = 1, group=None) -> "DeviceGallery":
        """fp32 [N, D] `.npy` (the reference's format) -> this rank's shard on the device."""
        arr = np.load(path, mmap_mode="r")
        if arr.ndim != 2 or arr.dtype != np.float32:
            raise ValueError(f"{path}: expected a float32 [N, D] array, got {arr.dtype} {arr.shape}")
        lo, hi = ShardedGallery.shard_bounds(arr.shape[0], world, rank)
        g = cls(engine, arr.shape[1], capacity=hi - lo, index_base=lo, group=group, normalized=normalize)
        g.append(arr[lo:hi], normalize=normalize)
        return g

    def save(self, path) -> None:
        path = Path(path)
        np.save(path, self.view().cpu().numpy())
        meta = {"rows": self.rows, "dim": self.dim, "index_base": self.index_base, "normalized": self.normalized, "dtype": "float16"}
        Path(str(path) + ".json").write_text(json.dumps(meta))

    @classmethod
    def load(cls, engine, path, group=None) -> "DeviceGallery":
        torch = _torch()
        path = Path(path)
        arr = np.load(path, mmap_mode="r")
        if arr.ndim != 2 or arr.dtype != np.float16:
            raise ValueError(f"{path}: expected a float16 [N, D] gallery file, got {arr.dtype} {arr.shape}")
        meta_p = Path(str(path) + ".json")
        meta = json.loads(meta_p.read_text()) if meta_p.exists() else {}
        g = cls(engine, arr.shape[1], capacity=arr.shape[0], index_base=int(meta.get("index_base", 0)), group=group,
                normalized=bool(meta.get("normalized", True)))
        per = max(1, cls.CHUNK_BYTES // (2 * g.dim))
        for lo in range(0, arr.shape[0], per):
            g._buf[lo:lo + per].copy_(torch.from_numpy(np.array(arr[lo:lo + per])))
        g.rows = int(arr.shape[0])
        return g

    # ------------------------------------------------------------------ search
    def search(self, queries_f16, k: int, events: list | None = None):
        """(scores float32 [Q, k], global indices int64 [Q, k]); across ranks when `group` spans several."""
        return ShardedGallery(self.engine, self.view(), self.index_base, self.group).search(queries_f16, k, events)
