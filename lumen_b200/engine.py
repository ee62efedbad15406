"""Thin host wrapper over one libb200clip handle (one per GPU).  PyTorch appears here only to own device /
pinned memory and to name the current CUDA stream; all arithmetic happens inside the library."""
from __future__ import annotations

import ctypes as C
import os
from collections.abc import Sequence

import numpy as np

from . import _lib
from .vit_config import OPENAI_MEAN, OPENAI_STD

MODE_SQUASH, MODE_SHORTEST_CROP = 0, 1
LAYOUT_NCHW, LAYOUT_PATCH = 0, 1


def _torch():
    import torch
    return torch


def pack_images(images: Sequence[np.ndarray], out: np.ndarray | None = None):
    """Packs decoded uint8 HWC RGB images at 16-byte aligned offsets (+16 spare bytes), as the C ABI expects.
    Returns (buffer, offsets[int64], hw[int32 Bx2], total_bytes)."""
    B = len(images)
    offsets = np.zeros(B, np.int64)
    hw = np.zeros((B, 2), np.int32)
    pos = 0
    for i, im in enumerate(images):
        if im.ndim != 3 or im.shape[2] != 3 or im.dtype != np.uint8:
            raise ValueError("images must be uint8 arrays of shape (H, W, 3)")
        offsets[i] = pos
        hw[i] = im.shape[:2]
        pos += (im.shape[0] * im.shape[1] * 3 + 15) & ~15
    total = pos + 16
    if out is None or out.size < total:
        out = np.empty(total, np.uint8)
    for i, im in enumerate(images):
        n = im.shape[0] * im.shape[1] * 3
        out[offsets[i]:offsets[i] + n] = np.ascontiguousarray(im).reshape(-1)
    return out, offsets, hw, total


class Engine:
    """Owns a b200clip handle.  `cfg` is a dict from vit_config.VIT_CONFIGS (or layers=0 for no encoder)."""

    def __init__(self, cfg: dict | None = None, device: int = 0, max_batch: int = 512, image_size: int | None = None,
                 mean=OPENAI_MEAN, std=OPENAI_STD):
        self.lib = _lib.load()
        cfg = dict(cfg or {})
        c = _lib.Config()
        c.device = device
        c.image_size = image_size or cfg.get("image_size", 224)
        c.patch = cfg.get("patch", 0)
        c.width = cfg.get("width", 0)
        c.layers = cfg.get("layers", 0)
        c.heads = cfg.get("heads", 0)
        c.mlp_dim = cfg.get("mlp_dim", 0)
        c.embed_dim = cfg.get("embed_dim", 0)
        c.activation = 0 if cfg.get("activation", "quick_gelu") == "quick_gelu" else 1
        c.max_batch = max_batch
        c.ln_eps = cfg.get("ln_eps", 1e-5)
        c.mean = (C.c_float * 3)(*mean)
        c.std = (C.c_float * 3)(*std)
        self.cfg, self.device, self.image_size = cfg, device, c.image_size
        self.embed_dim = c.embed_dim
        self.max_batch = max_batch
        self._h = C.c_void_p()
        _lib.check(self.lib.b200clip_create(C.byref(c), C.byref(self._h)))
        self._pin_in = None
        self.text_cfg = None
        self._jpeg_dev = None
        self._pin_out = None
        self._twin = None

    # ------------------------------------------------------------------ lifecycle
    def clone(self) -> "Engine":
        """A second engine on the same GPU sharing this one's image-tower weights, with activations and scratch of its own
        (b200clip_clone): two batches can then be in flight on two streams.  Weights must be loaded first."""
        other = object.__new__(Engine)
        other.lib, other.cfg, other.device, other.image_size = self.lib, self.cfg, self.device, self.image_size
        other.embed_dim, other.max_batch = self.embed_dim, self.max_batch
        other._h = C.c_void_p()
        _lib.check(self.lib.b200clip_clone(self._h, C.byref(other._h)))
        other._pin_in = other._jpeg_dev = other._pin_out = other._twin = None
        other.text_cfg = None
        return other

    def twin(self) -> "Engine":
        """The engine `embed_stream` alternates with (made on first use, closed with this one)."""
        if self._twin is None:
            self._twin = self.clone()
        return self._twin

    def close(self) -> None:
        if getattr(self, "_twin", None) is not None:
            self._twin.close()
            self._twin = None
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.b200clip_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # noqa: D105
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def _stream(self):
        torch = _torch()
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def sync(self) -> None:
        _lib.check(self.lib.b200clip_sync(self._h, self._stream()))

    @property
    def launches(self) -> int:
        return int(self.lib.b200clip_launch_count(self._h))

    def profile(self, enable: bool) -> None:
        _lib.check(self.lib.b200clip_profile(self._h, int(enable)))

    def profile_read(self) -> dict:
        ms = (C.c_float * 4)()
        n = (C.c_int64 * 4)()
        _lib.check(self.lib.b200clip_profile_read(self._h, ms, n, self._stream()))
        names = ("preprocess", "gemm", "norm", "attention")
        return {k: {"ms": float(ms[i]), "launches": int(n[i])} for i, k in enumerate(names)}

    # ------------------------------------------------------------------ weights
    def load_weights(self, weights: dict[str, np.ndarray]) -> None:
        for name, w in weights.items():
            if name.endswith("position_ids"):
                continue
            a = np.ascontiguousarray(w, dtype=np.float32)
            shape = (C.c_int64 * a.ndim)(*a.shape)
            _lib.check(self.lib.b200clip_load_weight(self._h, name.encode(), a.ctypes.data_as(C.c_void_p), shape, a.ndim))
        _lib.check(self.lib.b200clip_finalize_weights(self._h))

    # ------------------------------------------------------------------ N4: text tower
    def init_text(self, cfg: dict, max_batch: int = 256) -> None:
        """Adds a CLIP text tower (TEXT_CONFIGS entry) to this engine; weights follow through load_weights (text_model.* keys)."""
        c = _lib.TextConfig()
        c.context_length, c.vocab_size, c.width, c.layers = cfg["context_length"], cfg["vocab_size"], cfg["width"], cfg["layers"]
        c.heads, c.mlp_dim, c.embed_dim = cfg["heads"], cfg["mlp_dim"], cfg["embed_dim"]
        c.activation = 0 if cfg.get("activation", "quick_gelu") == "quick_gelu" else 1
        c.max_batch, c.ln_eps = int(max_batch), float(cfg.get("ln_eps", 1e-5))
        _lib.check(self.lib.b200clip_text_init(self._h, C.byref(c)))
        self.text_cfg = dict(cfg)

    def encode_text(self, ids) -> "np.ndarray":
        """Token ids int [B, context_length] (numpy or cuda tensor) -> float32 [B, D] unit vectors (numpy)."""
        torch = _torch()
        if getattr(self, "text_cfg", None) is None:
            raise _lib.B200ClipError(-3, "encode_text: init_text was not called")
        if not isinstance(ids, torch.Tensor):
            ids = torch.from_numpy(np.ascontiguousarray(ids, dtype=np.int32))
        if ids.ndim != 2 or ids.shape[1] != self.text_cfg["context_length"]:
            raise _lib.B200ClipError(-1, f"encode_text: ids must be [B, {self.text_cfg['context_length']}]")
        if ids.shape[0] == 0:
            return np.empty((0, self.text_cfg["embed_dim"]), np.float32)
        if int(ids.min()) < 0 or int(ids.max()) >= self.text_cfg["vocab_size"]:
            raise _lib.B200ClipError(-1, "encode_text: token id outside the vocabulary")
        ids = ids.to(device=f"cuda:{self.device}", dtype=torch.int32).contiguous()
        B = int(ids.shape[0])
        out = torch.empty((B, self.text_cfg["embed_dim"]), dtype=torch.float32, device=ids.device)
        _lib.check(self.lib.b200clip_encode_text(self._h, C.c_void_p(ids.data_ptr()), B, C.c_void_p(out.data_ptr()), self._stream()))
        return out.cpu().numpy()

    # ------------------------------------------------------------------ device-buffer API
    def preprocess(self, src_dev, offsets: np.ndarray, hw: np.ndarray, mode: int = MODE_SQUASH, layout: int = LAYOUT_NCHW,
                   bf16: bool = False):
        """src_dev: uint8 cuda tensor (packed).  Returns a cuda tensor: [B,3,S,S] or patch-major [B*n, kpad]."""
        torch = _torch()
        B, S = int(len(offsets)), self.image_size
        dt = torch.bfloat16 if bf16 else torch.float32
        if layout == LAYOUT_NCHW:
            out = torch.empty((B, 3, S, S), dtype=dt, device=src_dev.device)
        else:
            P = self.cfg["patch"]
            kpad = (3 * P * P + 63) // 64 * 64
            out = torch.zeros((B * (S // P) ** 2, kpad), dtype=dt, device=src_dev.device)
        offsets = np.ascontiguousarray(offsets, np.int64)
        hw = np.ascontiguousarray(hw, np.int32)
        _lib.check(self.lib.b200clip_preprocess(self._h, C.c_void_p(src_dev.data_ptr()), offsets.ctypes.data_as(C.c_void_p),
                                                hw.ctypes.data_as(C.c_void_p), B, mode, layout, int(bf16),
                                                C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def encode(self, patches_dev, B: int):
        torch = _torch()
        out = torch.empty((B, self.embed_dim), dtype=torch.float32, device=patches_dev.device)
        _lib.check(self.lib.b200clip_encode(self._h, C.c_void_p(patches_dev.data_ptr()), B, C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def embed_u8(self, src_dev, offsets: np.ndarray, hw: np.ndarray, mode: int = MODE_SQUASH, out=None):
        torch = _torch()
        B = int(len(offsets))
        if out is None:
            out = torch.empty((B, self.embed_dim), dtype=torch.float32, device=src_dev.device)
        offsets = np.ascontiguousarray(offsets, np.int64)
        hw = np.ascontiguousarray(hw, np.int32)
        _lib.check(self.lib.b200clip_embed_u8(self._h, C.c_void_p(src_dev.data_ptr()), offsets.ctypes.data_as(C.c_void_p),
                                              hw.ctypes.data_as(C.c_void_p), B, mode, C.c_void_p(out.data_ptr()), self._stream()))
        return out

    # ------------------------------------------------------------------ N2: encoded bytes -> device RGB -> embeddings
    @staticmethod
    def jpeg_probe(data: bytes):
        """(H, W, components) from the headers, or None when this decoder does not cover the stream (see include/b200clip.h)."""
        hw = (C.c_int32 * 2)()
        nc = C.c_int32()
        rc = _lib.load().b200clip_jpeg_probe(data, len(data), hw, C.byref(nc))
        return (int(hw[0]), int(hw[1]), int(nc.value)) if rc == 0 else None

    def decode_jpeg(self, images: Sequence[bytes], host_threads: int = 0, extra_bytes: int = 0):
        """Baseline JPEG byte strings -> packed ragged RGB in device memory (Pillow / libjpeg-turbo exact).
        Returns (uint8 device tensor, offsets int64[n], hw int32[n,2], status int32[n], used_bytes); status 1 marks images the
        device decoder does not cover (nothing written, hw = 0).  `extra_bytes` reserves room behind the decoded images."""
        torch = _torch()
        n = len(images)
        offsets, hw, status = np.zeros(n, np.int64), np.zeros((n, 2), np.int32), np.zeros(n, np.int32)
        need = 16
        for b in images:
            info = self.jpeg_probe(b)
            if info is not None:
                need += (info[0] * info[1] * 3 + 15) & ~15
        cap = need + int(extra_bytes)
        if self._jpeg_dev is None or self._jpeg_dev.numel() < cap:
            self._jpeg_dev = torch.empty(int(cap * 1.25), dtype=torch.uint8, device=f"cuda:{self.device}")
        ptrs = (C.c_char_p * max(n, 1))(*images)
        lens = (C.c_size_t * max(n, 1))(*[len(b) for b in images])
        total = C.c_int64()
        _lib.check(self.lib.b200clip_decode_jpeg(self._h, ptrs, lens, n, C.c_void_p(self._jpeg_dev.data_ptr()), self._jpeg_dev.numel(),
                                                 offsets.ctypes.data_as(C.c_void_p), hw.ctypes.data_as(C.c_void_p),
                                                 status.ctypes.data_as(C.c_void_p), C.byref(total), int(host_threads), self._stream()))
        return self._jpeg_dev, offsets, hw, status, int(total.value)

    def embed_bytes(self, images: Sequence[bytes], mode: int = MODE_SQUASH, host_decode=None, host_threads: int = 0) -> np.ndarray:
        """image_batch_to_vectors on encoded bytes: JPEGs the device decoder covers never exist as pixels on the host; the rest
        go through `host_decode(bytes) -> uint8 HWC` (the reference's PIL call) and are uploaded behind them."""
        torch = _torch()
        n = len(images)
        if n == 0:
            return np.empty((0, self.embed_dim), np.float32)
        probes = [self.jpeg_probe(b) for b in images]
        others = {i: np.array(host_decode(images[i]), dtype=np.uint8, order="C") for i in range(n) if probes[i] is None} if host_decode else {}
        if any(p is None for p in probes) and not host_decode:
            raise _lib.B200ClipError(-1, "embed_bytes: an image is not a baseline JPEG and no host decoder was given")
        extra = sum(((a.size + 15) & ~15) for a in others.values())
        dev, offsets, hw, status, used = self.decode_jpeg(images, host_threads, extra)
        # streams the device decoder found damaged after a clean header (status 1): the host decoder decides, per image, whether
        # they still decode (PIL accepts many slightly damaged files with a warning) or raise
        late = [i for i in range(n) if status[i] != 0 and i not in others]
        if late:
            if not host_decode:
                raise _lib.B200ClipError(-1, f"embed_bytes: image {late[0]} is damaged and no host decoder was given")
            for i in late:
                others[i] = np.array(host_decode(images[i]), dtype=np.uint8, order="C")
            need = used + sum(((a.size + 15) & ~15) for a in others.values()) + 16
            if dev.numel() < need:
                grown = torch.empty(int(need * 1.25), dtype=torch.uint8, device=dev.device)
                grown[:used].copy_(dev[:used])
                self._jpeg_dev = dev = grown
        for i, a in others.items():
            if a.ndim != 3 or a.shape[2] != 3:
                raise _lib.B200ClipError(-1, "embed_bytes: host decoder must return H x W x 3 uint8")
            dev[used:used + a.size].copy_(torch.from_numpy(a.reshape(-1)), non_blocking=False)
            offsets[i], hw[i] = used, a.shape[:2]
            used += (a.size + 15) & ~15
        out = self.embed_u8(dev, offsets, hw, mode)
        return out.cpu().numpy()

    # ------------------------------------------------------------------ host-buffer API (what the backend calls)
    def embed_host(self, images: Sequence[np.ndarray], mode: int = MODE_SQUASH) -> np.ndarray:
        """Decoded uint8 HWC images (host) -> float32 [B, D] unit vectors (host).  H2D, kernels, D2H, sync."""
        torch = _torch()
        B = len(images)
        if B == 0:
            return np.empty((0, self.embed_dim), np.float32)
        need = sum(((im.shape[0] * im.shape[1] * 3 + 15) & ~15) for im in images) + 16
        if self._pin_in is None or self._pin_in.numel() < need:
            self._pin_in = torch.empty(int(need * 1.25), dtype=torch.uint8, pin_memory=True)
        if self._pin_out is None or self._pin_out.numel() < B * self.embed_dim:
            self._pin_out = torch.empty(B * self.embed_dim, dtype=torch.float32, pin_memory=True)
        buf, offsets, hw, total = pack_images(images, self._pin_in.numpy())
        return self.embed_packed_host(self._pin_in, total, offsets, hw, mode)

    def embed_packed_host(self, pinned_u8, total: int, offsets: np.ndarray, hw: np.ndarray, mode: int = MODE_SQUASH) -> np.ndarray:
        torch = _torch()
        B = int(len(offsets))
        if self._pin_out is None or self._pin_out.numel() < B * self.embed_dim:
            self._pin_out = torch.empty(B * self.embed_dim, dtype=torch.float32, pin_memory=True)
        offsets = np.ascontiguousarray(offsets, np.int64)
        hw = np.ascontiguousarray(hw, np.int32)
        _lib.check(self.lib.b200clip_embed_host(self._h, C.c_void_p(pinned_u8.data_ptr()), int(total),
                                                offsets.ctypes.data_as(C.c_void_p), hw.ctypes.data_as(C.c_void_p), B, mode,
                                                C.c_void_p(self._pin_out.data_ptr()), self._stream()))
        return self._pin_out[:B * self.embed_dim].view(B, self.embed_dim).numpy().copy()

    def embed_stream(self, batches, mode: int = MODE_SQUASH):
        """Pipelined form of embed_packed_host for a stream of batches (serving / bulk indexing).

        `batches` yields (pinned_u8 tensor, total_bytes, offsets, hw) in the pack_images layout.  The host->device
        copy of batch i+1 runs on a side stream while batch i is preprocessed and encoded, results come back through
        two pinned buffers; one numpy [B, D] array is yielded per batch, in order.  Every byte still crosses PCIe
        inside the loop - only the waiting is overlapped.

        Two details matter for the overlap: both streams are non-default (the legacy default stream serialises
        against every other stream), and batch i is enqueued BEFORE the upload of batch i+1 - the copy engine serves
        host->device transfers in submission order, so the small descriptor upload of a batch must not queue up
        behind a big image upload that is still waiting for its buffer."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        caller = torch.cuda.current_stream(dev)
        # Two batches in flight: even batches on this engine, odd ones on its twin (same weights, own activations), each on a compute
        # stream of its own.  One handle serialises its calls; with two, the memory-bound kernels of one batch (LayerNorm, attention,
        # out-projection) run under the tensor-bound GEMMs of the other (+4-5 % images/s at batch 512).  B200_STREAM_TWIN=0: one engine.
        two = os.environ.get("B200_STREAM_TWIN", "1") != "0" and self.cfg.get("layers", 0) > 0
        engines = [self, self.twin()] if two else [self, self]
        computes = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)] if two else [torch.cuda.Stream(device=dev)] * 2
        copy = torch.cuda.Stream(device=dev)
        for c in set(computes):
            c.wait_stream(caller)
        NS = 4 if two else 2                                   # buffer slots: two per engine, so an upload never waits for the engine it feeds
        src = [None] * NS; out = [None] * NS; pin_out = [None] * NS; meta = [None] * NS
        ready = [torch.cuda.Event() for _ in range(NS)]        # H2D of the slot finished
        done = [torch.cuda.Event() for _ in range(NS)]         # D2H of the slot's result finished

        def stage(i, item):
            pinned, total, offsets, hw = item
            s, need = i % NS, int(total)
            if src[s] is None or src[s].numel() < need:
                src[s] = torch.empty(int(need * 1.25) + 64, dtype=torch.uint8, device=dev)
            with torch.cuda.stream(copy):
                if i >= NS:
                    # wait for the END of the batch that used the slot (its result download included), not just for its last read of
                    # the slot: released any earlier, this big upload grabs the copy engine just before the next batch's small
                    # descriptor upload becomes eligible and delays that batch's kernels by a whole transfer
                    copy.wait_event(done[s])
                src[s][:need].copy_(pinned[:need], non_blocking=True)
                ready[s].record(copy)
            meta[s] = (np.ascontiguousarray(offsets, np.int64), np.ascontiguousarray(hw, np.int32))

        def launch(i):
            s, e = i % NS, i & 1
            offsets, hw = meta[s]
            B = int(len(offsets))
            if out[s] is None or out[s].shape[0] < B:
                out[s] = torch.empty((B, self.embed_dim), dtype=torch.float32, device=dev)
                pin_out[s] = torch.empty((B, self.embed_dim), dtype=torch.float32, pin_memory=True)
            with torch.cuda.stream(computes[e]):
                computes[e].wait_event(ready[s])
                engines[e].embed_u8(src[s], offsets, hw, mode, out=out[s][:B])
                pin_out[s][:B].copy_(out[s][:B], non_blocking=True)
                done[s].record(computes[e])
            return B

        finished = False
        try:
            it = iter(batches)
            pending = []                                           # (index, B): launched, result not yet yielded
            ahead = NS // 2                                        # uploads staged ahead of the launches
            staged = []
            for _ in range(ahead):
                nxt = next(it, None)
                if nxt is None:
                    break
                stage(len(staged), nxt)
                staged.append(True)
            n_staged = len(staged)
            i = 0
            while i < n_staged:
                while pending and pending[0][0] <= i - NS:         # this batch reuses the result buffers of batch i - NS
                    pi, pB = pending.pop(0)
                    done[pi % NS].synchronize()
                    yield pin_out[pi % NS][:pB].numpy().copy()
                pending.append((i, launch(i)))
                nxt = next(it, None)
                if nxt is not None:
                    stage(n_staged, nxt)                           # upload overlaps the encodes in flight (its slot is guarded on the device)
                    n_staged += 1
                i += 1
            for pi, pB in pending:
                done[pi % NS].synchronize()
                yield pin_out[pi % NS][:pB].numpy().copy()
            for e, c in {id(c): (e, c) for e, c in zip(engines, computes)}.values():
                with torch.cuda.stream(c):
                    e.sync()
                caller.wait_stream(c)
            finished = True
        finally:
            if not finished:
                # the consumer stopped early (or a call failed): nothing may still be running on the side streams when the staging
                # buffers of this generator are released
                copy.synchronize()
                for c in set(computes):
                    c.synchronize()

    # ------------------------------------------------------------------ classify / search
    def set_labels(self, labels: np.ndarray, slot: int = 0) -> None:
        a = np.ascontiguousarray(labels, dtype=np.float32)
        _lib.check(self.lib.b200clip_set_labels(self._h, slot, a.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1], self._stream()))

    def classify(self, emb_dev, k: int, slot: int = 0, scale: float = 100.0, softmax: bool = True):
        torch = _torch()
        B = emb_dev.shape[0]
        vals = torch.empty((B, k), dtype=torch.float32, device=emb_dev.device)
        idx = torch.empty((B, k), dtype=torch.int64, device=emb_dev.device)
        _lib.check(self.lib.b200clip_classify(self._h, slot, C.c_void_p(emb_dev.data_ptr()), B, float(scale), int(softmax), k,
                                              C.c_void_p(vals.data_ptr()), C.c_void_p(idx.data_ptr()), self._stream()))
        return vals, idx

    def topk(self, queries_f16, gallery_f16, k: int, index_base: int = 0):
        torch = _torch()
        Q, D = queries_f16.shape
        N = gallery_f16.shape[0]
        assert queries_f16.dtype == torch.float16 and gallery_f16.dtype == torch.float16
        assert queries_f16.is_contiguous() and gallery_f16.is_contiguous()
        if N == 0:                                              # an empty shard (more ranks than rows): nothing to report
            return (torch.full((Q, k), float("-inf"), dtype=torch.float32, device=queries_f16.device),
                    torch.full((Q, k), -1, dtype=torch.int64, device=queries_f16.device))
        scores = torch.empty((Q, k), dtype=torch.float32, device=queries_f16.device)
        idx = torch.empty((Q, k), dtype=torch.int64, device=queries_f16.device)
        _lib.check(self.lib.b200clip_topk(self._h, C.c_void_p(queries_f16.data_ptr()), Q, C.c_void_p(gallery_f16.data_ptr()), N, D, k,
                                          int(index_base), C.c_void_p(scores.data_ptr()), C.c_void_p(idx.data_ptr()), self._stream()))
        return scores, idx

    def topk_keys(self, queries_f16, gallery_f16, k: int):
        """This shard's top-k as 64-bit exchange keys [Q, k] (int64 storage): score bits | ~(row inside the shard), 0 = empty."""
        torch = _torch()
        Q, D = queries_f16.shape
        assert queries_f16.dtype == torch.float16 and gallery_f16.dtype == torch.float16
        assert queries_f16.is_contiguous() and gallery_f16.is_contiguous()
        if gallery_f16.shape[0] == 0:                           # an empty shard contributes empty slots (key 0)
            return torch.zeros((Q, k), dtype=torch.int64, device=queries_f16.device)
        keys = torch.empty((Q, k), dtype=torch.int64, device=queries_f16.device)
        _lib.check(self.lib.b200clip_topk_keys(self._h, C.c_void_p(queries_f16.data_ptr()), Q, C.c_void_p(gallery_f16.data_ptr()), gallery_f16.shape[0],
                                               D, k, C.c_void_p(keys.data_ptr()), self._stream()))
        return keys

    def search_sharded(self, queries_f16, gallery_shard_f16, k: int, index_base: int, nccl_comm: int, world: int):
        """Sharded search in one C-ABI call on a raw ncclComm_t (an integer address): local top-k, grouped all-gather, merge.
        Every rank gets the global ([Q,k] f32 scores, [Q,k] i64 indices)."""
        torch = _torch()
        Q, D = queries_f16.shape
        N = 0 if gallery_shard_f16 is None else gallery_shard_f16.shape[0]
        assert queries_f16.dtype == torch.float16 and queries_f16.is_contiguous()
        assert N == 0 or (gallery_shard_f16.dtype == torch.float16 and gallery_shard_f16.is_contiguous())
        scores = torch.empty((Q, k), dtype=torch.float32, device=queries_f16.device)
        idx = torch.empty((Q, k), dtype=torch.int64, device=queries_f16.device)
        _lib.check(self.lib.b200clip_search_sharded(self._h, C.c_void_p(queries_f16.data_ptr()), Q,
                                                    C.c_void_p(gallery_shard_f16.data_ptr() if N else 0), N, D, k, int(index_base),
                                                    C.c_void_p(nccl_comm), world, C.c_void_p(scores.data_ptr()), C.c_void_p(idx.data_ptr()),
                                                    self._stream()))
        return scores, idx

    def merge_topk_keys(self, part_keys, part_bases):
        """part_keys [P,Q,k] exchange keys, part_bases [P] int64 (cuda): index_base of every part -> merged ([Q,k] f32, [Q,k] i64)."""
        torch = _torch()
        P, Q, k = part_keys.shape
        scores = torch.empty((Q, k), dtype=torch.float32, device=part_keys.device)
        idx = torch.empty((Q, k), dtype=torch.int64, device=part_keys.device)
        _lib.check(self.lib.b200clip_merge_topk_keys(self._h, C.c_void_p(part_keys.data_ptr()), C.c_void_p(part_bases.data_ptr()), P, Q, k,
                                                     C.c_void_p(scores.data_ptr()), C.c_void_p(idx.data_ptr()), self._stream()))
        return scores, idx

    def merge_topk(self, part_scores, part_idx):
        """part_scores [P,Q,k] f32, part_idx [P,Q,k] i64 (cuda) -> merged ([Q,k], [Q,k])."""
        torch = _torch()
        P, Q, k = part_scores.shape
        scores = torch.empty((Q, k), dtype=torch.float32, device=part_scores.device)
        idx = torch.empty((Q, k), dtype=torch.int64, device=part_scores.device)
        _lib.check(self.lib.b200clip_merge_topk(self._h, C.c_void_p(part_scores.data_ptr()), C.c_void_p(part_idx.data_ptr()), P, Q, k,
                                                C.c_void_p(scores.data_ptr()), C.c_void_p(idx.data_ptr()), self._stream()))
        return scores, idx

    # ------------------------------------------------------------------ building blocks (parity tests)
    def gemm(self, A_bf16, W_bf16, bias=None, epilogue: int = _lib.EPI_F32, activation: int = 0, out=None):
        torch = _torch()
        M, K = A_bf16.shape
        N = W_bf16.shape[0]
        if out is None:
            dt = torch.float32 if epilogue in (_lib.EPI_F32, _lib.EPI_BIAS_RES_F32) else torch.bfloat16
            out = torch.zeros((M, N), dtype=dt, device=A_bf16.device)
        _lib.check(self.lib.b200clip_gemm_bf16(self._h, C.c_void_p(A_bf16.data_ptr()), C.c_void_p(W_bf16.data_ptr()), M, N, K,
                                               C.c_void_p(bias.data_ptr()) if bias is not None else None,
                                               C.c_void_p(out.data_ptr()), epilogue, activation, self._stream()))
        return out

    def attention(self, qkv_bf16, B: int, S: int, width: int, heads: int, out=None, causal: bool = False):
        torch = _torch()
        if out is None:
            out = torch.empty((B * S, width), dtype=torch.bfloat16, device=qkv_bf16.device)
        _lib.check(self.lib.b200clip_attention(self._h, C.c_void_p(qkv_bf16.data_ptr()), C.c_void_p(out.data_ptr()), B, S, width, heads,
                                               int(causal), self._stream()))
        return out

    def layernorm(self, x_f32, gamma, beta, eps: float = 1e-5):
        torch = _torch()
        rows, d = x_f32.shape
        out = torch.empty((rows, d), dtype=torch.bfloat16, device=x_f32.device)
        _lib.check(self.lib.b200clip_layernorm(self._h, C.c_void_p(x_f32.data_ptr()), C.c_void_p(gamma.data_ptr()),
                                               C.c_void_p(beta.data_ptr()), C.c_void_p(out.data_ptr()), rows, d, float(eps), self._stream()))
        return out
