/*
 * libb200clip — C ABI of the B200-native lumen-clip image path.
 *
 * The reference (EdwinZhanCN/Lumen) is pure Python and has no FFI of its own; these entry points are what
 * a ctypes binding for its backend seam needs.  Each declaration cites the reference interface it stands
 * behind (paths relative to packages/lumen-clip/src/lumen_clip/).  See INTEGRATION.md for the binding.
 *
 * Conventions
 *   - every function returns 0 on success or a negative code (B200CLIP_E*); the message of the last failure
 *     on the calling thread is available from b200clip_last_error();
 *   - the caller owns every buffer it passes (PyTorch allocates device tensors and hands over data_ptr());
 *     the library owns only what lives inside a handle: weights, activations workspace, resize tables,
 *     search scratch;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream); calls only enqueue work
 *     unless stated otherwise; one handle per GPU; calls on one handle are serialised by an internal mutex,
 *     so the 10 gRPC worker threads of the reference server (server.py:294-297) may share it;
 *   - there is no CPU fallback: without a CUDA device b200clip_create fails.
 */
#ifndef B200CLIP_H
#define B200CLIP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200CLIP_OK        0
#define B200CLIP_EINVAL   -1   /* bad argument                       */
#define B200CLIP_ECUDA    -2   /* CUDA runtime / driver error        */
#define B200CLIP_ESTATE   -3   /* wrong call order, missing weights  */
#define B200CLIP_EDEVICE  -4   /* no usable sm_100 device            */
#define B200CLIP_EKERNEL  -5   /* device-side watchdog tripped       */

typedef struct b200clip_handle b200clip_handle;

/* Model + preprocessing description: what ModelResources provides (resources/loader.py:80-139) plus the
 * ViT shape.  layers == 0 builds a handle without an encoder (preprocess / classify / search only). */
typedef struct b200clip_config {
    int32_t device;          /* CUDA ordinal */
    int32_t image_size;      /* S: get_image_size() */
    int32_t patch;           /* P */
    int32_t width;           /* d */
    int32_t layers;          /* L */
    int32_t heads;
    int32_t mlp_dim;
    int32_t embed_dim;       /* D */
    int32_t activation;      /* 0 = QuickGELU (OpenAI / HF weights), 1 = exact GELU (LAION open_clip weights) */
    int32_t max_batch;       /* images per encoder pass; larger requests are chunked internally */
    float   ln_eps;
    float   mean[3];         /* get_normalization_stats() */
    float   std[3];
} b200clip_config;

/* N4 — text tower shapes (CLIPTextConfig): ViT-B/32's is context 77, vocab 49408, width 512, 8 heads, 12 layers, mlp 2048, D 512. */
typedef struct b200clip_text_config {
    int32_t context_length;  /* tokens per sequence (77); must exceed 64 (the causal mask lives in the tcgen05 attention kernel) */
    int32_t vocab_size;
    int32_t width;           /* heads * 64, a multiple of 128, at most 1024 */
    int32_t layers;
    int32_t heads;
    int32_t mlp_dim;
    int32_t embed_dim;       /* D: must equal the image tower's for classify / search */
    int32_t activation;      /* 0 = QuickGELU, 1 = exact GELU */
    int32_t max_batch;       /* sequences per pass; larger requests are chunked */
    float   ln_eps;
} b200clip_text_config;

/* lifecycle: BaseClipBackend.initialize()/close(), backends/base.py:126-134 */
int  b200clip_create(const b200clip_config* cfg, b200clip_handle** out);
/* A second handle on the same GPU that SHARES the image tower's weights with `src` (no copy; the device memory lives until the
 * last handle using it is destroyed) and owns its activations, descriptors and scratch.  No counterpart in the reference (one
 * ONNX session serves its ten worker threads one request at a time).  It exists so that two batches can be in flight: calls on one
 * handle are serialised, calls on two handles with two streams overlap, and the memory-bound kernels of one batch (LayerNorm,
 * attention, the out-projection) then run under the tensor-bound GEMMs of the other: +4-5 % images/s on ViT-B/32 at batch 512.
 * Weights are loaded through `src` only; a text tower and label sets are not cloned. */
int  b200clip_clone(b200clip_handle* src, b200clip_handle** out);
void b200clip_destroy(b200clip_handle* h);
const char* b200clip_last_error(void);
/* Synchronise `stream` and report a tripped device-side watchdog (B200CLIP_EKERNEL). */
int  b200clip_sync(b200clip_handle* h, void* stream);

/* Weight upload — replaces ort.InferenceSession(vision.onnx) / CLIPModel.from_pretrained
 * (backends/onnxrt_backend.py:166-170, backends/torch_backend.py:290-295).  `name` is a transformers key of the
 * vision tower (vision_model.embeddings.patch_embedding.weight, ...encoder.layers.N.self_attn.q_proj.weight,
 * visual_projection.weight, ...); data is host float32, C-contiguous.  q/k/v are fused and everything the GEMMs
 * read is cast to bf16 on upload. */
int b200clip_load_weight(b200clip_handle* h, const char* name, const float* host_data, const int64_t* shape, int32_t ndim);
/* Verifies that every tensor of the configured tower has been loaded. */
int b200clip_finalize_weights(b200clip_handle* h);

/* K1 — preprocess: the closure at backends/onnxrt_backend.py:410-431 (mode 0, squash) and the open_clip transform of
 * backends/torch_backend.py:201-204 (mode 1, shortest side + centre crop), on decoded uint8 RGB.
 *   src_dev      packed HWC images in device memory, 16-byte aligned, 16 spare bytes after the last image
 *   offsets      host array [B]: byte offset of each image inside src_dev (multiples of 16)
 *   hw           host array [B][2]: height, width
 *   out_dev      layout 0: [B,3,S,S]; layout 1: patch-major [B*(S/P)^2, kpad] with kpad = 3*P*P rounded up to 64
 *   out_bf16     0 = float32 output, 1 = bfloat16 output */
int b200clip_preprocess(b200clip_handle* h, const uint8_t* src_dev, const int64_t* offsets, const int32_t* hw,
                        int32_t B, int32_t mode, int32_t layout, int32_t out_bf16, void* out_dev, void* stream);

/* E1+E2 — encoder + L2 normalise: sess_vision.run + embedding/||embedding|| (backends/onnxrt_backend.py:485-489,
 * 541-546).  patches_dev is the bf16 patch-major output of b200clip_preprocess(layout 1); emb_dev is float32 [B,D]
 * unit vectors.  B may exceed max_batch (chunked). */
int b200clip_encode(b200clip_handle* h, const void* patches_dev, int32_t B, float* emb_dev, void* stream);

/* E3 — image_batch_to_vectors on decoded images already in device memory (backends/onnxrt_backend.py:497-551):
 * preprocess -> encode -> normalise in one call. */
int b200clip_embed_u8(b200clip_handle* h, const uint8_t* src_dev, const int64_t* offsets, const int32_t* hw,
                      int32_t B, int32_t mode, float* emb_dev, void* stream);

/* The same from HOST buffers (the call a Python backend makes): copies `src_host` (packed like src_dev,
 * total_bytes long, ideally pinned) to the device, runs embed_u8, copies [B,D] float32 back into emb_host and
 * synchronises. */
int b200clip_embed_host(b200clip_handle* h, const uint8_t* src_host, int64_t total_bytes, const int64_t* offsets,
                        const int32_t* hw, int32_t B, int32_t mode, float* emb_host, void* stream);

/* N4 — text tower (label-embedding bootstrap: CLIPModelManager._compute_label_embeddings, general_clip/clip_model.py:145-172, which
 * calls backend.text_batch_to_vectors -> backends/torch_backend.py text path = CLIPTextModelWithProjection: token + position
 * embedding, L x [LN1, causal MHA, LN2, MLP], final LN, hidden state at the EOT token (argmax of the ids), text_projection, L2 norm).
 * b200clip_text_init adds the tower to a handle (the image tower may be absent: cfg.layers == 0); its tensors go through
 * b200clip_load_weight under their transformers names (text_model.embeddings.token_embedding.weight, text_model.embeddings.
 * position_embedding.weight, text_model.encoder.layers.N.*, text_model.final_layer_norm.*, text_projection.weight).
 * b200clip_encode_text: ids_dev int32 [B, context_length] in device memory (tokenisation stays with the caller's tokenizer)
 * -> emb_dev float32 [B, D] unit vectors. */
int b200clip_text_init(b200clip_handle* h, const b200clip_text_config* cfg);
int b200clip_encode_text(b200clip_handle* h, const int32_t* ids_dev, int32_t B, float* emb_dev, void* stream);

/* P0 / N2 — JPEG bytes -> decoded RGB on the device, in front of K1.  Replaces `Image.open(io.BytesIO(b)).convert("RGB")`
 * (backends/onnxrt_backend.py:478,412; backends/torch_backend.py:570,635), i.e. Pillow -> libjpeg-turbo defaults (islow IDCT,
 * fancy chroma upsampling, fixed-point YCbCr->RGB), bit for bit.  jpegs[i] / lens[i]: n byte strings in host memory.  Huffman
 * decoding runs on `host_threads` host threads (0 = all cores); IDCT, upsampling and colour conversion run on the device and
 * write image i as H x W x 3 uint8 at dst_dev + offsets_out[i] (16-byte aligned, the layout b200clip_preprocess reads);
 * hw_out[2i], hw_out[2i+1] = H, W; *total_out = bytes used.  status_out[i] = 0 decoded, 1 = left to the caller's own decoder:
 * a flavour this one does not cover (progressive, CMYK, 4:1:1, not a JPEG, more pixels than PIL's MAX_IMAGE_PIXELS ...) or a
 * stream it found damaged; nothing usable is written for it and hw_out is 0.  One bad image never fails the batch: whether a
 * damaged file still decodes (with a warning) or raises stays the decision of the reference's PIL call.  b200clip_jpeg_probe
 * reads only the headers (no device work). */
int b200clip_jpeg_probe(const void* jpeg, size_t len, int32_t* hw, int32_t* ncomp);
int b200clip_decode_jpeg(b200clip_handle* h, const void* const* jpegs, const size_t* lens, int32_t n, uint8_t* dst_dev,
                         int64_t dst_capacity, int64_t* offsets_out, int32_t* hw_out, int32_t* status_out, int64_t* total_out,
                         int32_t host_threads, void* stream);

/* N4 — gallery rows for b200clip_topk: fp32 [rows, D] in device memory -> fp16 at dst_f16_dev (the caller's gallery buffer, already
 * offset to the first free row), divided by the row's L2 norm when normalize != 0.  Replaces the host-side
 * `text_embeddings / np.linalg.norm(..., axis=1, keepdims=True)` of BioCLIPModelManager.initialize (expert_bioclip/bioclip_model.py:
 * 251-256) applied to the fp32 `.npy` that ModelResources loads (lumen-resources loader.py:366). */
int b200clip_gallery_pack(b200clip_handle* h, const float* src_dev, int64_t rows, int32_t D, int32_t normalize, void* dst_f16_dev, void* stream);

/* Label / gallery upload for classify: replaces ModelResources.label_embeddings (resources/loader.py:366); rows are
 * L2-normalised once on the device (general_clip/clip_model.py:289-291 does it per call). slot 0 = labels, 1 = scenes. */
int b200clip_set_labels(b200clip_handle* h, int32_t slot, const float* host_labels, int32_t N, int32_t D, void* stream);

/* K11 — CLIPModelManager.classify_image / classify_scene (general_clip/clip_model.py:274-315, 341-357) for a batch of
 * image embeddings in device memory: cosine, optional softmax(scale * s), top-k (value desc, index asc).
 * out_values float32 [B,k], out_index int64 [B,k] (device). */
int b200clip_classify(b200clip_handle* h, int32_t slot, const float* emb_dev, int32_t B, float scale, int32_t softmax,
                      int32_t k, float* out_values_dev, int64_t* out_index_dev, void* stream);

/* K10 — BioCLIPModelManager.classify_image's search (expert_bioclip/bioclip_model.py:313-316), batched:
 * top-k of queries . gallery^T over this GPU's gallery shard, score descending / index ascending.
 * queries fp16 [Q,D], gallery fp16 [N,D] (device, row-major, D % 64 == 0, 16-byte aligned), k <= 768.
 * out_scores float32 [Q,k], out_index int64 [Q,k] = index_base + local row; empty slots are (-inf, -1).
 * Scratch lives in the handle and grows on demand (cudaMalloc on the first call of a given size). */
int b200clip_topk(b200clip_handle* h, const void* queries_f16_dev, int32_t Q, const void* gallery_f16_dev, int64_t N,
                  int32_t D, int32_t k, int64_t index_base, float* out_scores_dev, int64_t* out_index_dev, void* stream);

/* K12 — merge `parts` candidate lists (scores float32 [parts][Q][k], index int64 [parts][Q][k], e.g. the result of an
 * NCCL all-gather of every GPU's b200clip_topk output) into the global top-k with the same total order. */
int b200clip_merge_topk(b200clip_handle* h, const float* part_scores_dev, const int64_t* part_index_dev, int32_t parts,
                        int32_t Q, int32_t k, float* out_scores_dev, int64_t* out_index_dev, void* stream);

/* The same two steps in the form the sharded search exchanges between GPUs (gallery row-sharded, one all-gather of the
 * per-GPU candidates, merge on every rank; the reference function being sharded is bioclip_model.py:313-316).  b200clip_topk_keys
 * returns this GPU's top-k as ONE 64-bit key per candidate, out_keys [Q,k]: order-preserving score bits in the high word, ~(row index
 * inside this shard) in the low word, 0 = empty slot - 8 bytes per candidate instead of 4 + 8, so a single all-gather (NCCL, MPI, or
 * the caller's own transport: no torch type is involved) moves them.  b200clip_merge_topk_keys takes the gathered keys
 * [parts][Q][k] and the index_base of every part (device array [parts]) and writes the global top-k like b200clip_merge_topk. */
int b200clip_topk_keys(b200clip_handle* h, const void* queries_f16_dev, int32_t Q, const void* gallery_f16_dev, int64_t N,
                       int32_t D, int32_t k, uint64_t* out_keys_dev, void* stream);
int b200clip_merge_topk_keys(b200clip_handle* h, const uint64_t* part_keys_dev, const int64_t* part_bases_dev, int32_t parts,
                             int32_t Q, int32_t k, float* out_scores_dev, int64_t* out_index_dev, void* stream);

/* Sharded gallery search in ONE call for a caller that brings an NCCL communicator - SURVEY §8(b)'s b200clip_merge_topk(..., ncclComm_t);
 * the reference has no counterpart (bioclip_model.py:313-316 searches one in-process matrix).  Every rank passes the same queries and
 * its own row shard (N_local may be 0) with the global index of the shard's first row; the call runs the fused top-k on the shard, moves
 * the [Q,k] 64-bit exchange keys and the index bases with one grouped ncclAllGather on `stream`, merges, and leaves the global top-k
 * (scores fp32 [Q,k], indices int64 [Q,k], -1 = empty slot) on EVERY rank.  `nccl_comm` is an ncclComm_t of `world` ranks; NCCL is
 * resolved at run time from the libnccl the process has loaded (the library does not link against it): B200CLIP_ESTATE if there is none. */
int b200clip_search_sharded(b200clip_handle* h, const void* queries_f16_dev, int32_t Q, const void* gallery_shard_f16_dev, int64_t N_local,
                            int32_t D, int32_t k, int64_t index_base, void* nccl_comm, int32_t world, float* out_scores_dev,
                            int64_t* out_index_dev, void* stream);

/* Search without any allocation.  No counterpart in the reference: it exists for callers that are not allowed to
 * allocate device memory at the moment they search — a host thread whose stream is being captured into a CUDA graph,
 * a real-time serving loop that must not hit the allocator's locks, or a plug-in living inside another framework's
 * memory pool.  Same kernels, arguments and results as b200clip_topk, but no handle: the scratch (candidate lists and
 * per-split results) is a fixed 32 MiB array compiled into the library, which every GPU that loads the library keeps
 * until exit.  It runs on the CURRENT device.  Limits: k <= 256 and Q*k small enough for the array (roughly 6200
 * queries at k = 100); larger requests fail with B200CLIP_EINVAL and a message naming the limit.  The first call on
 * each device resolves the array's device address (cudaGetSymbolAddress) — make that first call outside any stream
 * capture.  All calls share the one array, so two calls whose kernels overlap on the GPU (different streams or
 * threads) race; serialise them.  It calls neither cudaMalloc nor cudaFree nor any allocating library routine. */
int b200clip_topk_static(const void* queries_f16_dev, int32_t Q, const void* gallery_f16_dev, int64_t N, int32_t D,
                         int32_t k, int64_t index_base, float* out_scores_dev, int64_t* out_index_dev, void* stream);
/* The handle-less call above has no b200clip_sync to report a tripped device-side watchdog, so this does it (no counterpart in the
 * reference either): synchronises `stream`, reads the watchdog words at the head of the built-in array on the CURRENT device and
 * returns B200CLIP_EKERNEL (clearing them) if the last b200clip_topk_static there gave up on a pipeline barrier - its results
 * are then garbage.  Allocates nothing; 8 bytes device -> host. */
int b200clip_topk_static_status(void* stream);

/* Building blocks exported for the parity tests (same kernels the calls above use). */
int b200clip_gemm_bf16(b200clip_handle* h, const void* A_bf16_dev, const void* W_bf16_dev, int32_t M, int32_t N, int32_t K,
                       const float* bias_dev, void* out_dev, int32_t epilogue, int32_t activation, void* stream);
int b200clip_attention(b200clip_handle* h, const void* qkv_bf16_dev, void* out_bf16_dev, int32_t B, int32_t S,
                       int32_t width, int32_t heads, int32_t causal, void* stream);
int b200clip_layernorm(b200clip_handle* h, const float* x_dev, const float* gamma_dev, const float* beta_dev,
                       void* out_bf16_dev, int32_t rows, int32_t width, float eps, void* stream);

/* Introspection used by bench.py: kernels launched by this handle so far. */
int64_t b200clip_launch_count(b200clip_handle* h);
/* Optional timing: while enabled, CUDA events bracket every kernel launch of the embed path on its stream.
 * profile_read synchronises and returns summed milliseconds / launch counts for the 4 classes
 * {0 preprocess, 1 tcgen05 GEMM, 2 LayerNorm/assembly/normalise, 3 attention}, then clears the records. */
int b200clip_profile(b200clip_handle* h, int32_t enable);
int b200clip_profile_read(b200clip_handle* h, float* ms_by_class, int64_t* launches_by_class, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CLIP_H */
