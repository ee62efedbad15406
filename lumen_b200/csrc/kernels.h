// Launchers shared between the translation units of libb200clip (internal; the public ABI is include/b200clip.h).
#pragma once
#include "common.cuh"

namespace b200 {

struct JpegState;
JpegState* jpeg_state_create(int* rc);
void jpeg_state_destroy(JpegState*);
int jpeg_probe(const void* buf, size_t len, int32_t* hw, int32_t* ncomp);
int jpeg_decode_batch(JpegState*, const void* const* bufs, const size_t* lens, int n, void* dst, size_t dst_capacity, int64_t* offsets_out,
                      int32_t* hw_out, int32_t* status_out, int64_t* total_out, int threads, cudaStream_t st, int* launches);

struct PreprocessState;
PreprocessState* preprocess_state_create(const float* mean, const float* stdv, int* rc);
void preprocess_state_destroy(PreprocessState*);
int preprocess_run(PreprocessState*, const uint8_t* src_dev, const long long* offsets, const int* hw, int B, int S, int mode,
                   int layout, int out_bf16, int P, int kpad, void* out_dev, cudaStream_t st);
int gemm_bf16_tc(const void* A, int lda, const void* W, int ldw, int M, int N, int K, const float* bias, void* out, int ldc,
                 int epi, int act, DeviceStatus* status, cudaStream_t st, int reverse = 0);
int launch_layernorm(const float* x, long long row_stride, const float* g, const float* b, void* out, int rows, int d, float eps, cudaStream_t st,
                     int forward = 0);
int launch_embed_preln(const float* patch, const float* cls, const float* pos, const float* g, const float* b, float* xres,
                       int B, int S, int d, float eps, cudaStream_t st);
int launch_l2_normalize(const float* in, float* out, int rows, int D, cudaStream_t st);
// last layer: one query (the CLS row) per (image, head) against all keys / values of the image
int launch_cls_attention(const void* q_cls_bf16, const void* qkv_bf16, void* out_bf16, int B, int S, int d, int heads, cudaStream_t st);
int launch_text_embed(const int* ids, const float* tok, const float* pos, float* x, int rows, int ctx, int d, int vocab, cudaStream_t st);
int launch_text_eot_gather(const int* ids, const float* x, float* out, int B, int ctx, int d, cudaStream_t st);
int launch_gallery_pack(const float* in, void* out_f16, long long rows, int D, int normalize, cudaStream_t st);
// tcgen05 attention (attention_tc.cu), head_dim 64: attention_tc_kernel for S > 64 (and the causal text tower),
// S <= 64 on tcgen05 (attention_tc.cu): two (image, head) problems per 128-row tile, four small CTAs per SM
int launch_attention_short(const void* qkv, void* out, int B, int S, int d, int heads, DeviceStatus* status, cudaStream_t st, int reverse = 0);
int launch_attention_auto(const void* qkv, void* out, int B, int S, int d, int heads, int causal, DeviceStatus* status, cudaStream_t st, int reverse = 0);
int launch_attention_tc(const void* qkv, void* out, int B, int S, int d, int heads, int causal, DeviceStatus* status, cudaStream_t st);
struct TopkPlan { int splits, qtiles, cpl, halves, room; long long rows_per_split; size_t cand_bytes, part_bytes, qstage_bytes; };
int topk_plan(int Q, long long N, int k, TopkPlan* plan);
int topk_run(const void* queries, const void* gallery, int Q, long long N, int D, int k, long long index_base, void* scratch,
             const TopkPlan& plan, float* out_s, long long* out_i, DeviceStatus* status, cudaStream_t st,
             unsigned long long* out_keys = nullptr);   // out_keys: emit 64-bit exchange keys (shard-local indices) instead of (out_s, out_i)
int merge_topk_keys_run(const unsigned long long* pkeys, const long long* bases_dev, int P, int Q, int k, float* out_s, long long* out_i, cudaStream_t st);
int merge_topk_run(const float* ps, const long long* pi, int P, int Q, int k, float* out_s, long long* out_i, cudaStream_t st);
int classify_run(const float* img, const float* labels_unit, int B, int N, int D, float scale, int softmax, int k,
                 float* out_v, long long* out_i, cudaStream_t st);

enum : int { EPI_F32 = 0, EPI_BIAS_BF16 = 1, EPI_BIAS_ACT_BF16 = 2, EPI_BIAS_RES_F32 = 3 };


}  // namespace b200
