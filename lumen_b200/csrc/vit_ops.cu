// K3 LayerNorm, token assembly (CLS + position + pre-LN), E2 L2-normalise, text-tower embedding / pooling, gallery packing.
// (K5 attention lives in attention_tc.cu.)
// The math follows what the reference's encoder call computes (onnxrt_backend.py:485 / torch_backend.py:652-666;
// restated in oracle/clip_oracle.py:vit_encode).  The residual stream is fp32, GEMM operands are bf16.
#include "common.cuh"
#include "kernels.h"

#include <climits>

namespace b200 {

// ------------------------------------------------------------------------------------------ LayerNorm
// One warp per row; the row lives in registers (d <= 1024, d % 128 == 0), two-pass variance in fp32.
constexpr int LN_MAX_V4 = 8;

__device__ __forceinline__ void ln_row(float4 (&v)[LN_MAX_V4], int nv, int d, float eps, float& mean, float& rstd) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAX_V4; ++i) if (i < nv) s += v[i].x + v[i].y + v[i].z + v[i].w;
    mean = warp_sum(s) / d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAX_V4; ++i) if (i < nv) {
        float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, e = v[i].w - mean;
        q += a * a + b * b + c * c + e * e;
    }
    rstd = rsqrtf(warp_sum(q) / d + eps);
}

// x: fp32 rows (row r at x + r*row_stride*d), out: bf16 [rows, d].
__global__ void __launch_bounds__(256) layernorm_bf16_kernel(const float* __restrict__ x, long long row_stride,
                                                             const float* __restrict__ gamma, const float* __restrict__ beta,
                                                             __nv_bfloat16* __restrict__ out, int rows, int d, float eps, int forward) {
    pdl_wait();
    pdl_trigger();
    // rows are walked from the end: the GEMM that just updated x finished with the last row blocks, so those are the
    // lines most likely still in L2 (forward != 0: that GEMM itself ran from the end)
    const int r0 = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int row = forward ? r0 : rows - 1 - r0, lane = threadIdx.x & 31;
    if (row < 0 || row >= rows) return;
    const int nv = d >> 7;
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * row_stride * d);
    float4 v[LN_MAX_V4];
#pragma unroll
    for (int i = 0; i < LN_MAX_V4; ++i) if (i < nv) v[i] = xr[i * 32 + lane];
    float mean, rstd;
    ln_row(v, nv, d, eps, mean, rstd);
    uint2* o = reinterpret_cast<uint2*>(out + (size_t)row * d);
#pragma unroll
    for (int i = 0; i < LN_MAX_V4; ++i) if (i < nv) {
        const float4 g = reinterpret_cast<const float4*>(gamma)[i * 32 + lane];
        const float4 b = reinterpret_cast<const float4*>(beta)[i * 32 + lane];
        __nv_bfloat162 lo = __floats2bfloat162_rn((v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y);
        __nv_bfloat162 hi = __floats2bfloat162_rn((v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w);
        uint2 pk; pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
        o[i * 32 + lane] = pk;
    }
}

// Token assembly: row (b, s) = LN_pre( (s == 0 ? cls : patch[b*n + s-1]) + pos[s] ) -> fp32 residual stream.
__global__ void __launch_bounds__(256) embed_preln_kernel(const float* __restrict__ patch, const float* __restrict__ cls,
                                                          const float* __restrict__ pos, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float* __restrict__ xres,
                                                          int B, int S, int d, float eps) {
    pdl_wait();
    pdl_trigger();
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= B * S) return;
    const int b = row / S, s = row - b * S;
    const int nv = d >> 7;
    const float4* src = reinterpret_cast<const float4*>(s == 0 ? cls : patch + ((size_t)b * (S - 1) + (s - 1)) * d);
    const float4* pr = reinterpret_cast<const float4*>(pos + (size_t)s * d);
    float4 v[LN_MAX_V4];
#pragma unroll
    for (int i = 0; i < LN_MAX_V4; ++i) if (i < nv) {
        const float4 a = src[i * 32 + lane], p = pr[i * 32 + lane];
        v[i] = make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
    }
    float mean, rstd;
    ln_row(v, nv, d, eps, mean, rstd);
    float4* o = reinterpret_cast<float4*>(xres + (size_t)row * d);
#pragma unroll
    for (int i = 0; i < LN_MAX_V4; ++i) if (i < nv) {
        const float4 g = reinterpret_cast<const float4*>(gamma)[i * 32 + lane];
        const float4 be = reinterpret_cast<const float4*>(beta)[i * 32 + lane];
        o[i * 32 + lane] = make_float4((v[i].x - mean) * rstd * g.x + be.x, (v[i].y - mean) * rstd * g.y + be.y,
                                       (v[i].z - mean) * rstd * g.z + be.z, (v[i].w - mean) * rstd * g.w + be.w);
    }
}

// v / ||v||_2 per row, no epsilon (onnxrt_backend.py:545-546).  One warp per row.
__global__ void __launch_bounds__(256) l2_normalize_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int D) {
    pdl_wait();
    pdl_trigger();
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* r = in + (size_t)row * D;
    float s = 0.f;
    for (int i = lane; i < D; i += 32) s += r[i] * r[i];
    const float nrm = sqrtf(warp_sum(s));
    for (int i = lane; i < D; i += 32) out[(size_t)row * D + i] = r[i] / nrm;
}

// N4 text tower (modeling_clip.py:CLIPTextEmbeddings): x[b, s] = token_embedding[ids[b, s]] + position_embedding[s], fp32 residual stream.
// One warp per token row; ids outside the vocabulary are clamped (the host wrapper validates them).
__global__ void __launch_bounds__(256) text_embed_kernel(const int* __restrict__ ids, const float* __restrict__ tok, const float* __restrict__ pos,
                                                         float* __restrict__ x, int rows, int ctx, int d, int vocab) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const int id = min(max(ids[row], 0), vocab - 1);
    const float4* t = reinterpret_cast<const float4*>(tok + (size_t)id * d);
    const float4* p = reinterpret_cast<const float4*>(pos + (size_t)(row % ctx) * d);
    float4* o = reinterpret_cast<float4*>(x + (size_t)row * d);
    for (int i = lane; i < d / 4; i += 32) {
        const float4 a = __ldg(t + i), b = __ldg(p + i);
        o[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    }
}
// Pooling of the text tower: the hidden state at the end-of-text token, i.e. at argmax(ids) (first maximum, as torch.argmax;
// the EOT id is the largest id of the CLIP vocabulary) -> compact [B, d] rows.  One warp per sequence.
__global__ void __launch_bounds__(256) text_eot_gather_kernel(const int* __restrict__ ids, const float* __restrict__ x, float* __restrict__ out,
                                                              int B, int ctx, int d) {
    const int b = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (b >= B) return;
    int best = INT_MIN, at = 0;
    for (int s = lane; s < ctx; s += 32) { const int v = ids[b * ctx + s]; if (v > best) { best = v; at = s; } }
    for (int o = 16; o; o >>= 1) {
        const int ob = __shfl_xor_sync(0xffffffffu, best, o), oa = __shfl_xor_sync(0xffffffffu, at, o);
        if (ob > best || (ob == best && oa < at)) { best = ob; at = oa; }
    }
    const float4* src = reinterpret_cast<const float4*>(x + ((size_t)b * ctx + at) * d);
    float4* dst = reinterpret_cast<float4*>(out + (size_t)b * d);
    for (int i = lane; i < d / 4; i += 32) dst[i] = src[i];
}

// N4: fp32 embedding rows -> fp16 gallery rows (what K10 reads), optionally divided by their L2 norm first
// (bioclip_model.py:251-256 normalises label matrices that are not unit length; resources/loader.py:366 loads them as fp32 .npy).
// One warp per row; HBM-bound: 4 B in + 2 B out per element.
__global__ void __launch_bounds__(256) gallery_pack_kernel(const float* __restrict__ in, __half* __restrict__ out, long long rows, int D, int normalize) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* r = in + row * D;
    __half* o = out + row * D;
    float inv = 1.f;
    if (normalize) {
        float s = 0.f;
        for (int i = lane; i < D; i += 32) s += r[i] * r[i];
        inv = sqrtf(warp_sum(s));
    }
    if ((D & 3) == 0 && ((reinterpret_cast<uintptr_t>(r) | reinterpret_cast<uintptr_t>(o)) & 15) == 0 && (D & 7) == 0) {
        for (int i = lane * 8; i < D; i += 256) {
            const float4 a = *reinterpret_cast<const float4*>(r + i), b = *reinterpret_cast<const float4*>(r + i + 4);
            __half2 h0, h1, h2, h3;
            if (normalize) {
                h0 = __floats2half2_rn(a.x / inv, a.y / inv); h1 = __floats2half2_rn(a.z / inv, a.w / inv);
                h2 = __floats2half2_rn(b.x / inv, b.y / inv); h3 = __floats2half2_rn(b.z / inv, b.w / inv);
            } else {
                h0 = __floats2half2_rn(a.x, a.y); h1 = __floats2half2_rn(a.z, a.w); h2 = __floats2half2_rn(b.x, b.y); h3 = __floats2half2_rn(b.z, b.w);
            }
            uint4 w;
            w.x = *reinterpret_cast<uint32_t*>(&h0); w.y = *reinterpret_cast<uint32_t*>(&h1);
            w.z = *reinterpret_cast<uint32_t*>(&h2); w.w = *reinterpret_cast<uint32_t*>(&h3);
            *reinterpret_cast<uint4*>(o + i) = w;
        }
    } else {
        for (int i = lane; i < D; i += 32) o[i] = __float2half_rn(normalize ? r[i] / inv : r[i]);
    }
}

// ------------------------------------------------------------------------------------------ last layer: CLS query only
// The tower's output is the projection of the CLS token's final hidden state (modeling_clip.py:684-694: pooled_output =
// last_hidden_state[:, 0]), so in the LAST transformer layer only the CLS row of every image is ever read again: its attention
// needs the keys and values of all tokens but a single query per (image, head).  One warp per (image, head):
//   scores   eight lanes per key (16 bytes = 8 dims each), four keys per step, partial dots reduced inside the lane group;
//   softmax  over the warp's score array in shared memory, fp32, max-subtracted like the full kernels;
//   output   lanes over the 64 head dims (2 each), one 128-byte V row per key.
// q: bf16 [B, d] (CLS rows only), qkv: bf16 [B*S, 3d] of which only the K and V columns are read, out: bf16 [B, d].
__global__ void __launch_bounds__(256) cls_attention_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ qkv,
                                                            __nv_bfloat16* __restrict__ out, int B, int S, int d, int heads) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ float cls_sc[];                       // [8 warps][S]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long item = (long long)blockIdx.x * 8 + warp;
    if (item >= (long long)B * heads) return;
    const int b = (int)(item / heads), h = (int)(item % heads);
    float* sc = cls_sc + (size_t)warp * S;
    const int g = lane >> 3, c = lane & 7;
    float qv[8];
    {
        const uint4 w = *reinterpret_cast<const uint4*>(q + (size_t)b * d + h * 64 + c * 8);
        const __nv_bfloat162* w2 = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(w2[i]); qv[2 * i] = f.x; qv[2 * i + 1] = f.y; }
    }
    const __nv_bfloat16* kbase = qkv + (size_t)b * S * 3 * d + d + h * 64;
    const __nv_bfloat16* vbase = kbase + d;
    float mx = -INFINITY;
    for (int jb = 0; jb < S; jb += 4) {
        const int j = jb + g;
        float part = 0.f;
        if (j < S) {
            const uint4 w = *reinterpret_cast<const uint4*>(kbase + (size_t)j * 3 * d + c * 8);
            const __nv_bfloat162* w2 = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(w2[i]); part = fmaf(qv[2 * i], f.x, part); part = fmaf(qv[2 * i + 1], f.y, part); }
        }
        part += __shfl_xor_sync(0xffffffffu, part, 1);
        part += __shfl_xor_sync(0xffffffffu, part, 2);
        part += __shfl_xor_sync(0xffffffffu, part, 4);
        part *= 0.125f;
        if (j < S) { if (c == 0) sc[j] = part; mx = fmaxf(mx, part); }
    }
    mx = warp_max(mx);
    __syncwarp();
    float sum = 0.f;
    for (int j = lane; j < S; j += 32) { const float e = __expf(sc[j] - mx); sc[j] = e; sum += e; }
    sum = warp_sum(sum);
    __syncwarp();
    float o0 = 0.f, o1 = 0.f;
    for (int j = 0; j < S; ++j) {
        const float pj = sc[j];
        const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(vbase + (size_t)j * 3 * d + 2 * lane));
        o0 = fmaf(pj, v.x, o0); o1 = fmaf(pj, v.y, o1);
    }
    const float inv = 1.f / sum;
    *reinterpret_cast<__nv_bfloat162*>(out + (size_t)b * d + h * 64 + 2 * lane) = __floats2bfloat162_rn(o0 * inv, o1 * inv);
}

// ------------------------------------------------------------------------------------------ launchers
int launch_cls_attention(const void* q, const void* qkv, void* out, int B, int S, int d, int heads, cudaStream_t st) {
    B200_REQUIRE(d == heads * 64 && B > 0 && S > 0, "cls_attention: head_dim must be 64");
    const size_t smem = (size_t)8 * S * sizeof(float);
    B200_REQUIRE(smem <= 48 * 1024, "cls_attention: sequence too long");
    const long long items = (long long)B * heads;
    B200_CUDA_OK(launch_pdl(cls_attention_kernel, dim3((unsigned)((items + 7) / 8)), dim3(256), smem, st, reinterpret_cast<const __nv_bfloat16*>(q),
                            reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(out), B, S, d, heads));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int launch_layernorm(const float* x, long long row_stride, const float* g, const float* b, void* out, int rows, int d, float eps, cudaStream_t st,
                     int forward) {
    B200_REQUIRE(d % 128 == 0 && d <= 128 * LN_MAX_V4, "layernorm: width must be a multiple of 128, at most 1024");
    B200_CUDA_OK(launch_pdl(layernorm_bf16_kernel, dim3((rows + 7) / 8), dim3(256), 0, st, x, row_stride, g, b, reinterpret_cast<__nv_bfloat16*>(out), rows, d, eps, forward));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int launch_embed_preln(const float* patch, const float* cls, const float* pos, const float* g, const float* b, float* xres,
                       int B, int S, int d, float eps, cudaStream_t st) {
    B200_REQUIRE(d % 128 == 0 && d <= 128 * LN_MAX_V4, "embed: width must be a multiple of 128, at most 1024");
    B200_CUDA_OK(launch_pdl(embed_preln_kernel, dim3((B * S + 7) / 8), dim3(256), 0, st, patch, cls, pos, g, b, xres, B, S, d, eps));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int launch_text_embed(const int* ids, const float* tok, const float* pos, float* x, int rows, int ctx, int d, int vocab, cudaStream_t st) {
    if (rows <= 0) return B200_OK;
    text_embed_kernel<<<(rows + 7) / 8, 256, 0, st>>>(ids, tok, pos, x, rows, ctx, d, vocab);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int launch_text_eot_gather(const int* ids, const float* x, float* out, int B, int ctx, int d, cudaStream_t st) {
    if (B <= 0) return B200_OK;
    text_eot_gather_kernel<<<(B + 7) / 8, 256, 0, st>>>(ids, x, out, B, ctx, d);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int launch_gallery_pack(const float* in, void* out_f16, long long rows, int D, int normalize, cudaStream_t st) {
    if (rows <= 0) return B200_OK;
    gallery_pack_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(in, static_cast<__half*>(out_f16), rows, D, normalize);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int launch_l2_normalize(const float* in, float* out, int rows, int D, cudaStream_t st) {
    if (rows == 0) return B200_OK;
    B200_CUDA_OK(launch_pdl(l2_normalize_kernel, dim3((rows + 7) / 8), dim3(256), 0, st, in, out, rows, D));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
}  // namespace b200
