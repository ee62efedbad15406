// K3 LayerNorm, token assembly (CLS + position + pre-LN), K5 attention, E2 L2-normalise.
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
                                                             __nv_bfloat16* __restrict__ out, int rows, int d, float eps) {
    pdl_wait();
    pdl_trigger();
    // rows are walked from the end: the GEMM that just updated x finished with the last row blocks, so those are the
    // lines most likely still in L2
    const int row = rows - 1 - (blockIdx.x * 8 + (threadIdx.x >> 5)), lane = threadIdx.x & 31;
    if (row < 0) return;
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

// ------------------------------------------------------------------------------------------ attention
// softmax(Q K^T / sqrt(64)) V for one (image, head, 64-query tile), head_dim fixed at 64 (all CLIP ViTs here).
// Flash-style: keys in tiles of 64 with an online softmax, bf16 mma.sync m16n8k16 with fp32 accumulators.
// (Attention is 1-9 % of the tower's FLOPs; the tcgen05 budget goes to the GEMMs.)
constexpr int ATT_PITCH = 72;   // bf16 elements per shared-memory row (64 + 8 pad: conflict-free ldmatrix)

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, bool valid) {
    const int n = valid ? 16 : 0;       // src-size 0 -> the 16 destination bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(n) : "memory");
}
// 64 rows x 64 bf16 (128 B per row) -> 512 16-byte vectors over 128 threads; rows past the end are zero-filled.
__device__ __forceinline__ void load_tile64_async(__nv_bfloat16* dst, const __nv_bfloat16* src, int row0, int nrows_valid, int ld) {
#pragma unroll
    for (int i = threadIdx.x; i < 512; i += 128) {
        const int r = i >> 3, c = (i & 7) * 8;
        const bool ok = row0 + r < nrows_valid;
        cp_async16_zfill(dst + r * ATT_PITCH + c, src + (size_t)(ok ? row0 + r : 0) * ld + c, ok);
    }
}

// Persistent CTAs walk a flat list of (image, head, 64-query tile) items; every step consumes one 64-key tile of
// K and V.  The next step's tiles (and the next item's Q tile) are fetched with cp.async into the other buffer
// while the current step computes, so global-load latency overlaps the MMAs and softmax.
constexpr int ATT_TILE = 64 * ATT_PITCH;          // bf16 elements per staged tile
constexpr int ATT_SMEM = 6 * ATT_TILE * 2;        // Q, K, V double buffered

__global__ void __launch_bounds__(128, 4) attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                                                           int S, int d, int heads, int n_qt, int total_items) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(16) __nv_bfloat16 att_sm[];
    __nv_bfloat16* sQ = att_sm;                    // [2][ATT_TILE]
    __nv_bfloat16* sK = att_sm + 2 * ATT_TILE;
    __nv_bfloat16* sV = att_sm + 4 * ATT_TILE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int ld = 3 * d;
    const int n_kt = (S + 63) >> 6;
    const float sl2 = 0.125f * 1.4426950408889634f;   // 64^-0.5 * log2(e)

    auto item_base = [&](int item, int& qt) -> const __nv_bfloat16* {
        qt = item % n_qt;
        const int bh = item / n_qt, h = bh % heads, b = bh / heads;
        return qkv + (size_t)b * S * ld + h * 64;
    };
    auto issue = [&](int item, int kt, int buf, int qbuf) {
        int qt;
        const __nv_bfloat16* base = item_base(item, qt);
        if (kt == 0) load_tile64_async(sQ + qbuf * ATT_TILE, base, qt * 64, S, ld);
        load_tile64_async(sK + buf * ATT_TILE, base + d, kt * 64, S, ld);
        load_tile64_async(sV + buf * ATT_TILE, base + 2 * d, kt * 64, S, ld);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    int item = blockIdx.x, kt = 0, step = 0, local_item = 0;
    if (item >= total_items) return;
    issue(item, 0, 0, 0);
    uint32_t qa[4][4];
    float o[8][4];
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    while (item < total_items) {
        // next step: same item / next key tile, or first tile of this CTA's next item
        int n_item = item, n_kt_idx = kt + 1, n_local = local_item;
        if (n_kt_idx == n_kt) { n_item = item + gridDim.x; n_kt_idx = 0; n_local = local_item + 1; }
        if (n_item < total_items) {
            issue(n_item, n_kt_idx, (step + 1) & 1, n_local & 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const __nv_bfloat16* cK = sK + (step & 1) * ATT_TILE;
        const __nv_bfloat16* cV = sV + (step & 1) * ATT_TILE;
        if (kt == 0) {
            const __nv_bfloat16* cQ = sQ + (local_item & 1) * ATT_TILE;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                ldsm_x4(qa[ks], cQ + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * ATT_PITCH + ks * 16 + (lane >> 4) * 8);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
            m0 = m1 = -INFINITY; l0 = l1 = 0.f;
        }
        const int k0 = kt * 64;
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                uint32_t kb[4];
                ldsm_x4(kb, cK + (np * 16 + (lane & 7) + (lane >> 4) * 8) * ATT_PITCH + ks * 16 + ((lane >> 3) & 1) * 8);
                mma_bf16(s[2 * np], qa[ks], kb[0], kb[1]);
                mma_bf16(s[2 * np + 1], qa[ks], kb[2], kb[3]);
            }
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int key = k0 + nt * 8 + 2 * t;
            if (key >= S) { s[nt][0] = -INFINITY; s[nt][2] = -INFINITY; }
            if (key + 1 >= S) { s[nt][1] = -INFINITY; s[nt][3] = -INFINITY; }
            mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float c0 = exp2f((m0 - mx0) * sl2), c1 = exp2f((m1 - mx1) * sl2);   // exp2f(-inf) = 0 on the first tile
        m0 = mx0; m1 = mx1;
        l0 *= c0; l1 *= c1;
#pragma unroll
        for (int i = 0; i < 8; ++i) { o[i][0] *= c0; o[i][1] *= c0; o[i][2] *= c1; o[i][3] *= c1; }
        uint32_t pa[4][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float p0 = exp2f((s[nt][0] - m0) * sl2), p1 = exp2f((s[nt][1] - m0) * sl2);
            const float p2 = exp2f((s[nt][2] - m1) * sl2), p3 = exp2f((s[nt][3] - m1) * sl2);
            l0 += p0 + p1; l1 += p2 + p3;
            pa[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
            pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
            for (int dp = 0; dp < 4; ++dp) {
                uint32_t vb[4];
                ldsm_x4_t(vb, cV + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * ATT_PITCH + dp * 16 + (lane >> 4) * 8);
                mma_bf16(o[2 * dp], pa[kk], vb[0], vb[1]);
                mma_bf16(o[2 * dp + 1], pa[kk], vb[2], vb[3]);
            }
        if (kt == n_kt - 1) {
            float t0 = l0, t1 = l1;
            t0 += __shfl_xor_sync(0xffffffffu, t0, 1); t0 += __shfl_xor_sync(0xffffffffu, t0, 2);
            t1 += __shfl_xor_sync(0xffffffffu, t1, 1); t1 += __shfl_xor_sync(0xffffffffu, t1, 2);
            const float i0 = 1.f / t0, i1 = 1.f / t1;
            const int qt = item % n_qt, bh = item / n_qt, h = bh % heads, b = bh / heads;
            const int r0 = qt * 64 + warp * 16 + g, r1 = r0 + 8;
            __nv_bfloat16* ob = out + (size_t)b * S * d + h * 64;
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) {
                if (r0 < S) *reinterpret_cast<uint32_t*>(ob + (size_t)r0 * d + dt * 8 + 2 * t) = pack_bf16(o[dt][0] * i0, o[dt][1] * i0);
                if (r1 < S) *reinterpret_cast<uint32_t*>(ob + (size_t)r1 * d + dt * 8 + 2 * t) = pack_bf16(o[dt][2] * i1, o[dt][3] * i1);
            }
        }
        __syncthreads();                               // this step's K/V (and Q) buffers may now be overwritten
        item = n_item; kt = n_kt_idx; local_item = n_local; ++step;
    }
}

// ------------------------------------------------------------------------------------------ launchers
int launch_layernorm(const float* x, long long row_stride, const float* g, const float* b, void* out, int rows, int d, float eps, cudaStream_t st) {
    B200_REQUIRE(d % 128 == 0 && d <= 128 * LN_MAX_V4, "layernorm: width must be a multiple of 128, at most 1024");
    B200_CUDA_OK(launch_pdl(layernorm_bf16_kernel, dim3((rows + 7) / 8), dim3(256), 0, st, x, row_stride, g, b, reinterpret_cast<__nv_bfloat16*>(out), rows, d, eps));
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
int launch_attention(const void* qkv, void* out, int B, int S, int d, int heads, cudaStream_t st) {
    B200_REQUIRE(d == heads * 64, "attention: head_dim must be 64");
    B200_REQUIRE(B > 0 && S > 0, "attention: empty batch");
    const int n_qt = (S + 63) / 64;
    const long long items = (long long)B * heads * n_qt;
    B200_REQUIRE(items < (1LL << 30), "attention: too many (image, head, tile) items");
    static bool attr_set[64] = {false};
    int dev = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev]) {
        B200_CUDA_OK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM));
        if (dev < 64) attr_set[dev] = true;
    }
    const int grid = (int)(items < 4LL * kNumSMs ? items : 4LL * kNumSMs);
    B200_CUDA_OK(launch_pdl(attention_kernel, dim3(grid), dim3(128), ATT_SMEM, st, reinterpret_cast<const __nv_bfloat16*>(qkv),
                            reinterpret_cast<__nv_bfloat16*>(out), S, d, heads, n_qt, (int)items));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

}  // namespace b200
