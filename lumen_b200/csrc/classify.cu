// K11 — zero-shot classify: cosine -> (optional) softmax(scale * s) -> top-k, one CTA per image.
//
// Replaces packages/lumen-clip/src/lumen_clip/general_clip/clip_model.py:274-315 (scale 100, softmax, top-k)
// and :341-357 (scene: scale 1, best entry).  All arithmetic is fp32 like the reference; the label matrix is
// row-normalised once when it is uploaded (the reference re-normalises it on every call, :289-291).
// Ordering: value descending, index ascending among equal values.
#include "common.cuh"
#include "kernels.h"

namespace b200 {

__global__ void __launch_bounds__(256) classify_kernel(const float* __restrict__ img, const float* __restrict__ labels,
                                                       int N, int D, float scale, int softmax, int k,
                                                       float* __restrict__ out_v, long long* __restrict__ out_i) {
    extern __shared__ __align__(16) float csm[];
    float* sc = csm;                    // [N] scores -> values
    float* qv = csm + N;                // [D] unit-norm query
    __shared__ float red_v[8];
    __shared__ int red_i[8];
    __shared__ float bcast;
    const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* q = img + (size_t)b * D;
    // ||q||
    float s = 0.f;
    for (int i = threadIdx.x; i < D; i += 256) { const float v = q[i]; s += v * v; }
    s = warp_sum(s);
    if (lane == 0) red_v[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) { float t = 0.f; for (int w = 0; w < 8; ++w) t += red_v[w]; bcast = sqrtf(t); }
    __syncthreads();
    const float nrm = bcast;
    for (int i = threadIdx.x; i < D; i += 256) qv[i] = q[i] / nrm;
    __syncthreads();
    // dot products: one warp per label row, lanes stride the row (coalesced)
    for (int n = warp; n < N; n += 8) {
        const float* t = labels + (size_t)n * D;
        float a = 0.f;
        for (int i = lane; i < D; i += 32) a = fmaf(qv[i], t[i], a);
        a = warp_sum(a);
        if (lane == 0) sc[n] = a;
    }
    __syncthreads();
    if (softmax) {
        float m = -INFINITY;
        for (int n = threadIdx.x; n < N; n += 256) m = fmaxf(m, sc[n] * scale);
        m = warp_max(m);
        if (lane == 0) red_v[warp] = m;
        __syncthreads();
        if (threadIdx.x == 0) { float t = red_v[0]; for (int w = 1; w < 8; ++w) t = fmaxf(t, red_v[w]); bcast = t; }
        __syncthreads();
        m = bcast;
        float e = 0.f;
        for (int n = threadIdx.x; n < N; n += 256) { const float v = expf(sc[n] * scale - m); sc[n] = v; e += v; }
        e = warp_sum(e);
        __syncthreads();
        if (lane == 0) red_v[warp] = e;
        __syncthreads();
        if (threadIdx.x == 0) { float t = 0.f; for (int w = 0; w < 8; ++w) t += red_v[w]; bcast = t; }
        __syncthreads();
        const float tot = bcast;
        for (int n = threadIdx.x; n < N; n += 256) sc[n] = sc[n] / tot;
        __syncthreads();
    }
    // k rounds of block arg-max (value descending, index ascending); winners are struck out with -inf
    for (int r = 0; r < k; ++r) {
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int n = threadIdx.x; n < N; n += 256) {
            const float v = sc[n];
            if (v > bv || (v == bv && n < bi)) { bv = v; bi = n; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { red_v[warp] = bv; red_i[warp] = bi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float v = red_v[0]; int i = red_i[0];
            for (int w = 1; w < 8; ++w) if (red_v[w] > v || (red_v[w] == v && red_i[w] < i)) { v = red_v[w]; i = red_i[w]; }
            const bool valid = i != 0x7fffffff && r < N;
            out_v[(size_t)b * k + r] = valid ? v : -INFINITY;
            out_i[(size_t)b * k + r] = valid ? i : -1;
            if (valid) sc[i] = -INFINITY;
        }
        __syncthreads();
    }
}

int classify_run(const float* img, const float* labels_unit, int B, int N, int D, float scale, int softmax, int k,
                 float* out_v, long long* out_i, cudaStream_t st) {
    B200_REQUIRE(B >= 0 && N > 0 && D > 0 && k > 0, "classify: bad sizes");
    if (B == 0) return B200_OK;
    const size_t smem = ((size_t)N + D) * 4;
    B200_REQUIRE(smem <= 200 * 1024, "classify: label set too large for one CTA (use the gallery search path)");
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(classify_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    classify_kernel<<<B, 256, smem, st>>>(img, labels_unit, N, D, scale, softmax, k, out_v, out_i);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

}  // namespace b200
