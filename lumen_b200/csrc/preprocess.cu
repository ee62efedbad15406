// K1 — fused uint8-RGB resize (Pillow-exact antialiased bicubic) + crop + mean/std normalise.
//
// Replaces, bit for bit on the integer part, the reference closure
//   packages/lumen-clip/src/lumen_clip/backends/onnxrt_backend.py:410-431   (mode 0: squash to SxS)
// and the open_clip / torchvision transform built at backends/torch_backend.py:201-204
//   (mode 1: shortest side -> S, centre crop).
//
// Pillow's resize is two separable passes over 22-bit fixed-point coefficients with the horizontal result
// rounded and clipped to uint8 before the vertical pass, so a single fused 2-D float filter is NOT
// equivalent (it misses by up to 22/255).  One CTA produces TR output rows of one image:
//   1. the input rows it needs are one contiguous byte span of the packed HWC image -> staged into shared
//      memory with 16-byte cp.async loads (double buffered),
//   2. horizontal pass (int32 MAC, clip) -> uint8 intermediate rows in shared memory,
//   3. vertical pass from shared memory, lookup of (v/255 - mean)/std (exact fp32, 3x256 table),
//   4. coalesced vector stores, either NCHW or directly in the patch-major [B*n, 3*P*P (+pad)] layout
//      that the patch-embed GEMM consumes through TMA (no im2col pass).
// HBM traffic is the algorithmic minimum: every source byte once (plus the row overlap between
// neighbouring tiles, which hits L2), every output element once.
#include "common.cuh"
#include "kernels.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <vector>

namespace b200 {

struct ImgDesc {
    long long src_off;                 // byte offset of this image in the packed uint8 buffer (16-byte aligned)
    int h_in, w_in;
    int top, left;                     // crop offset inside the resized image
    int ksize_h, ksize_v;
    const int* bounds_h; const int* kk_h;   // per output x of the resized width: (xmin, count), coefficients
    const int* bounds_v; const int* kk_v;
    int identity;                      // 0 = resize in preprocess_kernel, 1 = resized size == source size (pure crop + convert in
                                       // convert_kernel), 2 = too wide / too tall for a shared-memory tile (the two global-memory kernels),
                                       // 3 = more than 100 times taller than wide and shrinking vertically: vertical resize first, as PIL does
    int pad_[3];                       // sizeof == 80: pulled to the device as five 16-byte words
};

struct PreParams {
    const uint8_t* src;
    const ImgDesc* desc;
    const float* lut;                  // [3][256]
    void* out;
    int S, TR;
    int layout;                        // 0 = NCHW, 1 = patch-major
    int out_bf16;
    int P, npatch_side, kpad;          // patch-major geometry
    int inter_pitch;                   // bytes per intermediate row (S*3 rounded up to 16)
    int stage_bytes;                   // bytes per staging buffer
};

constexpr int PRECISION_BITS = 22;

__device__ __forceinline__ int clip8(int v) { v >>= PRECISION_BITS; return min(max(v, 0), 255); }

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Copies bytes [start, start+len) of `src` into `dst` at offset (start & 15), using aligned 16-byte loads.
__device__ __forceinline__ int stage_span(uint8_t* dst, const uint8_t* src, long long start, int len) {
    const long long a0 = start & ~15LL;
    const int prefix = (int)(start - a0);
    const int nvec = (prefix + len + 15) >> 4;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) cp_async16(dst + i * 16, src + a0 + (long long)i * 16);
    return prefix;
}

// Element index of (image b, channel 0, row y, column x) and the distance between channels, for both layouts.
__device__ __forceinline__ size_t out_index(const PreParams& p, int b, int y, int x, int& cstride) {
    if (p.layout == 0) { cstride = p.S * p.S; return (((size_t)b * 3) * p.S + y) * p.S + x; }
    cstride = p.P * p.P;
    return ((size_t)b * p.npatch_side * p.npatch_side + (size_t)(y / p.P) * p.npatch_side + x / p.P) * p.kpad + (y % p.P) * p.P + (x % p.P);
}

// Stores 8 consecutive x positions of one channel at element index `idx` (vec8: one 16-byte bf16 / two 16-byte
// fp32 stores).  When the patch size is not a multiple of 8 (e.g. 14) the 8 pixels may straddle a patch boundary:
// element pairs are then addressed one by one.
template <bool BF16>
__device__ __forceinline__ void store8(const PreParams& p, bool vec8, size_t idx, int b, int c, int y, int x, const float (&f)[8]) {
    if (BF16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out);
        if (vec8) {
            uint4 v;
            __nv_bfloat162 t0 = __floats2bfloat162_rn(f[0], f[1]), t1 = __floats2bfloat162_rn(f[2], f[3]);
            __nv_bfloat162 t2 = __floats2bfloat162_rn(f[4], f[5]), t3 = __floats2bfloat162_rn(f[6], f[7]);
            v.x = *reinterpret_cast<uint32_t*>(&t0); v.y = *reinterpret_cast<uint32_t*>(&t1);
            v.z = *reinterpret_cast<uint32_t*>(&t2); v.w = *reinterpret_cast<uint32_t*>(&t3);
            *reinterpret_cast<uint4*>(o + idx) = v;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                int cs;
                size_t e = out_index(p, b, y, x + i, cs);
                e += (size_t)c * cs;
                *reinterpret_cast<__nv_bfloat162*>(o + e) = __floats2bfloat162_rn(f[i], f[i + 1]);
            }
        }
    } else {
        float* o = reinterpret_cast<float*>(p.out);
        if (vec8) {
            *reinterpret_cast<float4*>(o + idx) = make_float4(f[0], f[1], f[2], f[3]);
            *reinterpret_cast<float4*>(o + idx + 4) = make_float4(f[4], f[5], f[6], f[7]);
        } else {
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                int cs;
                size_t e = out_index(p, b, y, x + i, cs);
                e += (size_t)c * cs;
                *reinterpret_cast<float2*>(o + e) = make_float2(f[i], f[i + 1]);
            }
        }
    }
}

__device__ __forceinline__ uint32_t byte_of(uint32_t w, int i) { return __byte_perm(w, 0, 0x4440 | i); }
// four unsigned bytes of a times four SIGNED bytes of b, accumulated into c
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// NW+1 aligned 32-bit words starting at the word that holds `p` -> NW words whose byte 0 is *p (little endian).
template <int NW>
__device__ __forceinline__ void load_bytes_aligned(const uint8_t* p, uint32_t (&al)[NW]) {
    const uint32_t sh = (smem_u32(p) & 3u) * 8u;
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(p - (smem_u32(p) & 3u));
    uint32_t prev = wp[0];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        const uint32_t next = wp[j + 1];
        al[j] = __funnelshift_r(prev, next, sh);
        prev = next;
    }
}

// Horizontal pass for one staged chunk of source rows.  A thread owns one output column: its (at most KMAX)
// coefficients sit in registers for all rows of the chunk; the 3*KMAX source bytes of a row are fetched as 32-bit
// words (one shared-memory load per four bytes), realigned with funnel shifts and split with PRMT, so the integer
// MACs, not the load/store unit, set the pace.  Coefficients beyond the real tap count are zero, so the extra
// bytes (still inside the staging allocation) contribute nothing.  KMAX == 0: generic byte loop for very wide filters.
template <int KMAX>
__device__ __forceinline__ void hpass_chunk(const ImgDesc& d, const uint8_t* __restrict__ rows, int rowbytes, int nr, int S,
                                            uint8_t* __restrict__ inter_rows, int inter_pitch) {
    for (int x = threadIdx.x; x < S; x += blockDim.x) {
        const int xo = d.left + x;
        const int xmin = __ldg(d.bounds_h + 2 * xo), cnt = __ldg(d.bounds_h + 2 * xo + 1);
        const int* k = d.kk_h + (size_t)xo * d.ksize_h;
        const uint8_t* px0 = rows + xmin * 3;
        uint8_t* o = inter_rows + x * 3;
        if (KMAX > 0) {
            // Exact 22-bit coefficients on the byte dot-product unit: w = l0 + 256 l1 + 65536 l2 with l0, l1 the low two bytes
            // (unsigned) and l2 = w >> 16 (signed, |l2| < 128 for normalised cubic weights), four taps per register and limb.
            // Per four taps and channel: two PRMT gather the channel's four bytes out of three source words, three DP4A
            // accumulate the limbs - 15 instructions per 12 multiply-adds where IMAD + PRMT per byte took 24.  Still integer
            // arithmetic end to end, so the bit-exactness tests against Pillow keep guarding it.
            constexpr int G = KMAX > 0 ? KMAX / 4 : 1;    // tap groups (KMAX is a multiple of 4)
            constexpr int NW = 3 * G;                     // source words per row: 12 bytes per tap group
            uint32_t L0[G], L1[G], L2[G];
#pragma unroll
            for (int g = 0; g < G; ++g) {
                L0[g] = L1[g] = L2[g] = 0u;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int t = 4 * g + i;
                    const int wt = t < cnt ? __ldg(k + t) : 0;
                    L0[g] |= ((uint32_t)wt & 255u) << (8 * i);
                    L1[g] |= (((uint32_t)wt >> 8) & 255u) << (8 * i);
                    L2[g] |= (((uint32_t)(wt >> 16)) & 255u) << (8 * i);
                }
            }
            for (int r = 0; r < nr; ++r) {
                uint32_t al[NW];
                load_bytes_aligned<NW>(px0 + r * rowbytes, al);
                uint32_t lo0 = 0, lo1 = 0, lo2 = 0, mi0 = 0, mi1 = 0, mi2 = 0;
                int hi0 = 0, hi1 = 0, hi2 = 0;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    const uint32_t w0 = al[3 * g], w1 = al[3 * g + 1], w2 = al[3 * g + 2];
                    const uint32_t r4 = __byte_perm(__byte_perm(w0, w1, 0x0630), w2, 0x5210);   // bytes 0, 3, 6, 9 of the 12
                    const uint32_t g4 = __byte_perm(__byte_perm(w0, w1, 0x0741), w2, 0x6210);   // bytes 1, 4, 7, 10
                    const uint32_t b4 = __byte_perm(__byte_perm(w0, w1, 0x0052), w2, 0x7410);   // bytes 2, 5, 8, 11
                    lo0 = __dp4a(r4, L0[g], lo0); mi0 = __dp4a(r4, L1[g], mi0); hi0 = dp4a_us(r4, L2[g], hi0);
                    lo1 = __dp4a(g4, L0[g], lo1); mi1 = __dp4a(g4, L1[g], mi1); hi1 = dp4a_us(g4, L2[g], hi1);
                    lo2 = __dp4a(b4, L0[g], lo2); mi2 = __dp4a(b4, L1[g], mi2); hi2 = dp4a_us(b4, L2[g], hi2);
                }
                const int a0 = (1 << (PRECISION_BITS - 1)) + (int)lo0 + (int)(mi0 << 8) + (hi0 << 16);
                const int a1 = (1 << (PRECISION_BITS - 1)) + (int)lo1 + (int)(mi1 << 8) + (hi1 << 16);
                const int a2 = (1 << (PRECISION_BITS - 1)) + (int)lo2 + (int)(mi2 << 8) + (hi2 << 16);
                o[0] = (uint8_t)clip8(a0); o[1] = (uint8_t)clip8(a1); o[2] = (uint8_t)clip8(a2);
                o += inter_pitch;
            }
        } else {
            for (int r = 0; r < nr; ++r) {
                const uint8_t* px = px0 + r * rowbytes;
                int a0 = 1 << (PRECISION_BITS - 1), a1 = a0, a2 = a0;
                for (int j = 0; j < cnt; ++j) {
                    const int w = __ldg(k + j);
                    a0 += px[3 * j] * w; a1 += px[3 * j + 1] * w; a2 += px[3 * j + 2] * w;
                }
                o[0] = (uint8_t)clip8(a0); o[1] = (uint8_t)clip8(a1); o[2] = (uint8_t)clip8(a2);
                o += inter_pitch;
            }
        }
    }
}

template <bool BF16>
__global__ void __launch_bounds__(256) preprocess_kernel(PreParams p) {
    extern __shared__ __align__(16) uint8_t sm[];
    float* lut = reinterpret_cast<float*>(sm);                  // 768 floats
    uint8_t* stage = sm + 3072;                                 // 2 x stage_bytes
    uint8_t* inter = stage + 2 * p.stage_bytes;                 // rows x inter_pitch

    const int b = blockIdx.y;
    const ImgDesc d = p.desc[b];
    const int S = p.S;
    const int y0 = blockIdx.x * p.TR, y1 = min(S, y0 + p.TR);
    for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i] = p.lut[i];
    const int rowbytes = d.w_in * 3;
    const int xvecs = S >> 3;                                   // groups of 8 output pixels (S % 8 == 0)
    const bool vec8 = p.layout == 0 || (p.P & 7) == 0;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (d.identity) return;                                     // handled by convert_kernel (1) or the global-memory pair (2)

    // ---- rows of the source this tile needs (vertical taps of its first and last output row)
    const int r_lo = d.bounds_v[2 * (d.top + y0)];
    const int r_hi = d.bounds_v[2 * (d.top + y1 - 1)] + d.bounds_v[2 * (d.top + y1 - 1) + 1];
    const int nrows = r_hi - r_lo;
    const int rpc = max(1, (p.stage_bytes - 32 - 128) / rowbytes);    // source rows per staged chunk
    const int nchunks = (nrows + rpc - 1) / rpc;

    int prefix_cur = stage_span(stage, p.src, d.src_off + (long long)r_lo * rowbytes, min(rpc, nrows) * rowbytes);
    cp_async_commit();
    for (int ch = 0; ch < nchunks; ++ch) {
        uint8_t* cur = stage + (ch & 1) * p.stage_bytes;
        const int row0 = r_lo + ch * rpc;
        const int nr = min(rpc, r_hi - row0);
        int prefix_next = 0;
        if (ch + 1 < nchunks) {
            const int rown = row0 + rpc;
            prefix_next = stage_span(stage + ((ch + 1) & 1) * p.stage_bytes, p.src,
                                     d.src_off + (long long)rown * rowbytes, min(rpc, r_hi - rown) * rowbytes);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        // ---- horizontal pass: one output column per thread, all rows of the chunk
        {
            const uint8_t* rows = cur + prefix_cur;
            uint8_t* irows = inter + (size_t)(row0 - r_lo) * p.inter_pitch;
            if (d.ksize_h <= 8) hpass_chunk<8>(d, rows, rowbytes, nr, S, irows, p.inter_pitch);
            else if (d.ksize_h <= 16) hpass_chunk<16>(d, rows, rowbytes, nr, S, irows, p.inter_pitch);
            else if (d.ksize_h <= 24) hpass_chunk<24>(d, rows, rowbytes, nr, S, irows, p.inter_pitch);
            else if (d.ksize_h <= 40) hpass_chunk<40>(d, rows, rowbytes, nr, S, irows, p.inter_pitch);
            else hpass_chunk<0>(d, rows, rowbytes, nr, S, irows, p.inter_pitch);
        }
        __syncthreads();
        prefix_cur = prefix_next;
    }
    // ---- vertical pass + normalise + store: a thread owns 8 pixels x 3 channels (24 contiguous, 8-byte aligned
    //      bytes of every intermediate row: three 64-bit loads per tap)
    for (int yy = y0 + wid; yy < y1; yy += 8) {
        const int yo = d.top + yy;
        const int ymin = __ldg(d.bounds_v + 2 * yo), cnt = __ldg(d.bounds_v + 2 * yo + 1);
        const int* k = d.kk_v + (size_t)yo * d.ksize_v;
        for (int xv = lane; xv < xvecs; xv += 32) {
            const uint8_t* col = inter + (size_t)(ymin - r_lo) * p.inter_pitch + xv * 24;
            int acc[24];
#pragma unroll
            for (int j = 0; j < 24; ++j) acc[j] = 1 << (PRECISION_BITS - 1);
            for (int t = 0; t < cnt; ++t) {
                const int w = __ldg(k + t);
                const uint2* rp = reinterpret_cast<const uint2*>(col + (size_t)t * p.inter_pitch);
                const uint2 q0 = rp[0], q1 = rp[1], q2 = rp[2];
                const uint32_t q[6] = {q0.x, q0.y, q1.x, q1.y, q2.x, q2.y};
#pragma unroll
                for (int j = 0; j < 24; ++j) acc[j] += (int)byte_of(q[j >> 2], j & 3) * w;
            }
            int cstride;
            const size_t idx0 = out_index(p, b, yy, xv * 8, cstride);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = lut[c * 256 + clip8(acc[3 * j + c])];
                store8<BF16>(p, vec8, idx0 + (size_t)c * cstride, b, c, yy, xv * 8, f);
            }
        }
    }
}

static_assert(sizeof(ImgDesc) == 80, "ImgDesc is pulled as five 16-byte words");
// Copies the batch's descriptors from pinned host memory (read through the unified address space) to the device.
__global__ void pull_desc_kernel(const uint4* __restrict__ host, uint4* __restrict__ dev, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dev[i] = host[i];
}

// Identity-size images (resized size == source size, e.g. 224x224 sources for a 224 model): crop + convert only.
// Persistent CTAs walk (image, row-tile) items; the next item's rows are fetched with cp.async while the current
// ones are converted, so the kernel streams at memory speed instead of paying one load latency per tile.
// The descriptor of an item (image geometry, 80 bytes in global memory) is needed before its rows can be requested; read on the spot
// it put one L2 round trip in front of every cp.async batch and another in front of every conversion (ncu: long-scoreboard stalls on
// exactly those loads).  So descriptors run two items ahead in registers.
struct ItemDesc { long long src_off; int w_in, top, left, identity; };
template <bool BF16>
__global__ void __launch_bounds__(256, 4) convert_kernel(PreParams p, int B, int tiles) {
    extern __shared__ __align__(16) uint8_t sm[];
    float* lut = reinterpret_cast<float*>(sm);
    uint8_t* stage = sm + 3072;                                 // 2 x stage_bytes
    for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i] = p.lut[i];
    const int S = p.S, xvecs = S >> 3;
    const bool vec8 = p.layout == 0 || (p.P & 7) == 0;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = B * tiles;
    auto load_desc = [&](int item) -> ItemDesc {
        ItemDesc r{0, 0, 0, 0, 0};
        if (item < total) {
            const ImgDesc& d = p.desc[item / tiles];
            r.src_off = d.src_off; r.w_in = d.w_in; r.top = d.top; r.left = d.left; r.identity = d.identity;
        }
        return r;
    };
    auto fetch = [&](int item, const ItemDesc& d, int buf) -> int {   // returns the byte offset of the first row inside the buffer
        const int y0 = (item % tiles) * p.TR;
        int prefix = 0;
        if (d.identity == 1) {
            const int rowbytes = d.w_in * 3, nr = min(S, y0 + p.TR) - y0;
            prefix = stage_span(stage + buf * p.stage_bytes, p.src, d.src_off + (long long)(d.top + y0) * rowbytes, nr * rowbytes);
        }
        cp_async_commit();
        return prefix;
    };
    int item = blockIdx.x, it = 0;
    if (item >= total) return;
    ItemDesc dc = load_desc(item);
    int prefix = fetch(item, dc, 0);
    ItemDesc dn = load_desc(item + gridDim.x);
    for (; item < total; item += gridDim.x, ++it) {
        const int next = item + gridDim.x;
        const ItemDesc dnn = load_desc(next + gridDim.x);          // consumed one iteration from now
        int prefix_next = 0;
        if (next < total) { prefix_next = fetch(next, dn, (it + 1) & 1); cp_async_wait<1>(); } else cp_async_wait<0>();
        __syncthreads();
        const int b = item / tiles, y0 = (item - b * tiles) * p.TR;
        if (dc.identity == 1) {
            const int rowbytes = dc.w_in * 3, nr = min(S, y0 + p.TR) - y0;
            const uint8_t* base = stage + (it & 1) * p.stage_bytes + prefix + dc.left * 3;
            for (int r = wid; r < nr; r += 8)
                for (int xv = lane; xv < xvecs; xv += 32) {
                    uint32_t al[6];
                    load_bytes_aligned<6>(base + r * rowbytes + xv * 24, al);
                    int cstride;
                    const size_t idx0 = out_index(p, b, y0 + r, xv * 8, cstride);
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        float f[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = lut[c * 256 + byte_of(al[(3 * j + c) >> 2], (3 * j + c) & 3)];
                        store8<BF16>(p, vec8, idx0 + (size_t)c * cstride, b, c, y0 + r, xv * 8, f);
                    }
                }
        }
        __syncthreads();
        prefix = prefix_next; dc = dn; dn = dnn;
    }
}

// Images no shared-memory tile can hold (a source row wider than a staging buffer, or a vertical reduction whose taps span more rows
// than fit): the same two integer passes through a global-memory intermediate, one thread per intermediate / output pixel.  Rare by
// construction (panoramas, 20k-pixel scans), so simple rather than fast; the arithmetic is identical, tap for tap.
__global__ void __launch_bounds__(256) hpass_global_kernel(const uint8_t* __restrict__ src, ImgDesc d, int S, int r_lo, int nrows,
                                                           uint8_t* __restrict__ inter) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)nrows * S) return;
    const int r = (int)(i / S), x = (int)(i - (long long)r * S), xo = d.left + x;
    const int xmin = d.bounds_h[2 * xo], cnt = d.bounds_h[2 * xo + 1];
    const int* k = d.kk_h + (size_t)xo * d.ksize_h;
    const uint8_t* px = src + d.src_off + ((long long)(r_lo + r) * d.w_in + xmin) * 3;
    int a0 = 1 << (PRECISION_BITS - 1), a1 = a0, a2 = a0;
    for (int j = 0; j < cnt; ++j) {
        const int w = k[j];
        a0 += px[3 * j] * w; a1 += px[3 * j + 1] * w; a2 += px[3 * j + 2] * w;
    }
    uint8_t* o = inter + ((size_t)r * S + x) * 3;
    o[0] = (uint8_t)clip8(a0); o[1] = (uint8_t)clip8(a1); o[2] = (uint8_t)clip8(a2);
}
template <bool BF16>
__global__ void __launch_bounds__(256) vpass_global_kernel(PreParams p, ImgDesc d, int b, int r_lo, const uint8_t* __restrict__ inter) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, S = p.S;
    if (i >= S * S) return;
    const int y = i / S, x = i - y * S, yo = d.top + y;
    const int ymin = d.bounds_v[2 * yo], cnt = d.bounds_v[2 * yo + 1];
    const int* k = d.kk_v + (size_t)yo * d.ksize_v;
    const uint8_t* col = inter + ((size_t)(ymin - r_lo) * S + x) * 3;
    int a0 = 1 << (PRECISION_BITS - 1), a1 = a0, a2 = a0;
    for (int t = 0; t < cnt; ++t) {
        const int w = k[t];
        const uint8_t* q = col + (size_t)t * S * 3;
        a0 += q[0] * w; a1 += q[1] * w; a2 += q[2] * w;
    }
    int cstride;
    const size_t idx = out_index(p, b, y, x, cstride);
    const float f0 = p.lut[clip8(a0)], f1 = p.lut[256 + clip8(a1)], f2 = p.lut[512 + clip8(a2)];
    if (BF16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out);
        o[idx] = __float2bfloat16_rn(f0); o[idx + cstride] = __float2bfloat16_rn(f1); o[idx + 2 * (size_t)cstride] = __float2bfloat16_rn(f2);
    } else {
        float* o = reinterpret_cast<float*>(p.out);
        o[idx] = f0; o[idx + cstride] = f1; o[idx + 2 * (size_t)cstride] = f2;
    }
}

// PIL.Image.resize runs the VERTICAL resize first (a complete resize, rounded to 8 bits) when the source is more than 100 times taller
// than wide and shrinks vertically; the bytes differ from horizontal-first.  Such strips take these two kernels: columns c_lo.. of the
// source that the crop window's horizontal taps touch -> S rows of uint8, then the horizontal pass + normalise.  One thread per element.
__global__ void __launch_bounds__(256) vfirst_v_kernel(const uint8_t* __restrict__ src, ImgDesc d, int S, int c_lo, int ncols,
                                                       uint8_t* __restrict__ inter) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S * ncols) return;
    const int y = i / ncols, x = i - y * ncols, yo = d.top + y;
    const int ymin = d.bounds_v[2 * yo], cnt = d.bounds_v[2 * yo + 1];
    const int* k = d.kk_v + (size_t)yo * d.ksize_v;
    const uint8_t* col = src + d.src_off + ((long long)ymin * d.w_in + (c_lo + x)) * 3;
    int a0 = 1 << (PRECISION_BITS - 1), a1 = a0, a2 = a0;
    for (int t = 0; t < cnt; ++t) {
        const int w = k[t];
        const uint8_t* q = col + (long long)t * d.w_in * 3;
        a0 += q[0] * w; a1 += q[1] * w; a2 += q[2] * w;
    }
    uint8_t* o = inter + ((size_t)y * ncols + x) * 3;
    o[0] = (uint8_t)clip8(a0); o[1] = (uint8_t)clip8(a1); o[2] = (uint8_t)clip8(a2);
}
template <bool BF16>
__global__ void __launch_bounds__(256) vfirst_h_kernel(PreParams p, ImgDesc d, int b, int c_lo, int ncols, const uint8_t* __restrict__ inter) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, S = p.S;
    if (i >= S * S) return;
    const int y = i / S, x = i - y * S, xo = d.left + x;
    const int xmin = d.bounds_h[2 * xo], cnt = d.bounds_h[2 * xo + 1];
    const int* k = d.kk_h + (size_t)xo * d.ksize_h;
    const uint8_t* px = inter + ((size_t)y * ncols + (xmin - c_lo)) * 3;
    int a0 = 1 << (PRECISION_BITS - 1), a1 = a0, a2 = a0;
    for (int j = 0; j < cnt; ++j) {
        const int w = k[j];
        a0 += px[3 * j] * w; a1 += px[3 * j + 1] * w; a2 += px[3 * j + 2] * w;
    }
    int cstride;
    const size_t idx = out_index(p, b, y, x, cstride);
    const float f0 = p.lut[clip8(a0)], f1 = p.lut[256 + clip8(a1)], f2 = p.lut[512 + clip8(a2)];
    if (BF16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out);
        o[idx] = __float2bfloat16_rn(f0); o[idx + cstride] = __float2bfloat16_rn(f1); o[idx + 2 * (size_t)cstride] = __float2bfloat16_rn(f2);
    } else {
        float* o = reinterpret_cast<float*>(p.out);
        o[idx] = f0; o[idx + cstride] = f1; o[idx + 2 * (size_t)cstride] = f2;
    }
}

// ------------------------------------------------------------------------------------------ host side
struct AxisTable {
    int ksize = 0; int* d_bounds = nullptr; int* d_kk = nullptr; double scale = 1.0;   // both arrays live in one allocation (d_bounds)
    unsigned long long last_use = 0;
    std::vector<int> h_bounds;          // host copy of (xmin, count) per output position: the driver needs the rows a crop window touches
};

// Pillow's precompute_coeffs + normalize_coeffs_8bpc for the bicubic filter (a = -0.5, support 2), in double.
static double keys_cubic(double x) {
    const double a = -0.5;
    if (x < 0.0) x = -x;
    if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
    if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
    return 0.0;
}

static int build_axis_table(int in_size, int out_size, AxisTable* t, cudaStream_t st) {
    const double scale = (double)in_size / out_size;
    const double fscale = scale < 1.0 ? 1.0 : scale;
    const double support = 2.0 * fscale;
    const int ksize = (int)std::ceil(support) * 2 + 1;
    std::vector<int> bounds(2 * (size_t)out_size), kk((size_t)out_size * ksize, 0);
    std::vector<double> w(ksize);
    for (int xx = 0; xx < out_size; ++xx) {
        const double center = (xx + 0.5) * scale, ss = 1.0 / fscale;
        int xmin = (int)(center - support + 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = (int)(center + support + 0.5);
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        double ww = 0.0;
        for (int x = 0; x < xmax; ++x) { w[x] = keys_cubic((x + xmin - center + 0.5) * ss); ww += w[x]; }
        for (int x = 0; x < xmax; ++x) {
            double v = w[x];
            if (ww != 0.0) v /= ww;
            kk[(size_t)xx * ksize + x] = v < 0 ? (int)(-0.5 + v * (1 << PRECISION_BITS)) : (int)(0.5 + v * (1 << PRECISION_BITS));
        }
        bounds[2 * xx] = xmin; bounds[2 * xx + 1] = xmax;
    }
    // one allocation, one copy per table (bounds first, coefficients behind them, 16-byte aligned)
    const size_t nb = (bounds.size() + 3) & ~(size_t)3;
    std::vector<int> blob(nb + kk.size(), 0);
    std::copy(bounds.begin(), bounds.end(), blob.begin());
    std::copy(kk.begin(), kk.end(), blob.begin() + nb);
    B200_CUDA_OK(cudaMalloc(&t->d_bounds, blob.size() * sizeof(int)));
    t->d_kk = t->d_bounds + nb;
    const cudaError_t e = cudaMemcpyAsync(t->d_bounds, blob.data(), blob.size() * sizeof(int), cudaMemcpyHostToDevice, st);
    const cudaError_t e2 = e == cudaSuccess ? cudaStreamSynchronize(st) : e;   // the host vector dies at return
    if (e2 != cudaSuccess) { cudaFree(t->d_bounds); t->d_bounds = t->d_kk = nullptr; B200_CUDA_OK(e2); }
    t->ksize = ksize; t->scale = scale;
    t->h_bounds = std::move(bounds);
    return B200_OK;
}

struct PreprocessState {
    std::map<std::pair<int, int>, AxisTable> tables;
    unsigned long long call_id = 0;     // stamps the tables a call uses (least-recently-used eviction)
    uint8_t* d_inter = nullptr; size_t inter_cap = 0;   // global-memory intermediate of the oversize path
    float* d_lut = nullptr;
    // descriptors: filled in one of four pinned host buffers, pulled into d_desc by a small kernel (zero-copy
    // reads).  No copy-engine transfer, so a caller's large uploads on another stream cannot delay this call.
    ImgDesc* d_desc = nullptr; ImgDesc* h_desc[4] = {nullptr, nullptr, nullptr, nullptr}; int desc_cap = 0; int desc_slot = 0;
    cudaEvent_t desc_free[4] = {nullptr, nullptr, nullptr, nullptr};
    ~PreprocessState() {
        for (auto& kv : tables) cudaFree(kv.second.d_bounds);
        cudaFree(d_lut); cudaFree(d_desc); cudaFree(d_inter);
        for (int i = 0; i < 4; ++i) { if (h_desc[i]) cudaFreeHost(h_desc[i]); if (desc_free[i]) cudaEventDestroy(desc_free[i]); }
    }
};

PreprocessState* preprocess_state_create(const float* mean, const float* stdv, int* rc) {
    auto* s = new PreprocessState();
    std::vector<float> lut(768);
    for (int c = 0; c < 3; ++c)
        for (int v = 0; v < 256; ++v) lut[c * 256 + v] = ((float)v / 255.0f - mean[c]) / stdv[c];   // fp32, reference order
    cudaError_t e = cudaMalloc(&s->d_lut, 768 * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(s->d_lut, lut.data(), 768 * sizeof(float), cudaMemcpyHostToDevice);
    for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&s->desc_free[i], cudaEventDisableTiming);
    if (e != cudaSuccess) { *rc = fail(B200_ECUDA, std::string("preprocess state: ") + cudaGetErrorString(e)); delete s; return nullptr; }
    *rc = B200_OK;
    return s;
}
void preprocess_state_destroy(PreprocessState* s) { delete s; }

// A long-running service meets an unbounded number of (source, target) sizes; the cache holds the most recently used ones.  When it
// is full, the older half is dropped - after a device synchronise, because a kernel still in flight may be reading those tables -
// except the tables the current call has already picked up.  (std::map nodes are stable: pointers handed out stay valid.)
constexpr size_t kMaxAxisTables = 4096;
static int get_table(PreprocessState* s, int in_size, int out_size, const AxisTable** out, cudaStream_t st) {
    auto key = std::make_pair(in_size, out_size);
    auto it = s->tables.find(key);
    if (it == s->tables.end()) {
        if (s->tables.size() >= kMaxAxisTables) {
            std::vector<std::pair<unsigned long long, std::pair<int, int>>> old;
            for (auto& kv : s->tables) if (kv.second.last_use != s->call_id) old.push_back({kv.second.last_use, kv.first});
            std::sort(old.begin(), old.end());
            const size_t drop = std::min(old.size(), s->tables.size() - kMaxAxisTables / 2);
            if (drop > 0) {
                B200_CUDA_OK(cudaDeviceSynchronize());
                for (size_t i = 0; i < drop; ++i) {
                    auto victim = s->tables.find(old[i].second);
                    cudaFree(victim->second.d_bounds);
                    s->tables.erase(victim);
                }
            }
        }
        AxisTable t;
        int rc = build_axis_table(in_size, out_size, &t, st);
        if (rc) return rc;
        it = s->tables.emplace(key, std::move(t)).first;
    }
    it->second.last_use = s->call_id;
    *out = &it->second;
    return B200_OK;
}

template <bool BF16> static int launch_pre(const PreParams& p, int tiles, int B, size_t smem, cudaStream_t st) {
    if (smem > 48 * 1024)
        B200_CUDA_OK(cudaFuncSetAttribute(preprocess_kernel<BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    preprocess_kernel<BF16><<<dim3(tiles, B), 256, smem, st>>>(p);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

// src_dev: packed uint8 HWC images, image i at byte offset offsets[i] (multiple of 16; 16 spare bytes after the last image).
// hw: [B][2] = (height, width).  mode 0 = squash, 1 = shortest side + centre crop.
// layout 0: out[B,3,S,S]; layout 1: out[B*n, kpad] patch-major with K index (c, i, j) and zero padding left untouched.
int preprocess_run(PreprocessState* s, const uint8_t* src_dev, const long long* offsets, const int* hw, int B, int S,
                   int mode, int layout, int out_bf16, int P, int kpad, void* out_dev, cudaStream_t st) {
    B200_REQUIRE(B >= 0 && S > 0 && S % 8 == 0, "preprocess: target size must be a positive multiple of 8");
    if (B == 0) return B200_OK;
    B200_REQUIRE(src_dev && offsets && hw && out_dev, "preprocess: null buffer");
    B200_REQUIRE(mode == 0 || mode == 1, "preprocess: mode must be 0 (squash) or 1 (shortest side + centre crop)");
    B200_REQUIRE(layout == 0 || (layout == 1 && P > 0 && S % P == 0 && kpad >= 3 * P * P && P % 2 == 0),
                 "preprocess: bad patch-major geometry");
    if (B > s->desc_cap) {
        B200_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(s->d_desc); s->d_desc = nullptr;
        for (int i = 0; i < 4; ++i) { if (s->h_desc[i]) cudaFreeHost(s->h_desc[i]); s->h_desc[i] = nullptr; }
        int cap = B < 512 ? 512 : B;
        B200_CUDA_OK(cudaMalloc(&s->d_desc, sizeof(ImgDesc) * cap));
        for (int i = 0; i < 4; ++i) B200_CUDA_OK(cudaMallocHost(&s->h_desc[i], sizeof(ImgDesc) * cap));
        s->desc_cap = cap;
    }
    s->desc_slot = (s->desc_slot + 1) & 3;
    s->call_id += 1;
    ImgDesc* hdesc = s->h_desc[s->desc_slot];
    B200_CUDA_OK(cudaEventSynchronize(s->desc_free[s->desc_slot]));   // the call four back has pulled its descriptors
    const int inter_pitch = (S * 3 + 15) & ~15;
    constexpr int kMaxStage = 65536;                       // largest staging buffer of the tiled kernel
    auto rows_needed_for = [&](double vs, int tr) { return (int)std::ceil(vs * tr + 4.0 * (vs < 1 ? 1 : vs)) + 3; };
    std::vector<const AxisTable*> vtab(B, nullptr), htab(B, nullptr);
    int max_rowbytes = 0, max_ident_rowbytes = 0;
    for (int i = 0; i < B; ++i) {
        const int h = hw[2 * i], w = hw[2 * i + 1];
        B200_REQUIRE(h > 0 && w > 0, "preprocess: image with non-positive size");
        B200_REQUIRE((offsets[i] & 15) == 0, "preprocess: image offsets must be multiples of 16 bytes");
        int new_h = S, new_w = S, top = 0, left = 0;
        if (mode == 1) {   // torchvision Resize(int) + CenterCrop: int(S*long/short), round-half-even offsets
            if (w <= h) { new_w = S; new_h = (int)((long long)S * h / w); } else { new_h = S; new_w = (int)((long long)S * w / h); }
            top = (int)std::nearbyint((new_h - S) / 2.0);
            left = (int)std::nearbyint((new_w - S) / 2.0);
        }
        ImgDesc& d = hdesc[i];
        d.src_off = offsets[i]; d.h_in = h; d.w_in = w; d.top = top; d.left = left; d.pad_[0] = d.pad_[1] = d.pad_[2] = 0;
        d.identity = (new_h == h && new_w == w) ? 1 : 0;
        d.bounds_h = d.kk_h = d.bounds_v = d.kk_v = nullptr; d.ksize_h = d.ksize_v = 0;
        if (!d.identity) {
            const AxisTable *th, *tv;
            int rc = get_table(s, w, new_w, &th, st); if (rc) return rc;
            rc = get_table(s, h, new_h, &tv, st); if (rc) return rc;
            d.bounds_h = th->d_bounds; d.kk_h = th->d_kk; d.ksize_h = th->ksize;
            d.bounds_v = tv->d_bounds; d.kk_v = tv->d_kk; d.ksize_v = tv->ksize;
            vtab[i] = tv; htab[i] = th;
            if ((long long)h > (long long)w * 100 && new_h < h) d.identity = 3;      // PIL resizes such strips vertically first (see vfirst_v_kernel)
            else if (w * 3 + 32 + 128 > kMaxStage) d.identity = 2;      // a source row does not fit a staging buffer
            else if (w * 3 > max_rowbytes) max_rowbytes = w * 3;
        } else if (w * 3 > max_ident_rowbytes) max_ident_rowbytes = w * 3;
    }
    // Staging: two buffers of `stage_bytes` (>= one source row + slack for the zero-weight over-reads of the register-tiled
    // horizontal pass); what is left of ~200 KB holds the uint8 intermediate rows of one tile.
    int stage_bytes = 16384;
    while (stage_bytes < max_rowbytes + 32 + 128) stage_bytes *= 2;
    const size_t budget = 200 * 1024 - 3072 - 2 * (size_t)stage_bytes;
    double max_vscale = 1.0; bool any_resize = false, any_oversize = false;
    for (int i = 0; i < B; ++i) {
        ImgDesc& d = hdesc[i];
        if (d.identity == 0 && (size_t)rows_needed_for(vtab[i]->scale, 1) * inter_pitch > budget) d.identity = 2;   // taps of ONE output row do not fit
        if (d.identity == 0) { any_resize = true; if (vtab[i]->scale > max_vscale) max_vscale = vtab[i]->scale; }
        any_oversize |= d.identity >= 2;
    }
    pull_desc_kernel<<<(B * 5 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const uint4*>(hdesc), reinterpret_cast<uint4*>(s->d_desc), B * 5);
    B200_CUDA_OK(cudaGetLastError());
    B200_CUDA_OK(cudaEventRecord(s->desc_free[s->desc_slot], st));
    int rc = B200_OK;
    if (max_ident_rowbytes > 0) {
        // identity images: TR rows per item, one staging buffer per item (double buffered)
        int TR = 16;
        while (TR > 1 && (size_t)TR * max_ident_rowbytes + 48 > 24 * 1024) TR /= 2;      // two buffers per CTA, four CTAs per SM
        const int ident_stage = (TR * max_ident_rowbytes + 48 + 15) & ~15;
        B200_REQUIRE(ident_stage <= 96 * 1024, "preprocess: source rows too wide");
        const size_t smem = 3072 + 2 * (size_t)ident_stage;
        const int tiles = (S + TR - 1) / TR;
        PreParams p{src_dev, s->d_desc, s->d_lut, out_dev, S, TR, layout, out_bf16, P, layout ? S / P : 0, kpad, inter_pitch, ident_stage};
        const long long total = (long long)B * tiles;
        const int grid = (int)std::min<long long>(total, 8LL * kNumSMs);
        if (out_bf16) {
            if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(convert_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            convert_kernel<true><<<grid, 256, smem, st>>>(p, B, tiles);
        } else {
            if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(convert_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            convert_kernel<false><<<grid, 256, smem, st>>>(p, B, tiles);
        }
        B200_CUDA_OK(cudaGetLastError());
    }
    if (any_oversize) {
        // one image at a time through the global-memory pair; the intermediate holds every source row the crop window touches
        PreParams p{src_dev, s->d_desc, s->d_lut, out_dev, S, 1, layout, out_bf16, P, layout ? S / P : 0, kpad, inter_pitch, 0};
        for (int i = 0; i < B; ++i) {
            const ImgDesc& d = hdesc[i];
            if (d.identity == 3) {
                const std::vector<int>& bh = htab[i]->h_bounds;
                const int c_lo = bh[2 * d.left], c_hi = bh[2 * (d.left + S - 1)] + bh[2 * (d.left + S - 1) + 1], ncols = c_hi - c_lo;
                const size_t need = (size_t)S * ncols * 3;
                if (need > s->inter_cap) {
                    B200_CUDA_OK(cudaStreamSynchronize(st));
                    cudaFree(s->d_inter); s->d_inter = nullptr; s->inter_cap = 0;
                    B200_CUDA_OK(cudaMalloc(&s->d_inter, need + need / 4));
                    s->inter_cap = need + need / 4;
                }
                vfirst_v_kernel<<<(S * ncols + 255) / 256, 256, 0, st>>>(src_dev, d, S, c_lo, ncols, s->d_inter);
                B200_CUDA_OK(cudaGetLastError());
                if (out_bf16) vfirst_h_kernel<true><<<(S * S + 255) / 256, 256, 0, st>>>(p, d, i, c_lo, ncols, s->d_inter);
                else vfirst_h_kernel<false><<<(S * S + 255) / 256, 256, 0, st>>>(p, d, i, c_lo, ncols, s->d_inter);
                B200_CUDA_OK(cudaGetLastError());
                continue;
            }
            if (d.identity != 2) continue;
            const std::vector<int>& bv = vtab[i]->h_bounds;
            const int r_lo = bv[2 * d.top], r_hi = bv[2 * (d.top + S - 1)] + bv[2 * (d.top + S - 1) + 1];
            const size_t need = (size_t)(r_hi - r_lo) * S * 3;
            if (need > s->inter_cap) {
                B200_CUDA_OK(cudaStreamSynchronize(st));
                cudaFree(s->d_inter); s->d_inter = nullptr; s->inter_cap = 0;
                B200_CUDA_OK(cudaMalloc(&s->d_inter, need + need / 4));
                s->inter_cap = need + need / 4;
            }
            const long long n1 = (long long)(r_hi - r_lo) * S;
            hpass_global_kernel<<<(unsigned)((n1 + 255) / 256), 256, 0, st>>>(src_dev, d, S, r_lo, r_hi - r_lo, s->d_inter);
            B200_CUDA_OK(cudaGetLastError());
            if (out_bf16) vpass_global_kernel<true><<<(S * S + 255) / 256, 256, 0, st>>>(p, d, i, r_lo, s->d_inter);
            else vpass_global_kernel<false><<<(S * S + 255) / 256, 256, 0, st>>>(p, d, i, r_lo, s->d_inter);
            B200_CUDA_OK(cudaGetLastError());
        }
    }
    if (!any_resize) return rc;
    // Tile height: as many output rows as keep the uint8 intermediate within ~64 KB (whole tile within 200 KB).
    auto rows_needed = [&](int tr) { return rows_needed_for(max_vscale, tr); };
    int TR = 32;
    while (TR > 1 && ((size_t)rows_needed(TR) * inter_pitch > 65536)) TR /= 2;
    B200_REQUIRE((size_t)rows_needed(TR) * inter_pitch <= budget, "preprocess: internal error: tile does not fit (oversize images should have been diverted)");
    const size_t smem = 3072 + 2 * (size_t)stage_bytes + (size_t)rows_needed(TR) * inter_pitch;
    PreParams p{src_dev, s->d_desc, s->d_lut, out_dev, S, TR, layout, out_bf16, P, layout ? S / P : 0, kpad, inter_pitch, stage_bytes};
    const int tiles = (S + TR - 1) / TR;
    return out_bf16 ? launch_pre<true>(p, tiles, B, smem, st) : launch_pre<false>(p, tiles, B, smem, st);
}

}  // namespace b200
