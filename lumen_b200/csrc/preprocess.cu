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
    int identity;                      // resized size == source size: pure crop + convert
    int pad_;
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

template <int VW, bool BF16>
__device__ __forceinline__ void store_vec(const PreParams& p, int b, int c, int y, int x, const float (&f)[VW]) {
    size_t idx;
    if (p.layout == 0) idx = (((size_t)b * 3 + c) * p.S + y) * p.S + x;
    else idx = ((size_t)b * p.npatch_side * p.npatch_side + (size_t)(y / p.P) * p.npatch_side + x / p.P) * p.kpad +
               (size_t)c * p.P * p.P + (y % p.P) * p.P + (x % p.P);
    if (BF16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + idx;
        if (VW == 8) {
            uint4 v;
            __nv_bfloat162 t0 = __floats2bfloat162_rn(f[0], f[1]), t1 = __floats2bfloat162_rn(f[2], f[3]);
            __nv_bfloat162 t2 = __floats2bfloat162_rn(f[4], f[5]), t3 = __floats2bfloat162_rn(f[6], f[7]);
            v.x = *reinterpret_cast<uint32_t*>(&t0); v.y = *reinterpret_cast<uint32_t*>(&t1);
            v.z = *reinterpret_cast<uint32_t*>(&t2); v.w = *reinterpret_cast<uint32_t*>(&t3);
            *reinterpret_cast<uint4*>(o) = v;
        } else {
#pragma unroll
            for (int i = 0; i < VW; i += 2) *reinterpret_cast<__nv_bfloat162*>(o + i) = __floats2bfloat162_rn(f[i], f[i + 1]);
        }
    } else {
        float* o = reinterpret_cast<float*>(p.out) + idx;
        if (VW == 8) {
            *reinterpret_cast<float4*>(o) = make_float4(f[0], f[1], f[2], f[3]);
            *reinterpret_cast<float4*>(o + 4) = make_float4(f[4], f[5], f[6], f[7]);
        } else {
#pragma unroll
            for (int i = 0; i < VW; i += 2) *reinterpret_cast<float2*>(o + i) = make_float2(f[i], f[i + 1]);
        }
    }
}

template <int VW, bool BF16>
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
    const int xvecs = S / VW;

    if (d.identity) {
        // Crop + convert only: stage source rows [top+y0, top+y1) (one contiguous span) and write the planes.
        const int nr = y1 - y0;
        const int prefix = stage_span(stage, p.src, d.src_off + (long long)(d.top + y0) * rowbytes, nr * rowbytes);
        cp_async_commit();
        cp_async_wait<0>();
        __syncthreads();
        const uint8_t* base = stage + prefix + d.left * 3;
        for (int i = threadIdx.x; i < nr * 3 * xvecs; i += blockDim.x) {
            const int xv = i % xvecs, c = (i / xvecs) % 3, r = i / (3 * xvecs);
            const uint8_t* px = base + r * rowbytes + xv * VW * 3 + c;
            float f[VW];
#pragma unroll
            for (int j = 0; j < VW; ++j) f[j] = lut[c * 256 + px[j * 3]];
            store_vec<VW, BF16>(p, b, c, y0 + r, xv * VW, f);
        }
        return;
    }

    // ---- rows of the source this tile needs (vertical taps of its first and last output row)
    const int r_lo = d.bounds_v[2 * (d.top + y0)];
    const int r_hi = d.bounds_v[2 * (d.top + y1 - 1)] + d.bounds_v[2 * (d.top + y1 - 1) + 1];
    const int nrows = r_hi - r_lo;
    const int rpc = max(1, (p.stage_bytes - 32) / rowbytes);    // source rows per staged chunk
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
        // ---- horizontal pass: (row, x) -> 3 channels
        for (int i = threadIdx.x; i < nr * S; i += blockDim.x) {
            const int r = i / S, x = i - r * S;
            const int xo = d.left + x;
            const int xmin = __ldg(d.bounds_h + 2 * xo), cnt = __ldg(d.bounds_h + 2 * xo + 1);
            const int* k = d.kk_h + (size_t)xo * d.ksize_h;
            const uint8_t* px = cur + prefix_cur + r * rowbytes + xmin * 3;
            int a0 = 1 << (PRECISION_BITS - 1), a1 = a0, a2 = a0;
            for (int j = 0; j < cnt; ++j) {
                const int w = __ldg(k + j);
                a0 += px[3 * j] * w; a1 += px[3 * j + 1] * w; a2 += px[3 * j + 2] * w;
            }
            uint8_t* o = inter + (size_t)(row0 - r_lo + r) * p.inter_pitch + x * 3;
            o[0] = (uint8_t)clip8(a0); o[1] = (uint8_t)clip8(a1); o[2] = (uint8_t)clip8(a2);
        }
        __syncthreads();
        prefix_cur = prefix_next;
    }
    // ---- vertical pass + normalise + store
    for (int i = threadIdx.x; i < (y1 - y0) * 3 * xvecs; i += blockDim.x) {
        const int xv = i % xvecs, c = (i / xvecs) % 3, y = y0 + i / (3 * xvecs);
        const int yo = d.top + y;
        const int ymin = __ldg(d.bounds_v + 2 * yo), cnt = __ldg(d.bounds_v + 2 * yo + 1);
        const int* k = d.kk_v + (size_t)yo * d.ksize_v;
        const uint8_t* col = inter + (size_t)(ymin - r_lo) * p.inter_pitch + xv * VW * 3 + c;
        int acc[VW];
#pragma unroll
        for (int j = 0; j < VW; ++j) acc[j] = 1 << (PRECISION_BITS - 1);
        for (int t = 0; t < cnt; ++t) {
            const int w = __ldg(k + t);
            const uint8_t* row = col + (size_t)t * p.inter_pitch;
#pragma unroll
            for (int j = 0; j < VW; ++j) acc[j] += row[j * 3] * w;
        }
        float f[VW];
#pragma unroll
        for (int j = 0; j < VW; ++j) f[j] = lut[c * 256 + clip8(acc[j])];
        store_vec<VW, BF16>(p, b, c, y, xv * VW, f);
    }
}

// ------------------------------------------------------------------------------------------ host side
struct AxisTable { int ksize = 0; int* d_bounds = nullptr; int* d_kk = nullptr; double scale = 1.0; };

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
    B200_CUDA_OK(cudaMalloc(&t->d_bounds, bounds.size() * sizeof(int)));
    B200_CUDA_OK(cudaMalloc(&t->d_kk, kk.size() * sizeof(int)));
    B200_CUDA_OK(cudaMemcpyAsync(t->d_bounds, bounds.data(), bounds.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    B200_CUDA_OK(cudaMemcpyAsync(t->d_kk, kk.data(), kk.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    B200_CUDA_OK(cudaStreamSynchronize(st));   // host vectors die at return
    t->ksize = ksize; t->scale = scale;
    return B200_OK;
}

struct PreprocessState {
    std::map<std::pair<int, int>, AxisTable> tables;
    float* d_lut = nullptr;
    ImgDesc* d_desc = nullptr; ImgDesc* h_desc = nullptr; int desc_cap = 0;
    cudaEvent_t desc_free = nullptr;
    ~PreprocessState() {
        for (auto& kv : tables) { cudaFree(kv.second.d_bounds); cudaFree(kv.second.d_kk); }
        cudaFree(d_lut); cudaFree(d_desc);
        if (h_desc) cudaFreeHost(h_desc);
        if (desc_free) cudaEventDestroy(desc_free);
    }
};

PreprocessState* preprocess_state_create(const float* mean, const float* stdv, int* rc) {
    auto* s = new PreprocessState();
    std::vector<float> lut(768);
    for (int c = 0; c < 3; ++c)
        for (int v = 0; v < 256; ++v) lut[c * 256 + v] = ((float)v / 255.0f - mean[c]) / stdv[c];   // fp32, reference order
    cudaError_t e = cudaMalloc(&s->d_lut, 768 * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(s->d_lut, lut.data(), 768 * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->desc_free, cudaEventDisableTiming);
    if (e != cudaSuccess) { *rc = fail(B200_ECUDA, std::string("preprocess state: ") + cudaGetErrorString(e)); delete s; return nullptr; }
    *rc = B200_OK;
    return s;
}
void preprocess_state_destroy(PreprocessState* s) { delete s; }

static int get_table(PreprocessState* s, int in_size, int out_size, const AxisTable** out, cudaStream_t st) {
    auto key = std::make_pair(in_size, out_size);
    auto it = s->tables.find(key);
    if (it == s->tables.end()) {
        if (s->tables.size() >= 4096) return fail(B200_ESTATE, "preprocess: more than 4096 distinct (source, target) sizes cached");
        AxisTable t;
        int rc = build_axis_table(in_size, out_size, &t, st);
        if (rc) return rc;
        it = s->tables.emplace(key, t).first;
    }
    *out = &it->second;
    return B200_OK;
}

template <int VW, bool BF16> static int launch_pre(const PreParams& p, int tiles, int B, size_t smem, cudaStream_t st) {
    static size_t configured = 0;
    if (smem > configured) {
        B200_CUDA_OK(cudaFuncSetAttribute(preprocess_kernel<VW, BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    preprocess_kernel<VW, BF16><<<dim3(tiles, B), 256, smem, st>>>(p);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

// src_dev: packed uint8 HWC images, image i at byte offset offsets[i] (multiple of 16; 16 spare bytes after the last image).
// hw: [B][2] = (height, width).  mode 0 = squash, 1 = shortest side + centre crop.
// layout 0: out[B,3,S,S]; layout 1: out[B*n, kpad] patch-major with K index (c, i, j) and zero padding left untouched.
int preprocess_run(PreprocessState* s, const uint8_t* src_dev, const long long* offsets, const int* hw, int B, int S,
                   int mode, int layout, int out_bf16, int P, int kpad, void* out_dev, cudaStream_t st) {
    B200_REQUIRE(B >= 0 && S > 0 && S % 2 == 0, "preprocess: target size must be positive and even");
    if (B == 0) return B200_OK;
    B200_REQUIRE(src_dev && offsets && hw && out_dev, "preprocess: null buffer");
    B200_REQUIRE(mode == 0 || mode == 1, "preprocess: mode must be 0 (squash) or 1 (shortest side + centre crop)");
    B200_REQUIRE(layout == 0 || (layout == 1 && P > 0 && S % P == 0 && kpad >= 3 * P * P && P % 2 == 0),
                 "preprocess: bad patch-major geometry");
    if (B > s->desc_cap) {
        if (s->d_desc) { cudaFree(s->d_desc); cudaFreeHost(s->h_desc); s->d_desc = nullptr; s->h_desc = nullptr; }
        int cap = B < 512 ? 512 : B;
        B200_CUDA_OK(cudaMalloc(&s->d_desc, sizeof(ImgDesc) * cap));
        B200_CUDA_OK(cudaMallocHost(&s->h_desc, sizeof(ImgDesc) * cap));
        s->desc_cap = cap;
    } else {
        B200_CUDA_OK(cudaEventSynchronize(s->desc_free));    // previous upload out of the pinned staging area
    }
    const int inter_pitch = (S * 3 + 15) & ~15;
    double max_vscale = 1.0; int max_rowbytes = 0; bool any_resize = false; int max_ident_rowbytes = 0;
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
        ImgDesc& d = s->h_desc[i];
        d.src_off = offsets[i]; d.h_in = h; d.w_in = w; d.top = top; d.left = left; d.pad_ = 0;
        d.identity = (new_h == h && new_w == w) ? 1 : 0;
        d.bounds_h = d.kk_h = d.bounds_v = d.kk_v = nullptr; d.ksize_h = d.ksize_v = 0;
        if (!d.identity) {
            const AxisTable *th, *tv;
            int rc = get_table(s, w, new_w, &th, st); if (rc) return rc;
            rc = get_table(s, h, new_h, &tv, st); if (rc) return rc;
            d.bounds_h = th->d_bounds; d.kk_h = th->d_kk; d.ksize_h = th->ksize;
            d.bounds_v = tv->d_bounds; d.kk_v = tv->d_kk; d.ksize_v = tv->ksize;
            if (tv->scale > max_vscale) max_vscale = tv->scale;
            any_resize = true;
            if (w * 3 > max_rowbytes) max_rowbytes = w * 3;
        } else if (w * 3 > max_ident_rowbytes) max_ident_rowbytes = w * 3;
    }
    // Tile height: as many output rows as keep the uint8 intermediate within ~64 KB (whole tile within 200 KB).
    int stage_bytes = 16384;
    while (stage_bytes < max_rowbytes + 32) stage_bytes *= 2;
    B200_REQUIRE(stage_bytes <= 65536, "preprocess: source rows wider than 21834 pixels are not supported");
    auto rows_needed = [&](int tr) { return (int)std::ceil(max_vscale * tr + 4.0 * (max_vscale < 1 ? 1 : max_vscale)) + 3; };
    int TR = 16;
    const size_t budget = 200 * 1024 - 3072 - 2 * (size_t)stage_bytes;
    if (any_resize) {
        while (TR > 1 && ((size_t)rows_needed(TR) * inter_pitch > 65536)) TR /= 2;
        B200_REQUIRE((size_t)rows_needed(TR) * inter_pitch <= budget, "preprocess: vertical downscale factor too large for one tile");
    }
    while (TR > 1 && (size_t)TR * max_ident_rowbytes + 32 > 2 * (size_t)stage_bytes) TR /= 2;
    B200_REQUIRE((size_t)TR * max_ident_rowbytes + 32 <= 2 * (size_t)stage_bytes, "preprocess: source rows too wide");
    const size_t inter_bytes = any_resize ? (size_t)rows_needed(TR) * inter_pitch : 0;
    const size_t smem = 3072 + 2 * (size_t)stage_bytes + inter_bytes;

    B200_CUDA_OK(cudaMemcpyAsync(s->d_desc, s->h_desc, sizeof(ImgDesc) * B, cudaMemcpyHostToDevice, st));
    B200_CUDA_OK(cudaEventRecord(s->desc_free, st));
    PreParams p{src_dev, s->d_desc, s->d_lut, out_dev, S, TR, layout, out_bf16, P, layout ? S / P : 0, kpad, inter_pitch, stage_bytes};
    const int tiles = (S + TR - 1) / TR;
    const bool vw8 = (S % 8 == 0) && (layout == 0 || P % 8 == 0);
    if (vw8) return out_bf16 ? launch_pre<8, true>(p, tiles, B, smem, st) : launch_pre<8, false>(p, tiles, B, smem, st);
    return out_bf16 ? launch_pre<2, true>(p, tiles, B, smem, st) : launch_pre<2, false>(p, tiles, B, smem, st);
}

}  // namespace b200
