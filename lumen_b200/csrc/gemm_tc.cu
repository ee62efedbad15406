// K2/K4/K6/K7/K8/K9 — one persistent, warp-specialised tcgen05 GEMM with fused epilogues.
//
//   C[M,N] = epilogue( A[M,K] (bf16, K-major) x W[N,K]^T (bf16, K-major = torch Linear layout) )
//
// Replaces the encoder matmuls the reference reaches through onnxruntime / torch
// (packages/lumen-clip/src/lumen_clip/backends/onnxrt_backend.py:485, torch_backend.py:652-666).
//
// Layout per CTA (320 threads, one CTA per SM, grid = min(#tiles, 148)):
//   warp 0      TMA producer: 128x64 A tile + BNx64 W tile per stage, SWIZZLE_128B, mbarrier complete_tx
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128 x BN x 16, fp32 accumulate in TMEM)
//   warps 2..9  epilogue: tcgen05.ld -> bias / activation / convert -> swizzled 128-byte boxes in smem -> TMA store
//               (residual epilogue: TMA reduce-add)
// TMEM holds two BN-column accumulators so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "common.cuh"
#include "kernels.h"

#include <cstdlib>
#include <mutex>

namespace b200 {

enum : int { ACT_QUICK_GELU = 0, ACT_GELU = 1 };

struct GemmParams {
    int M, N, K;          // K is a multiple of 64 (operands are zero-padded in K)
    const float* bias;    // [N] or nullptr
    void* out;            // bf16 or fp32, row pitch ldc elements
    int ldc;
    int epi, act;
    DeviceStatus* status;
    int reverse;          // walk the tiles from the last row block to the first (see gemm_bf16_tc)
};

constexpr int BM = 128, BK = 64;
constexpr int EPI_WARPS = 8;                  // two warps per TMEM lane quarter, alternating 16-column sub-chunks
constexpr int GEMM_THREADS = 64 + 32 * EPI_WARPS;

template <int BN> struct GemmCfg {
    static constexpr int STAGES = BN == 256 ? 4 : (BN == 192 ? 4 : 6);
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = BN == 64 ? 128 : BN == 128 ? 256 : 512;   // two accumulators, allocation must be a power of two
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * 16384 + 256 + 1024;
};

template <int ACT> __device__ __forceinline__ float apply_act(float v) {
    if (ACT == ACT_QUICK_GELU) {       // x * sigmoid(1.702 x) = x * (0.5 + 0.5 tanh(0.851 x)): one MUFU op
        float t;
        asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.851f * v));
        return v * fmaf(0.5f, t, 0.5f);
    }
    return 0.5f * v * (1.0f + erff(v * 0.70710678118654752f));
}
__device__ __forceinline__ void red_add_v4(float* addr, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ uint2 pack4_bf16(float a, float b, float c, float d) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    uint2 r; r.x = *reinterpret_cast<uint32_t*>(&lo); r.y = *reinterpret_cast<uint32_t*>(&hi);
    return r;
}

// Epilogue of one accumulator (128 rows x BN columns in this CTA's TMEM), executed by all eight epilogue warps.
// The tile leaves in boxes of 128 rows x 128 bytes (64 bf16 or 32 fp32 columns): a thread owns one row (its TMEM
// lane) and, with the warp that shares its lane quarter, half of the box's columns.  It adds the bias, applies the
// activation, converts, and writes four 16-byte chunks into a shared-memory slot laid out exactly as the output
// tensor map's SWIZZLE_128B box (chunk j of row r at r*128 + ((j ^ (r & 7)) << 4): conflict-free).  One thread then
// hands the slot to the TMA engine: a plain tensor store, or for the residual epilogue a tensor REDUCE-ADD
// (x += acc + bias performed by the memory system, one add per element, so still deterministic).  Global memory is
// therefore written in full 128-byte lines by the copy engine instead of 32-byte pieces by the LSU, which is what
// bounded every K = 768 GEMM before (62 us floor for the 118 MB QKV output).  Two slots alternate; `box_seq` counts
// boxes across tiles; `release()` returns the accumulator to the MMA warp once this warp has read its share.
constexpr int EPI_SLOT_BYTES = 128 * 128;
constexpr int EPI_THREADS = 32 * EPI_WARPS;

template <int BN, int EPI, int ACT, class Release>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, const CUtensorMap* tmC, uint8_t* slots, int& box_seq, int q, int half,
                                              int lane, bool issuer, uint32_t tsrc, int m0, int n0, Release&& release) {
    constexpr bool OUT_F32 = (EPI == EPI_F32 || EPI == EPI_BIAS_RES_F32);
    constexpr int BOXC = OUT_F32 ? 32 : 64;            // columns per 128-byte box
    constexpr int NV = BOXC / 2;                         // columns per thread and box
    const int nbox = min(BN / BOXC, (p.N - n0 + BOXC - 1) / BOXC);
    const int r = q * 32 + lane;                         // row inside the tile = TMEM lane
#pragma unroll 1
    for (int ob = 0; ob < nbox; ++ob, ++box_seq) {
        uint8_t* slot = slots + (box_seq & 1) * EPI_SLOT_BYTES;
        const int col0 = n0 + ob * BOXC + half * NV;
        uint32_t v[NV];
        if (NV == 32) tmem_ld_32x32(tsrc + ob * BOXC + half * NV, reinterpret_cast<uint32_t(&)[32]>(v));
        else tmem_ld_32x16(tsrc + ob * BOXC + half * NV, reinterpret_cast<uint32_t(&)[16]>(v));
        float4 b4[NV / 4];
#pragma unroll
        for (int i = 0; i < NV / 4; ++i) {
            b4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.bias != nullptr && col0 + 4 * i < p.N) b4[i] = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 4 * i));
        }
        if (issuer) bulk_wait_read<1>();                 // the store that used this slot two boxes ago has read it
        named_bar_sync(1, EPI_THREADS);
        tmem_ld_wait();
        if (ob == nbox - 1) release();
        float f[NV];
#pragma unroll
        for (int i = 0; i < NV / 4; ++i) {
            f[4 * i] = __uint_as_float(v[4 * i]) + b4[i].x; f[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + b4[i].y;
            f[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + b4[i].z; f[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + b4[i].w;
        }
        if (EPI == EPI_BIAS_ACT_BF16) {
#pragma unroll
            for (int i = 0; i < NV; ++i) f[i] = apply_act<ACT>(f[i]);
        }
        uint8_t* rowp = slot + r * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            uint4 c;
            if (OUT_F32) {
                c = make_uint4(__float_as_uint(f[4 * i]), __float_as_uint(f[4 * i + 1]), __float_as_uint(f[4 * i + 2]), __float_as_uint(f[4 * i + 3]));
            } else {
                const uint2 lo = pack4_bf16(f[8 * i], f[8 * i + 1], f[8 * i + 2], f[8 * i + 3]);
                const uint2 hi = pack4_bf16(f[8 * i + 4], f[8 * i + 5], f[8 * i + 6], f[8 * i + 7]);
                c = make_uint4(lo.x, lo.y, hi.x, hi.y);
            }
            *reinterpret_cast<uint4*>(rowp + (((half * 4 + i) ^ (r & 7)) << 4)) = c;
        }
        fence_proxy_async();                             // generic-proxy writes -> visible to the TMA (async proxy)
        named_bar_sync(2, EPI_THREADS);
        if (issuer) {
            if (EPI == EPI_BIAS_RES_F32) tma_reduce_add_2d(tmC, slot, n0 + ob * BOXC, m0);
            else tma_store_2d(tmC, slot, n0 + ob * BOXC, m0);
            bulk_commit();
        }
    }
}

template <int BN, int EPI, int ACT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmC, GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* epi_slots = smem + STAGES * Cfg::STAGE_BYTES;    // 2 x 16 KB output boxes, 1024-byte aligned
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES + 2 * 16384);
    uint64_t* full = bars;                  // [STAGES]  TMA -> MMA
    uint64_t* empty = bars + STAGES;        // [STAGES]  MMA -> TMA
    uint64_t* acc_full = bars + 2 * STAGES; // [2]       MMA -> epilogue
    uint64_t* acc_empty = acc_full + 2;     // [2]       epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (p.N + BN - 1) / BN;
    const int tiles = ((p.M + BM - 1) / BM) * tiles_n;
    const int kblocks = p.K / BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmC);
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], EPI_WARPS); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                              // everything above overlapped the previous kernel's tail; operands and outputs are touched below
    pdl_trigger();

    if (warp == 0) {
        if (lane == 0) {
            int s = 0; uint32_t ph = 0; bool ok = true;
            for (int t = blockIdx.x; t < tiles && ok; t += gridDim.x) {
                const int tt = p.reverse ? tiles - 1 - t : t;
                const int m0 = (tt / tiles_n) * BM, n0 = (tt % tiles_n) * BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (!(ok = mbar_wait(&empty[s], ph ^ 1, p.status, 1))) break;
                    uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
                    mbar_arrive_expect_tx(&full[s], Cfg::STAGE_BYTES);
                    tma_load_2d(sa, &tmA, &full[s], kb * BK, m0);
                    tma_load_2d(sa + Cfg::A_BYTES, &tmB, &full[s], kb * BK, n0);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc(BM, BN, /*bf16*/ 1);
            int s = 0; uint32_t ph = 0; int it = 0; bool ok = true;
            for (int t = blockIdx.x; t < tiles && ok; t += gridDim.x, ++it) {
                const int as = it & 1; const uint32_t aph = (it >> 1) & 1;
                if (!(ok = mbar_wait(&acc_empty[as], aph ^ 1, p.status, 2))) break;
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (!(ok = mbar_wait(&full[s], ph, p.status, 3))) break;
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + Cfg::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)   // +32 bytes along K inside the 128-byte swizzled row
                        umma_f16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit(&empty[s]);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
                if (ok) umma_commit(&acc_full[as]);
            }
        }
    } else {
        const int q = warp & 3;                       // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;             // which 16-column sub-chunks (even / odd) this warp stores
        const bool issuer = threadIdx.x == 64;         // first epilogue thread issues the TMA stores
        int it = 0, box_seq = 0; bool ok = true;
        for (int t = blockIdx.x; t < tiles && ok; t += gridDim.x, ++it) {
            const int tt = p.reverse ? tiles - 1 - t : t;
            const int m0 = (tt / tiles_n) * BM, n0 = (tt % tiles_n) * BN;
            const int as = it & 1; const uint32_t aph = (it >> 1) & 1;
            if (!(ok = mbar_wait(&acc_full[as], aph, p.status, 4))) break;
            tc_fence_after();
            epilogue_tile<BN, EPI, ACT>(p, &tmC, epi_slots, box_seq, q, half, lane, issuer,
                                        tmem_base + ((uint32_t)(q * 32) << 16) + as * BN, m0, n0, [&] {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[as]);
            });
        }
        if (issuer) bulk_wait<0>();                    // every output box has left shared memory and is written
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { __syncwarp(); tmem_dealloc(tmem_base, Cfg::TMEM_COLS); }
}

// ------------------------------------------------------------------------------------------ CTA-pair variant
// Same pipeline on a cluster of two CTAs (tcgen05 cta_group::2): one UMMA covers 256 rows x BN columns, CTA rank r
// holds rows m0 + 128 r of A and HALF of the W tile (BN/2 rows), so every SM reads 64 B/clk of operands from
// shared memory instead of 96 and a stage costs 32 KB instead of 48 (six stages).  The leader (rank 0) issues the
// MMAs; its commits arrive on the stage / accumulator barriers of both CTAs; both CTAs' TMA loads credit the
// leader's "full" barrier; the epilogue warps of both CTAs hand accumulators back on the leader's barrier.
template <int BN> struct Gemm2Cfg {
    static constexpr int STAGES = 6;
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = (BN / 2) * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * 16384 + 256 + 1024;
};

template <int BN, int EPI, int ACT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, GemmParams p) {
    using Cfg = Gemm2Cfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* epi_slots = smem + STAGES * Cfg::STAGE_BYTES;    // 2 x 16 KB output boxes, 1024-byte aligned
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES + 2 * 16384);
    uint64_t* full = bars;                  // [STAGES]  used in the leader only
    uint64_t* empty = bars + STAGES;        // [STAGES]  one per CTA, signalled by the leader's multicast commit
    uint64_t* acc_full = bars + 2 * STAGES; // [2]       one per CTA, multicast commit
    uint64_t* acc_empty = acc_full + 2;     // [2]       leader only: 2 x EPI_WARPS arrivals
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int tiles_n = (p.N + BN - 1) / BN;
    const int tiles = ((p.M + 2 * BM - 1) / (2 * BM)) * tiles_n;
    const int kblocks = p.K / BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmC);
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 2 * EPI_WARPS); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS);
    tc_fence_before();
    cluster_sync_all();                      // barriers of both CTAs initialised before any remote arrive / TMA credit
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                              // everything above overlapped the previous kernel's tail; operands and outputs are touched below
    pdl_trigger();

    if (warp == 0) {
        if (lane == 0) {
            int s = 0; uint32_t ph = 0; bool ok = true;
            for (int t = pair; t < tiles && ok; t += npairs) {
                const int tt = p.reverse ? tiles - 1 - t : t;
                const int m0 = (tt / tiles_n) * 2 * BM + (int)rank * BM, n0 = (tt % tiles_n) * BN + (int)rank * (BN / 2);
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (!(ok = mbar_wait(&empty[s], ph ^ 1, p.status, 21))) break;
                    uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
                    if (rank == 0) mbar_arrive_expect_tx(&full[s], 2 * Cfg::STAGE_BYTES);
                    tma_load_2d_pair(sa, &tmA, &full[s], kb * BK, m0);
                    tma_load_2d_pair(sa + Cfg::A_BYTES, &tmB, &full[s], kb * BK, n0);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0 && lane == 0) {
            constexpr uint32_t idesc = umma_idesc(2 * BM, BN, /*bf16*/ 1);
            int s = 0; uint32_t ph = 0; int it = 0; bool ok = true;
            for (int t = pair; t < tiles && ok; t += npairs, ++it) {
                const int as = it & 1; const uint32_t aph = (it >> 1) & 1;
                if (!(ok = mbar_wait(&acc_empty[as], aph ^ 1, p.status, 22))) break;
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (!(ok = mbar_wait(&full[s], ph, p.status, 23))) break;
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + Cfg::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) umma_f16_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit_pair(&empty[s]);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
                if (ok) umma_commit_pair(&acc_full[as]);
            }
        }
    } else {
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const bool issuer = threadIdx.x == 64;
        const uint32_t leader_acc_empty0 = mapa_u32(smem_u32(&acc_empty[0]), 0);
        int it = 0, box_seq = 0; bool ok = true;
        for (int t = pair; t < tiles && ok; t += npairs, ++it) {
            const int tt = p.reverse ? tiles - 1 - t : t;
            const int m0 = (tt / tiles_n) * 2 * BM + (int)rank * BM, n0 = (tt % tiles_n) * BN;
            const int as = it & 1; const uint32_t aph = (it >> 1) & 1;
            if (!(ok = mbar_wait(&acc_full[as], aph, p.status, 24))) break;
            tc_fence_after();
            epilogue_tile<BN, EPI, ACT>(p, &tmC, epi_slots, box_seq, q, half, lane, issuer,
                                        tmem_base + ((uint32_t)(q * 32) << 16) + as * BN, m0, n0, [&] {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(leader_acc_empty0 + as * 8);
            });
        }
        if (issuer) bulk_wait<0>();
    }
    tc_fence_before();
    cluster_sync_all();                      // neither CTA may exit (or free TMEM) while its partner can still touch it
    if (warp == 1) { __syncwarp(); tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS); }
}

// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t row_pitch_bytes,
                 uint32_t box_rows, bool fp16) {
    return make_tmap_2d_ex(out, base, rows, cols, row_pitch_bytes, box_rows, fp16 ? 1 : 0);
}

// dtype: 0 bf16, 1 fp16, 2 fp32.  The box is always 128 bytes wide (64 or 32 elements) with SWIZZLE_128B.
int make_tmap_2d_ex(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t row_pitch_bytes,
                    uint32_t box_rows, int dtype) {
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    });
    if (!g_encode) return fail(B200_EDEVICE, "cuTensorMapEncodeTiled is not available from the CUDA driver");
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (row_pitch_bytes & 15))
        return fail(B200_EINVAL, "TMA operands need 16-byte aligned base and row pitch");
    // A pass re-encodes the same ~200 descriptors (fixed activation buffers and weights); at small batch the encoder calls were a
    // third of the host time per pass, and the host is what bounds that regime.  Encoding is a pure function of the key.
    struct Key {
        const void* base; uint64_t rows, cols, pitch; uint32_t box_rows; int dtype;
        bool operator==(const Key& o) const { return base == o.base && rows == o.rows && cols == o.cols && pitch == o.pitch && box_rows == o.box_rows && dtype == o.dtype; }
    };
    struct Slot { Key k; CUtensorMap m; bool used = false; };
    static thread_local Slot cache[1024];
    const Key key{base, rows, cols, row_pitch_bytes, box_rows, dtype};
    uint64_t hsh = reinterpret_cast<uintptr_t>(base) * 0x9E3779B97F4A7C15ull ^ (rows * 0xC2B2AE3D27D4EB4Full) ^ (cols << 20) ^ (row_pitch_bytes << 7) ^ box_rows ^ ((uint64_t)dtype << 40);
    Slot& slot = cache[(hsh >> 32) & 1023];
    if (slot.used && slot.k == key) { *out = slot.m; return B200_OK; }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {row_pitch_bytes};
    cuuint32_t box[2] = {dtype == 2 ? 32u : 64u, box_rows};
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = dtype == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : dtype == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    CUresult r = g_encode(out, dt, 2,
                          const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(B200_ECUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    slot.k = key; slot.m = *out; slot.used = true;
    return B200_OK;
}

template <int BN, int EPI, int ACT> static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    static bool attr_set[64] = {false};
    int dev = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev]) {
        B200_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        if (dev < 64) attr_set[dev] = true;
    }
    const int tiles = ((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN);
    const int grid = tiles < kNumSMs ? tiles : kNumSMs;
    B200_CUDA_OK(launch_pdl(gemm_tc_kernel<BN, EPI, ACT>, dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, st, ta, tb, tc, p));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
template <int BN, int EPI, int ACT> static int launch_gemm2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p, cudaStream_t st) {
    using Cfg = Gemm2Cfg<BN>;
    static bool attr_set[64] = {false};
    int dev = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev]) {
        B200_CUDA_OK(cudaFuncSetAttribute(gemm_tc2_kernel<BN, EPI, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        if (dev < 64) attr_set[dev] = true;
    }
    const int tiles = ((p.M + 2 * BM - 1) / (2 * BM)) * ((p.N + BN - 1) / BN);
    const int pairs = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
    B200_CUDA_OK(launch_pdl(gemm_tc2_kernel<BN, EPI, ACT>, dim3(2 * pairs), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, st, ta, tb, tc, p));   // cluster dims are compiled in
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
static int dispatch_gemm2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p, cudaStream_t st) {
    switch (p.epi) {
        case EPI_F32: return launch_gemm2<256, EPI_F32, 0>(ta, tb, tc, p, st);
        case EPI_BIAS_BF16: return launch_gemm2<256, EPI_BIAS_BF16, 0>(ta, tb, tc, p, st);
        case EPI_BIAS_RES_F32: return launch_gemm2<256, EPI_BIAS_RES_F32, 0>(ta, tb, tc, p, st);
        default: return p.act == ACT_QUICK_GELU ? launch_gemm2<256, EPI_BIAS_ACT_BF16, ACT_QUICK_GELU>(ta, tb, tc, p, st)
                                                : launch_gemm2<256, EPI_BIAS_ACT_BF16, ACT_GELU>(ta, tb, tc, p, st);
    }
}
template <int BN> static int dispatch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p, cudaStream_t st) {
    switch (p.epi) {
        case EPI_F32: return launch_gemm<BN, EPI_F32, 0>(ta, tb, tc, p, st);
        case EPI_BIAS_BF16: return launch_gemm<BN, EPI_BIAS_BF16, 0>(ta, tb, tc, p, st);
        case EPI_BIAS_RES_F32: return launch_gemm<BN, EPI_BIAS_RES_F32, 0>(ta, tb, tc, p, st);
        default: return p.act == ACT_QUICK_GELU ? launch_gemm<BN, EPI_BIAS_ACT_BF16, ACT_QUICK_GELU>(ta, tb, tc, p, st)
                                                : launch_gemm<BN, EPI_BIAS_ACT_BF16, ACT_GELU>(ta, tb, tc, p, st);
    }
}

// A: bf16 [M, lda] (lda >= K, K % 64 == 0, zero padded), W: bf16 [N, ldw].
int gemm_bf16_tc(const void* A, int lda, const void* W, int ldw, int M, int N, int K, const float* bias, void* out,
                 int ldc, int epi, int act, DeviceStatus* status, cudaStream_t st, int reverse) {
    B200_REQUIRE(M > 0 && N > 0 && K > 0 && K % BK == 0, "gemm: K must be a positive multiple of 64 (pad the operands)");
    B200_REQUIRE(N % 4 == 0 && ldc % 4 == 0, "gemm: N and the output pitch must be multiples of 4");
    // 128x256 tiles read 96 B/clk of shared memory per MMA against 128 B/clk for 128x128, so prefer them unless
    // they waste clearly more tile-rounds on the 148 SMs.
    auto rounds = [&](int bn) {
        long t = (long)((M + BM - 1) / BM) * ((N + bn - 1) / bn);
        return (double)((t + kNumSMs - 1) / kNumSMs) * bn;
    };
    int bn = 256;
    // one or two row tiles (single-image requests, small micro-batches): the pass is a chain of short weight-streaming kernels, and
    // how many SMs pull on the weights is what counts: 64-wide tiles put four times as many CTAs on them as 256-wide ones
    static const int small_bn = [] { const char* e = getenv("B200_GEMM_SMALL_BN"); return e ? atoi(e) : 64; }();
    if (M <= 2 * BM && small_bn == 64 && N % 64 == 0) bn = 64;
    else if (N <= 128) bn = 128;
    else {
        // measured on B200: 192- and 128-wide tiles only pay off when they save well over 5 % of the tile rounds
        const double c256 = rounds(256), c192 = rounds(192) * 1.08, c128 = rounds(128) * 1.15;
        if (c192 < c256 && c192 <= c128) bn = 192;
        else if (c128 < c256 && c128 < c192) bn = 128;
    }
    CUtensorMap ta, tb, tc;
    const bool out_f32 = epi == EPI_F32 || epi == EPI_BIAS_RES_F32;
    int rcc = make_tmap_2d_ex(&tc, out, M, N, (uint64_t)ldc * (out_f32 ? 4 : 2), BM, out_f32 ? 2 : 0);   // 128-row x 128-byte output boxes
    if (rcc) return rcc;
    static const int force = [] { const char* e = getenv("B200_GEMM_PAIR"); return e ? atoi(e) : -1; }();   // -1 auto, 0 never, 1 always
    // measured on B200 (M = 25,600): the CTA pair wins on every encoder shape except the small out-proj (N = K = 768)
    const bool pair = force == 1 || (force != 0 && M >= 2048 && N >= 256 && !(N <= 768 && K <= 768));
    if (pair) {
        GemmParams p2{M, N, K, bias, out, ldc, epi, act, status, reverse};
        int rc2 = make_tmap_2d(&ta, A, M, K, (uint64_t)lda * 2, BM, false);
        if (rc2) return rc2;
        rc2 = make_tmap_2d(&tb, W, N, K, (uint64_t)ldw * 2, 128, false);      // each CTA loads half of the 256-wide W tile
        if (rc2) return rc2;
        return dispatch_gemm2(ta, tb, tc, p2, st);
    }
    int rc = make_tmap_2d(&ta, A, M, K, (uint64_t)lda * 2, BM, false);
    if (rc) return rc;
    rc = make_tmap_2d(&tb, W, N, K, (uint64_t)ldw * 2, bn, false);
    if (rc) return rc;
    GemmParams p{M, N, K, bias, out, ldc, epi, act, status, reverse};
    return bn == 64 ? dispatch_gemm<64>(ta, tb, tc, p, st) : bn == 128 ? dispatch_gemm<128>(ta, tb, tc, p, st) :
           bn == 192 ? dispatch_gemm<192>(ta, tb, tc, p, st) : dispatch_gemm<256>(ta, tb, tc, p, st);
}

}  // namespace b200
