// K5 on the 5th-generation tensor cores: softmax(Q K^T / 8) V per (image, head), head_dim 64, bf16 in / bf16 out.
//
// The mma.sync attention in vit_ops.cu is bound by the legacy HMMA rate (32 % of a ViT-B/16 step, 44 % of ViT-L/14@336);
// this kernel issues the two matmuls as tcgen05.mma with accumulators in TMEM.
//
// Tile = 128 query rows x 128 keys.
//   pack == 1 (S > 64): 128 queries of one (image, head); keys are walked in steps of 128 with an online softmax.
//   pack == 2 (S <= 64): two heads of one image share a tile (rows 0..63 head h, rows 64..127 head h+1; same for the
//                        keys); a block-diagonal mask keeps them apart, so short sequences (S = 50) still fill the MMA.
// Per step:   S = Q K^T            UMMA 128x128x16 x4, A = Q tile, B = K tile, both K-major SWIZZLE_128B (TMA)
//             P = exp2(...)         four softmax warps, thread = query row = TMEM lane; P goes to shared memory as the
//                                   K-major A operand of the second matmul (bf16, swizzled like a TMA tile)
//             O_step = P V          UMMA 128x64x16 x8, B = V tile as loaded ([key][dim], dim contiguous) = MN-major
//             O += alpha-rescaled   in registers of the softmax threads (64 fp32 per row)
// One persistent CTA per SM runs TWO such tiles at a time ("slots": own Q/K/V/P shared memory, own TMEM columns, own
// softmax warpgroup), so the tensor core works on one slot while the other slot's softmax runs (the exp2 stream on the
// SFUs is what bounds this kernel, not the MMAs).  warp 0 = TMA producer for both slots, warp 1 = MMA issuer for both,
// warps 2..5 = softmax of slot 0, warps 6..9 = softmax of slot 1.  TMEM: scores at columns 0 / 128, outputs at 256 / 320.
#include "common.cuh"
#include "kernels.h"

#include <cstdlib>

namespace b200 {

constexpr int AT_ROWS = 128, AT_DH = 64;
constexpr int AT_TILE_BYTES = AT_ROWS * AT_DH * 2;            // 16 KB: one 128 x 64 bf16 tile (Q, K or V)
constexpr int AT_P_BYTES = 2 * AT_TILE_BYTES;                 // 128 x 128 bf16 probabilities = two K-blocks of 64 keys
// per slot: Q | K0 V0 | K1 V1 | P
constexpr int AT_OFF_KV = AT_TILE_BYTES, AT_OFF_P = 5 * AT_TILE_BYTES, AT_SLOT_BYTES = 5 * AT_TILE_BYTES + AT_P_BYTES;
constexpr int AT_SMEM = 2 * AT_SLOT_BYTES + 256 + 1024;
constexpr int AT_THREADS = 320;
static_assert(AT_SMEM <= 227 * 1024, "attention_tc: shared memory plan exceeds one SM");

struct AttnParams {
    int S, d, heads, B;
    int pack;            // heads per tile (1 or 2)
    int n_qt, n_kt;      // query tiles per (image, head group), key steps per item
    int items;
    __nv_bfloat16* out;  // [B*S, d]
    DeviceStatus* status;
};

// MN-major B operand (V tile: 128 keys x 64 dims, one 128-byte row per key, SWIZZLE_128B): 8-key groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;                       // leading (MN) byte offset: a single 64-element group, unused
    d |= (uint64_t)(1024 >> 4) << 32;             // stride (K) byte offset between 8-key groups
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// idesc with fp32 accumulate, bf16 operands, A K-major, B K-major (b_mn = 0) or MN-major (b_mn = 1)
__host__ __device__ constexpr uint32_t attn_idesc(int M, int N, int b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// 64 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld_row64(uint32_t taddr, uint32_t (&v)[64]) {
    tmem_ld_32x32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
    tmem_ld_32x32(taddr + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
    tmem_ld_wait();
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

struct AttnItem { int b, h0, qt; };
__device__ __forceinline__ AttnItem attn_item(const AttnParams& p, int item) {
    const int hgroups = p.heads / p.pack, rest = item / p.n_qt;
    return {rest / hgroups, (rest % hgroups) * p.pack, item % p.n_qt};
}

// One softmax step of one row: scores (TMEM) -> running max / sum, probabilities (bf16) into the row's P bytes.
// MASKED = false: all 128 key columns are valid.  MASKED = true: only columns [c_lo, c_hi) inside 64-column chunks [ch0, ch1).
template <bool MASKED>
__device__ __forceinline__ void softmax_row(uint32_t t_s, uint8_t* p_row, int r, float sl2, int c_lo, int c_hi, int ch0, int ch1,
                                            float& m, float& l, float& alpha) {
    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    for (int c = MASKED ? ch0 : 0; c < (MASKED ? ch1 : 2); ++c) {
        uint32_t v[64];
        tmem_ld_row64(t_s + c * 64, v);
#pragma unroll
        for (int j = 0; j < 64; ++j) {
            float x = __uint_as_float(v[j]);
            if (MASKED) { const int col = c * 64 + j; x = (col >= c_lo && col < c_hi) ? x : -INFINITY; }
            mx4[j & 3] = fmaxf(mx4[j & 3], x);
        }
    }
    const float mx = fmaxf(m, fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * sl2);
    alpha = ex2_approx(m - mx);                    // 0 on the first step (m = -inf)
    float s4[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = MASKED ? ch0 : 0; c < (MASKED ? ch1 : 2); ++c) {
        uint32_t v[64];
        tmem_ld_row64(t_s + c * 64, v);
        uint8_t* blk = p_row + c * AT_TILE_BYTES;  // K-block of 64 keys
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float e[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                e[j] = ex2_approx(fmaf(__uint_as_float(v[8 * i + j]), sl2, -mx));
                if (MASKED) { const int col = c * 64 + 8 * i + j; e[j] = (col >= c_lo && col < c_hi) ? e[j] : 0.f; }
                s4[j & 3] += e[j];
            }
            uint4 w;
            w.x = pack_bf16x2(e[0], e[1]); w.y = pack_bf16x2(e[2], e[3]); w.z = pack_bf16x2(e[4], e[5]); w.w = pack_bf16x2(e[6], e[7]);
            *reinterpret_cast<uint4*>(blk + ((i ^ (r & 7)) << 4)) = w;
        }
    }
    l = l * alpha + ((s4[0] + s4[1]) + (s4[2] + s4[3]));
    m = mx;
}

__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * AT_SLOT_BYTES);
    // per slot (stride 10): q_full, q_empty, kv_full[2], kv_empty[2], s_full, p_full, o_full, o_free
    enum { Q_FULL = 0, Q_EMPTY = 1, KV_FULL = 2, KV_EMPTY = 4, S_FULL = 6, P_FULL = 7, O_FULL = 8, O_FREE = 9, NBAR = 10 };
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NBAR);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int R = AT_ROWS / p.pack;                    // tile rows (and keys per step) per head
    const int n_pairs = (p.items + 1) >> 1;            // CTA c works on item pairs c, c + grid, ...; slot = item & 1
    const int my_pairs = blockIdx.x < n_pairs ? (n_pairs - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    // the very last pair may hold a single item (slot 1 empty); it belongs to the CTA that owns pair n_pairs - 1
    const bool odd_tail = (p.items & 1) && my_pairs > 0 && ((n_pairs - 1 - (int)blockIdx.x) % (int)gridDim.x == 0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQKV);
        for (int s = 0; s < 2; ++s) {
            uint64_t* b = bars + s * NBAR;
            for (int i = 0; i < NBAR; ++i) mbar_init(&b[i], (i == P_FULL || i == O_FREE) ? 4 : 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    for (int i = threadIdx.x; i < 2 * (AT_P_BYTES / 16); i += AT_THREADS) {
        const int s = i / (AT_P_BYTES / 16), j = i % (AT_P_BYTES / 16);
        reinterpret_cast<uint4*>(smem + s * AT_SLOT_BYTES + AT_OFF_P)[j] = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async();                               // the zeroed off-diagonal blocks are read by the tensor core
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            bool ok = true;
            for (int li = 0; li < my_pairs && ok; ++li) {
                const int pair = blockIdx.x + li * gridDim.x;
                const int n_slots = (li == my_pairs - 1 && odd_tail) ? 1 : 2;
                for (int s = 0; s < n_slots && ok; ++s) {
                    const AttnItem it = attn_item(p, 2 * pair + s);
                    uint64_t* b = bars + s * NBAR;
                    uint8_t* base = smem + s * AT_SLOT_BYTES;
                    if (!(ok = mbar_wait(&b[Q_EMPTY], (li & 1) ^ 1, p.status, 31))) break;
                    mbar_arrive_expect_tx(&b[Q_FULL], AT_TILE_BYTES);
                    for (int u = 0; u < p.pack; ++u)
                        tma_load_2d(base + u * R * 128, &tmQKV, &b[Q_FULL], (it.h0 + u) * AT_DH, it.b * p.S + it.qt * R);
                }
                for (int kt = 0; kt < p.n_kt && ok; ++kt) {
                    const int g = li * p.n_kt + kt, st = g & 1;
                    for (int s = 0; s < n_slots && ok; ++s) {
                        const AttnItem it = attn_item(p, 2 * pair + s);
                        uint64_t* b = bars + s * NBAR;
                        uint8_t* k_dst = smem + s * AT_SLOT_BYTES + AT_OFF_KV + st * 2 * AT_TILE_BYTES;
                        if (!(ok = mbar_wait(&b[KV_EMPTY + st], ((g >> 1) & 1) ^ 1, p.status, 32))) break;
                        mbar_arrive_expect_tx(&b[KV_FULL + st], 2 * AT_TILE_BYTES);
                        for (int u = 0; u < p.pack; ++u) {
                            const int row = it.b * p.S + kt * R;
                            tma_load_2d(k_dst + u * R * 128, &tmQKV, &b[KV_FULL + st], p.d + (it.h0 + u) * AT_DH, row);
                            tma_load_2d(k_dst + AT_TILE_BYTES + u * R * 128, &tmQKV, &b[KV_FULL + st], 2 * p.d + (it.h0 + u) * AT_DH, row);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = attn_idesc(128, 128, 0), idesc_o = attn_idesc(128, 64, 1);
            const int G0 = my_pairs * p.n_kt, G1 = (my_pairs - (odd_tail ? 1 : 0)) * p.n_kt;   // steps per slot
            bool ok = true;
            // S = Q K^T of step g of a slot; the slot's score columns are free once P of step g - 1 was written
            auto issue_s = [&](int s, int g) -> bool {
                uint64_t* b = bars + s * NBAR;
                uint8_t* base = smem + s * AT_SLOT_BYTES;
                const int kt = g % p.n_kt, st = g & 1;
                if (kt == 0 && !mbar_wait(&b[Q_FULL], (g / p.n_kt) & 1, p.status, 33)) return false;
                if (!mbar_wait(&b[KV_FULL + st], (g >> 1) & 1, p.status, 34)) return false;
                tc_fence_after();
                const uint64_t dq = umma_desc_sw128(smem_u32(base));
                const uint64_t dk = umma_desc_sw128(smem_u32(base + AT_OFF_KV + st * 2 * AT_TILE_BYTES));
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tmem_base + s * 128, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(&b[S_FULL]);
                if (kt == p.n_kt - 1) umma_commit(&b[Q_EMPTY]);
                return true;
            };
            // O_step = P V of step g; needs P in shared memory and the previous O tile read out of TMEM
            auto issue_pv = [&](int s, int g) -> bool {
                uint64_t* b = bars + s * NBAR;
                uint8_t* base = smem + s * AT_SLOT_BYTES;
                const int st = g & 1;
                if (!mbar_wait(&b[P_FULL], g & 1, p.status, 35)) return false;
                if (g > 0 && !mbar_wait(&b[O_FREE], (g - 1) & 1, p.status, 36)) return false;
                tc_fence_after();
                const uint64_t dv = umma_desc_mn_sw128(smem_u32(base + AT_OFF_KV + st * 2 * AT_TILE_BYTES + AT_TILE_BYTES));
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint64_t dp = umma_desc_sw128(smem_u32(base + AT_OFF_P + (k >> 2) * AT_TILE_BYTES)) + 2 * (k & 3);
                    umma_f16(tmem_base + 256 + s * 64, dp, dv + (uint64_t)k * (2048 >> 4), idesc_o, k != 0);
                }
                umma_commit(&b[O_FULL]);
                umma_commit(&b[KV_EMPTY + st]);
                return true;
            };
            for (int s = 0; s < 2 && ok; ++s) if ((s ? G1 : G0) > 0) ok = issue_s(s, 0);
            for (int g = 0; g < G0 && ok; ++g)
                for (int s = 0; s < 2 && ok; ++s) {
                    const int Gs = s ? G1 : G0;
                    if (g >= Gs) continue;
                    ok = issue_pv(s, g);
                    if (ok && g + 1 < Gs) ok = issue_s(s, g + 1);
                }
        }
    } else {
        const int s = (warp - 2) >> 2, q = warp & 3, r = q * 32 + lane;     // slot, TMEM lane quarter, tile row = TMEM lane
        const int hsel = r / R;                        // which head of the tile this row belongs to (0 when pack == 1)
        const float sl2 = 0.125f * 1.4426950408889634f;
        const uint32_t t_s = tmem_base + ((uint32_t)(q * 32) << 16) + s * 128;
        const uint32_t t_o = tmem_base + ((uint32_t)(q * 32) << 16) + 256 + s * 64;
        uint64_t* b = bars + s * NBAR;
        uint8_t* sP = smem + s * AT_SLOT_BYTES + AT_OFF_P;
        uint8_t* p_row = sP + r * 128;
        const int my_items = my_pairs - ((s == 1 && odd_tail) ? 1 : 0);
        int g = 0; bool ok = true;
        for (int li = 0; li < my_items && ok; ++li) {
            const AttnItem it = attn_item(p, 2 * (blockIdx.x + li * gridDim.x) + s);
            float o[64];
#pragma unroll
            for (int c = 0; c < 64; ++c) o[c] = 0.f;
            float m = -INFINITY, l = 0.f;
            for (int kt = 0; kt < p.n_kt; ++kt, ++g) {
                if (!(ok = mbar_wait(&b[S_FULL], g & 1, p.status, 37))) break;
                tc_fence_after();
                float alpha;
                const int left = p.S - kt * 128;
                if (p.pack == 1 && left >= 128) softmax_row<false>(t_s, p_row, r, sl2, 0, 128, 0, 2, m, l, alpha);
                else if (p.pack == 1) softmax_row<true>(t_s, p_row, r, sl2, 0, left, 0, 2, m, l, alpha);
                else softmax_row<true>(t_s, p_row, r, sl2, hsel * 64, hsel * 64 + p.S, hsel, hsel + 1, m, l, alpha);
                fence_proxy_async();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&b[P_FULL]);
                if (!(ok = mbar_wait(&b[O_FULL], g & 1, p.status, 38))) break;
                tc_fence_after();
                {
                    uint32_t v[64];
                    tmem_ld_row64(t_o, v);
#pragma unroll
                    for (int c = 0; c < 64; ++c) o[c] = fmaf(o[c], alpha, __uint_as_float(v[c]));
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&b[O_FREE]);
            }
            if (!ok) break;
            // normalise, stage this row (128 B, swizzled) in its own P block, then write the warp's 32 rows coalesced
            const float inv = 1.f / l;
            uint8_t* st_row = p_row + (p.pack == 2 ? hsel : 0) * AT_TILE_BYTES;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                uint4 w;
                w.x = pack_bf16x2(o[8 * i] * inv, o[8 * i + 1] * inv); w.y = pack_bf16x2(o[8 * i + 2] * inv, o[8 * i + 3] * inv);
                w.z = pack_bf16x2(o[8 * i + 4] * inv, o[8 * i + 5] * inv); w.w = pack_bf16x2(o[8 * i + 6] * inv, o[8 * i + 7] * inv);
                *reinterpret_cast<uint4*>(st_row + ((i ^ (r & 7)) << 4)) = w;
            }
            __syncwarp();
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const int rr = q * 32 + t * 4 + (lane >> 3), ch = lane & 7;
                const int hs = rr / R, qrow = it.qt * R + (rr % R);
                if (qrow < p.S) {
                    const uint4 w = *reinterpret_cast<const uint4*>(sP + rr * 128 + (p.pack == 2 ? hs : 0) * AT_TILE_BYTES + ((ch ^ (rr & 7)) << 4));
                    *reinterpret_cast<uint4*>(p.out + (size_t)(it.b * p.S + qrow) * p.d + (it.h0 + hs) * AT_DH + ch * 8) = w;
                }
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

// B200_ATTN_TC: 0 = always the mma.sync kernel, 1 (default) = this kernel for S > 64, 2 = also for S <= 64 (two heads per tile).
bool attention_tc_supported(int S, int d, int heads) {
    static const int on = [] { const char* e = getenv("B200_ATTN_TC"); return e ? atoi(e) : 1; }();
    if (!on || heads <= 0 || d != heads * AT_DH || S <= 0) return false;
    return S > 64 || (on >= 2 && (heads % 2) == 0);
}

int launch_attention_tc(const void* qkv, void* out, int B, int S, int d, int heads, DeviceStatus* status, cudaStream_t st) {
    if (B <= 0) return B200_OK;
    AttnParams p;
    p.S = S; p.d = d; p.heads = heads; p.B = B;
    p.pack = S <= 64 ? 2 : 1;
    p.n_qt = p.pack == 2 ? 1 : (S + AT_ROWS - 1) / AT_ROWS;
    p.n_kt = p.n_qt;
    p.items = B * (heads / p.pack) * p.n_qt;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.status = status;
    CUtensorMap tm;
    int rc = make_tmap_2d(&tm, qkv, (uint64_t)B * S, (uint64_t)3 * d, (uint64_t)3 * d * 2, AT_ROWS / p.pack, false);
    if (rc) return rc;
    B200_CUDA_OK(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
    const int n_pairs = (p.items + 1) / 2;
    const int grid = n_pairs < kNumSMs ? n_pairs : kNumSMs;
    attention_tc_kernel<<<grid, AT_THREADS, AT_SMEM, st>>>(tm, p);
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

}  // namespace b200
