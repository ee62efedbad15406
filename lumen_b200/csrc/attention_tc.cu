// K5 on the 5th-generation tensor cores: softmax(Q K^T / 8) V per (image, head), head_dim 64, bf16 in / bf16 out.
//
// Both matmuls are tcgen05.mma with accumulators in TMEM.  Two kernels: attention_tc_kernel (this one) for S > 64, and
// attention_short_kernel (further down) for S <= 64.
//
// Tile = 128 query rows of one (image, head), walked over the keys in steps of 64 with an online softmax.
// Per step g: S[g&1] = Q K^T     UMMA 128x64x16 x4 into one of TWO score buffers in TMEM: the tensor core computes the scores
//                                of step g + 1 while the softmax warps work on step g
//             P[g&1] = exp2(..)  softmax warps, thread = query row = TMEM lane; P goes back to TMEM (bf16 pairs, 32 columns,
//                                double buffered) and is the A operand of the second matmul straight from there
//             O (+)= P V         UMMA 128x64x16 x4, B = V tile as loaded ([key][dim], dim contiguous) = MN-major
//             O accumulates in TMEM across the key steps.  The softmax reference maximum is only moved when a row's maximum
//             grows by more than 2^8 (probabilities stay <= 256, exact in bf16/fp32 terms); then the row's O is rescaled in
//             place (tcgen05.ld -> multiply -> tcgen05.st).  After the first key steps this is rare.
// One persistent CTA per SM runs TWO such tiles at a time ("slots": own Q/K/V shared memory, own 256 TMEM columns, own
// warps), independent pipelines that share the SM: warps 0 / 10 = TMA producers, warps 1 / 11 = MMA issuers,
// warps 2..5 / 6..9 = softmax (slot 0 / 1).  TMEM per slot: S0 0..63, S1 64..127, P0 128..159, P1 160..191, O 192..255.
// The exp2 stream on the SFUs (16 per clock and SM) is what bounds this kernel, not the MMAs, and a scheduler only sees
// two softmax warps (one per slot).  So the softmax thread is software-pipelined: the TMEM loads of step g + 1 (ready
// early, thanks to the second score buffer) and its row maximum are interleaved with the exponentials of step g, and the
// SFU pipe always has work queued while loads are in flight.
#include "common.cuh"
#include "kernels.h"

#include <cstdlib>

namespace b200 {

constexpr int AT_ROWS = 128, AT_DH = 64, AT_KEYS = 64;        // query rows per tile, head dim, keys per step
constexpr int AT_Q_BYTES = AT_ROWS * AT_DH * 2;               // 16 KB
constexpr int AT_KV_BYTES = AT_KEYS * AT_DH * 2;              // 8 KB: K or V tile of one step
constexpr int AT_STAGES = 4;                                  // K/V ring depth per slot
// per slot: Q0 | Q1 | (K V) x AT_STAGES.  The Q buffer of a finished item doubles as its output staging area.
constexpr int AT_OFF_KV = 2 * AT_Q_BYTES, AT_SLOT_BYTES = 2 * AT_Q_BYTES + AT_STAGES * 2 * AT_KV_BYTES;
constexpr int AT_TM_S = 0, AT_TM_P = 128, AT_TM_O = 192, AT_TM_SLOT = 256;   // TMEM columns inside a slot
constexpr int AT_BAR_BYTES = 512;
constexpr int AT_SMEM = 2 * AT_SLOT_BYTES + AT_BAR_BYTES + 1024;
constexpr int AT_THREADS = 384;
static_assert(AT_SMEM <= 227 * 1024, "attention_tc: shared memory plan exceeds one SM");

struct AttnParams {
    int S, d, heads, B;
    int n_qt, n_kt;      // query tiles per (image, head), key steps per item
    int items;
    int causal;          // 1: key j is visible to query i only when j <= i (text tower)
    __nv_bfloat16* out;  // [B*S, d]
    DeviceStatus* status;
};

// MN-major B operand (V tile: 64 keys x 64 dims, one 128-byte row per key, SWIZZLE_128B): 8-key groups 1024 B apart.
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
    const int rest = item / p.n_qt;
    return {rest / p.heads, rest % p.heads, item % p.n_qt};
}

// Bounded mbarrier wait with a suspend-time hint: the thread parks in hardware instead of spinning through the issue slots
// the softmax warps need.
__device__ __forceinline__ bool mbar_wait_parked(uint64_t* bar, uint32_t parity, DeviceStatus* st, int where) {
#pragma unroll 1
    for (uint32_t it = 0; it < (1u << 16); ++it) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(100000u) : "memory");
        if (ok) return true;
    }
    if (st) { st->code = 1; st->where = where; }
    return false;
}
// Bounded spinning wait for the latency-critical hand-offs (MMA issuer, scores of the next step)
__device__ __forceinline__ bool mbar_wait_spin(uint64_t* bar, uint32_t parity, DeviceStatus* st, int where) {
#pragma unroll 1
    for (uint32_t it = 0; it < (1u << 26); ++it)
        if (mbar_try_wait(bar, parity)) return true;
    if (st) { st->code = 1; st->where = where; }
    return false;
}
// 32 registers per thread <-> 32 consecutive fp32 columns of the thread's TMEM lane
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        :: "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
           "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
           "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
           "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}
// 16 registers per thread -> 16 consecutive 32-bit columns (32 bf16 probabilities) of the thread's TMEM lane
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        :: "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
           "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]: A = 128 lanes x 16 bf16 (8 columns of bf16 pairs), lane = row
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32_nowait(uint32_t taddr, uint32_t (&v)[32]) { tmem_ld_32x32(taddr, v); }
__device__ __forceinline__ void mask_tail(uint32_t (&v)[32], int lim) {     // keep columns [0, lim) of this 32-column chunk
    if (lim < 32) {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j >= lim) v[j] = 0xff800000u;       // -inf
    }
}

// 12 warps = 3 on each of the SM's four register files (16 K registers each): 168 registers per thread is the ceiling
__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * AT_SLOT_BYTES);
    // per slot: q_full[2], q_empty[2], kv_full[STAGES], kv_empty[STAGES], s_full[2], p_full[2], o_full[2] (P V of a step done), o_free (per item).
    // p_full is split by step parity: the scores of step g + 1 are ready early, so a fast warp may finish step g + 1 before a slow
    // one has arrived for step g; it cannot reach step g + 2 (those scores need P of step g complete).  o_full is split the same way
    // because the softmax warps do not wait for every P V: S_FULL(g) only proves P V(g - 2) done, so a single barrier could still
    // be one phase behind the one a parity wait assumes.
    enum { Q_FULL = 0, Q_EMPTY = 2, KV_FULL = 4, KV_EMPTY = 4 + AT_STAGES, S_FULL = 4 + 2 * AT_STAGES, P_FULL = S_FULL + 2,
           O_FULL = P_FULL + 2, O_FREE = O_FULL + 2, NBAR = O_FREE + 1 };
    static_assert(2 * NBAR * 8 + 8 <= AT_BAR_BYTES, "attention_tc: barrier area too small");
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NBAR);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_pairs = (p.items + 1) >> 1;            // CTA c works on item pairs c, c + grid, ...; slot = item & 1
    const int my_pairs = blockIdx.x < n_pairs ? (n_pairs - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    // the very last pair may hold a single item (slot 1 empty); it belongs to the CTA that owns pair n_pairs - 1
    const bool odd_tail = (p.items & 1) && my_pairs > 0 && ((n_pairs - 1 - (int)blockIdx.x) % (int)gridDim.x == 0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQKV);
        for (int s = 0; s < 2; ++s) {
            uint64_t* b = bars + s * NBAR;
            for (int i = 0; i < NBAR; ++i) mbar_init(&b[i], (i == Q_EMPTY || i == Q_EMPTY + 1 || i == P_FULL || i == P_FULL + 1 || i == O_FREE) ? 4 : 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_trigger();

    if (warp == 0 || warp == 10) {
        // ---- TMA producer of one slot: Q (two 64-row boxes) once per item, K and V of a step through the ring
        if (lane == 0) {
            const int s = warp == 0 ? 0 : 1;
            const int my_items = my_pairs - ((s == 1 && odd_tail) ? 1 : 0);
            uint64_t* b = bars + s * NBAR;
            uint8_t* base = smem + s * AT_SLOT_BYTES;
            bool ok = true;
            for (int li = 0, g = 0; li < my_items && ok; ++li) {
                const AttnItem it = attn_item(p, 2 * (blockIdx.x + li * gridDim.x) + s);
                const int qi = li & 1, row0 = it.b * p.S;
                if (!(ok = mbar_wait_parked(&b[Q_EMPTY + qi], ((li >> 1) & 1) ^ 1, p.status, 31))) break;
                mbar_arrive_expect_tx(&b[Q_FULL + qi], AT_Q_BYTES);
                for (int u = 0; u < 2; ++u)
                    tma_load_2d(base + qi * AT_Q_BYTES + u * 8192, &tmQKV, &b[Q_FULL + qi], it.h0 * AT_DH, row0 + it.qt * AT_ROWS + u * 64);
                for (int kt = 0; kt < p.n_kt; ++kt, ++g) {
                    const int st = g % AT_STAGES;
                    const int row = row0 + kt * AT_KEYS, col = it.h0 * AT_DH;
                    uint8_t* k_dst = base + AT_OFF_KV + st * 2 * AT_KV_BYTES;
                    if (!(ok = mbar_wait_parked(&b[KV_EMPTY + st], ((g / AT_STAGES) & 1) ^ 1, p.status, 32))) break;
                    mbar_arrive_expect_tx(&b[KV_FULL + st], 2 * AT_KV_BYTES);
                    tma_load_2d(k_dst, &tmQKV, &b[KV_FULL + st], p.d + col, row);
                    tma_load_2d(k_dst + AT_KV_BYTES, &tmQKV, &b[KV_FULL + st], 2 * p.d + col, row);
                }
            }
        }
    } else if (warp == 1 || warp == 11) {
        // ---- MMA issuer of one slot
        if (lane == 0) {
            const int s = warp == 1 ? 0 : 1;
            constexpr uint32_t idesc_o = attn_idesc(128, 64, 1);
            const int G = (my_pairs - ((s == 1 && odd_tail) ? 1 : 0)) * p.n_kt;     // steps of this slot
            uint64_t* b = bars + s * NBAR;
            uint8_t* base = smem + s * AT_SLOT_BYTES;
            const uint32_t tm = tmem_base + s * AT_TM_SLOT;
            bool ok = true;
            // keys a step really holds, rounded up to the MMA's K granule
            auto keys16 = [&](int g) {
                const int left = p.S - (g % p.n_kt) * AT_KEYS;
                return left >= AT_KEYS ? AT_KEYS : ((left + 15) & ~15);
            };
            // S[g & 1] = Q K^T of step g; that score buffer is free once P of step g - 2 was written
            auto issue_s = [&](int g) -> bool {
                const int kt = g % p.n_kt, item = g / p.n_kt, qi = item & 1, st = g % AT_STAGES;
                if (kt == 0 && !mbar_wait_parked(&b[Q_FULL + qi], (item >> 1) & 1, p.status, 33)) return false;
                if (!mbar_wait_parked(&b[KV_FULL + st], (g / AT_STAGES) & 1, p.status, 34)) return false;
                tc_fence_after();
                const uint64_t dq = umma_desc_sw128(smem_u32(base + qi * AT_Q_BYTES));
                const uint64_t dk = umma_desc_sw128(smem_u32(base + AT_OFF_KV + st * 2 * AT_KV_BYTES));
                const uint32_t idesc_s = attn_idesc(128, keys16(g), 0);
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tm + AT_TM_S + (g & 1) * 64, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(&b[S_FULL + (g & 1)]);
                return true;
            };
            // O (+)= P[g & 1] V of step g, P read from TMEM; the first step of an item overwrites O, so the previous item's O must have been read out
            auto issue_pv = [&](int g) -> bool {
                const int st = g % AT_STAGES, nk = keys16(g) >> 4, kt = g % p.n_kt, item = g / p.n_kt;
                if (kt == 0 && item > 0 && !mbar_wait_parked(&b[O_FREE], (item - 1) & 1, p.status, 36)) return false;
                tc_fence_after();
                const uint64_t dv = umma_desc_mn_sw128(smem_u32(base + AT_OFF_KV + st * 2 * AT_KV_BYTES + AT_KV_BYTES));
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (k < nk)
                        umma_f16_ts(tm + AT_TM_O, tm + AT_TM_P + (g & 1) * 32 + k * 8, dv + (uint64_t)k * (2048 >> 4), idesc_o, (kt != 0 || k != 0) ? 1u : 0u);
                }
                umma_commit(&b[O_FULL + (g & 1)]);
                umma_commit(&b[KV_EMPTY + st]);
                return true;
            };
            if (G > 0) ok = issue_s(0);
            if (ok && G > 1) ok = issue_s(1);
            for (int g = 0; g < G && ok; ++g) {
                if (!(ok = mbar_wait_spin(&b[P_FULL + (g & 1)], (g >> 1) & 1, p.status, 35))) break;
                // P(g) is written, so S[g & 1] is free: the scores the softmax warps will ask for next go first
                if (g + 2 < G && !(ok = issue_s(g + 2))) break;
                ok = issue_pv(g);
            }
        }
    } else {
        // ---- softmax: slot, TMEM lane quarter, tile row (one thread per row)
        const int s = (warp - 2) >> 2, q = warp & 3, r = q * 32 + lane;
        const float sl2 = 0.125f * 1.4426950408889634f;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + s * AT_TM_SLOT;
        const uint32_t t_o = t_lane + AT_TM_O;
        uint64_t* b = bars + s * NBAR;
        uint8_t* base = smem + s * AT_SLOT_BYTES;
        const int my_items = my_pairs - ((s == 1 && odd_tail) ? 1 : 0);
        int g = 0; bool ok = true;
        float m = 0.f, l = 0.f;                                         // m: reference maximum the probabilities are taken against
        float s4[4];
        // 32 probabilities exp2(s * sl2 - m) -> bf16 pairs into columns [c16, c16 + 16) of P[gg & 1]
        auto exp32 = [&](const uint32_t (&v)[32], int gg, int c16) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float e0 = ex2_approx(fmaf(__uint_as_float(v[2 * i]), sl2, -m)), e1 = ex2_approx(fmaf(__uint_as_float(v[2 * i + 1]), sl2, -m));
                s4[(2 * i) & 3] += e0; s4[(2 * i + 1) & 3] += e1;
                w[i] = pack_bf16x2(e0, e1);
            }
            tmem_st_32x16(t_lane + AT_TM_P + (gg & 1) * 32 + c16, w);
        };
        auto max32 = [&](const uint32_t (&v)[32]) {
            float a = fmaxf(__uint_as_float(v[0]), __uint_as_float(v[1])), c = fmaxf(__uint_as_float(v[2]), __uint_as_float(v[3]));
#pragma unroll
            for (int j = 4; j < 32; j += 4) {
                a = fmaxf(a, fmaxf(__uint_as_float(v[j]), __uint_as_float(v[j + 1])));
                c = fmaxf(c, fmaxf(__uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
            }
            return fmaxf(a, c);
        };
        auto publish = [&](int gg) {                                    // P of step gg is complete in TMEM
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&b[P_FULL + (gg & 1)]);
        };
        // One pipelined step: (lo, hi) hold the masked scores of step g and m is already valid for them; the scores of
        // step g + 1 are fetched into (nlo, nhi) and reduced to the next maximum while the SFUs work through step g.
        auto step = [&](uint32_t (&lo)[32], uint32_t (&hi)[32], uint32_t (&nlo)[32], uint32_t (&nhi)[32], int kt, int qrow) -> bool {
            const bool has_next = kt + 1 < p.n_kt;
            int nv_next = min(AT_KEYS, p.S - (kt + 1) * AT_KEYS);
            if (p.causal) nv_next = min(nv_next, qrow + 1 - (kt + 1) * AT_KEYS);   // may be <= 0: the whole step is masked for this row
            const uint32_t t_next = t_lane + AT_TM_S + ((g + 1) & 1) * 64;
            s4[0] = s4[1] = s4[2] = s4[3] = 0.f;
            exp32(lo, g, 0);
            if (has_next) {
                // S(g + 1) was started when P(g - 1) was published, one half step ago: by now it is (nearly) there
                if (!mbar_wait_spin(&b[S_FULL + ((g + 1) & 1)], ((g + 1) >> 1) & 1, p.status, 37)) return false;
                tc_fence_after();
                tmem_ld_32x32(t_next, nlo);
                tmem_ld_32x32(t_next + 32, nhi);
            }
            // the last key step of an image tower usually holds a handful of keys (S = 197: 5, S = 257: 1, S = 577: 1): the P V MMA of
            // the step then reads 16 or 32 probability columns only (keys16), and the upper 32 need not be computed at all.  (Causal
            // rows of one tile differ in how many keys they see, so there the columns are read and must be written.)
            if (p.causal || p.S - kt * AT_KEYS > 32) exp32(hi, g, 16);
            l += (s4[0] + s4[1]) + (s4[2] + s4[3]);
            publish(g);
            if (has_next) {
                tmem_ld_wait();
                mask_tail(nlo, nv_next);
                mask_tail(nhi, nv_next - 32);
                const float mx = fmaxf(m, fmaxf(max32(nlo), max32(nhi)) * sl2);
                const bool grow = mx - m > 8.f;
                if (__any_sync(0xffffffffu, grow)) {
                    // rare: move the reference; P V of step g (just released) must land before O is touched
                    if (!mbar_wait_parked(&b[O_FULL + (g & 1)], (g >> 1) & 1, p.status, 38)) return false;
                    tc_fence_after();
                    const float alpha = grow ? ex2_approx(m - mx) : 1.f;
#pragma unroll 1
                    for (int h = 0; h < 2; ++h) {
                        tmem_ld_32x32(t_o + h * 32, lo);                // (step g's scores are dead: reuse the registers)
                        tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 32; ++c) lo[c] = __float_as_uint(__uint_as_float(lo[c]) * alpha);
                        tmem_st_32x32(t_o + h * 32, lo);
                    }
                    tmem_st_wait();
                    l *= alpha;
                    if (grow) m = mx;
                }
            }
            return true;
        };
        for (int li = 0; li < my_items && ok; ++li) {
            const AttnItem it = attn_item(p, 2 * (blockIdx.x + li * gridDim.x) + s);
            const int qi = li & 1;
            uint32_t a_lo[32], a_hi[32], b_lo[32], b_hi[32];
            l = 0.f;
            // A warp whose 32 rows all lie past the end of the sequence (the last query tile of S = 257 holds ONE real row) has
            // nothing to compute: it only keeps the barrier protocol going, paced by the score barriers so that it can never arrive
            // twice in one phase.  Its SFU and issue slots go to the other slot's warps.
            if (it.qt * AT_ROWS + q * 32 >= p.S) {
                for (int kt = 0; kt < p.n_kt; ++kt, ++g) {
                    if (!(ok = mbar_wait_parked(&b[S_FULL + (g & 1)], (g >> 1) & 1, p.status, 37))) break;
                    tc_fence_after();
                    publish(g);
                }
                if (!ok) break;
                if (lane == 0) { mbar_arrive(&b[O_FREE]); mbar_arrive(&b[Q_EMPTY + qi]); }
                continue;
            }
            // first step of the item: load, mask, take the maximum as the reference
            if (!(ok = mbar_wait_parked(&b[S_FULL + (g & 1)], (g >> 1) & 1, p.status, 37))) break;
            tc_fence_after();
            const uint32_t t_s = t_lane + AT_TM_S + (g & 1) * 64;
            tmem_ld_32x32(t_s, a_lo);
            tmem_ld_32x32(t_s + 32, a_hi);
            tmem_ld_wait();
            const int qrow = it.qt * AT_ROWS + r;
            const int nv = p.causal ? min(min(AT_KEYS, p.S), qrow + 1) : min(AT_KEYS, p.S);
            mask_tail(a_lo, nv);
            mask_tail(a_hi, nv - 32);
            m = fmaxf(max32(a_lo), max32(a_hi)) * sl2;
            for (int kt = 0; kt < p.n_kt && ok; kt += 2) {
                if (!(ok = step(a_lo, a_hi, b_lo, b_hi, kt, qrow))) break;
                ++g;
                if (kt + 1 < p.n_kt) { if (!(ok = step(b_lo, b_hi, a_lo, a_hi, kt + 1, qrow))) break; ++g; }
            }
            if (!ok) break;
            if (!(ok = mbar_wait_parked(&b[O_FULL + ((g - 1) & 1)], ((g - 1) >> 1) & 1, p.status, 39))) break;
            tc_fence_after();
            // the row's 64 output columns out of TMEM (then the O columns go back to the MMA warp), normalised, staged as one
            // swizzled 128-byte row in the item's Q buffer (its MMAs are done)
            const float inv = 1.f / l;
            uint8_t* stage = base + qi * AT_Q_BYTES;
            tmem_ld_32x32(t_o, a_lo);
            tmem_ld_32x32(t_o + 32, a_hi);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&b[O_FREE]);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t (&v)[32] = i < 4 ? a_lo : a_hi;
                const int o = (i & 3) * 8;
                uint4 w;
                w.x = pack_bf16x2(__uint_as_float(v[o]) * inv, __uint_as_float(v[o + 1]) * inv);
                w.y = pack_bf16x2(__uint_as_float(v[o + 2]) * inv, __uint_as_float(v[o + 3]) * inv);
                w.z = pack_bf16x2(__uint_as_float(v[o + 4]) * inv, __uint_as_float(v[o + 5]) * inv);
                w.w = pack_bf16x2(__uint_as_float(v[o + 6]) * inv, __uint_as_float(v[o + 7]) * inv);
                *reinterpret_cast<uint4*>(stage + r * 128 + ((i ^ (r & 7)) << 4)) = w;
            }
            __syncwarp();
#pragma unroll
            for (int t = 0; t < 8; ++t) {                               // the warp's 32 rows, four 128-byte rows per instruction
                const int rr = q * 32 + t * 4 + (lane >> 3), ch = lane & 7;
                const int qrow = it.qt * AT_ROWS + rr;
                if (qrow < p.S) {
                    const uint4 w = *reinterpret_cast<const uint4*>(stage + rr * 128 + ((ch ^ (rr & 7)) << 4));
                    *reinterpret_cast<uint4*>(p.out + (size_t)(it.b * p.S + qrow) * p.d + it.h0 * AT_DH + ch * 8) = w;
                }
            }
            fence_proxy_async();                                        // the buffer goes back to the TMA producer
            __syncwarp();
            if (lane == 0) mbar_arrive(&b[Q_EMPTY + qi]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

// ------------------------------------------------------------------------------------------ short sequences (S <= 64)
// ViT-B/32 has 50 tokens: 6,144 (image, head) problems of 50 x 50 per layer at batch 512 - 1 % of the FLOPs, and HBM-bound by rights
// (118 MB of qkv in, 39 MB out: 24 us at the measured peak).  What bounds a per-problem pipeline there is not a pipe but the chain
// of hand-offs (TMA -> MMA -> softmax -> MMA -> store), a few microseconds long, so this kernel runs MANY short chains side by
// side instead of a deep one: four small CTAs per SM, each owning 128 TMEM columns, each walking its own list of tiles.
//   tile   = two (image, head) problems: rows 0..63 of the 128-row MMA are the queries of the first, rows 64..127 of the second
//   S      = Q K^T as ONE UMMA 128 x 128 x 16 x 4: Q = two stacked 64-row boxes, K likewise; the off-diagonal 64 x 64 blocks are
//            scores against the other problem's keys and are never read
//   P      = exp2(..) by 128 threads (thread = row = TMEM lane), written back over the score columns as bf16 pairs: keys of the
//            row's own problem get probabilities, the other problem's keys get 0, so that
//   O      = P V as ONE UMMA chain 128 x 64 x 16 x 8 over the two stacked V boxes (A = P straight from TMEM) leaves both
//            problems' outputs in columns 64..127 (whose scores are dead by then)
// Rows / keys past the sequence end inside a 64-row box belong to the next image (or are out of bounds = 0): keys are masked
// before the softmax (probability 0), query rows are computed and dropped at the store.
constexpr int SA_THREADS = 160;                               // warps 0..3 softmax (TMEM lane quarters), warp 4 = TMA + MMA issue
constexpr int SA_BOX = 64 * AT_DH * 2;                        // one 64-row x 128-byte box: 8 KB
constexpr int SA_SMEM = 6 * SA_BOX + 128 + 1024;

struct ShortAttnParams {
    int S, d, heads, items, tiles;
    __nv_bfloat16* out;
    DeviceStatus* status;
    int reverse;             // walk the (image, head) problems from the last image to the first
};

__global__ void __launch_bounds__(SA_THREADS, 4)
attention_short_kernel(const __grid_constant__ CUtensorMap tmQKV, ShortAttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sQ = smem;                      // 2 boxes: queries of problem 0 | problem 1; reused as the output staging area
    uint8_t* sK = smem + 2 * SA_BOX;
    uint8_t* sV = smem + 4 * SA_BOX;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * SA_BOX);
    uint64_t *ld_full = bars, *s_full = bars + 1, *p_full = bars + 2, *o_full = bars + 3, *q_free = bars + 4;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tmQKV);
        mbar_init(ld_full, 1); mbar_init(s_full, 1); mbar_init(p_full, 4); mbar_init(o_full, 1); mbar_init(q_free, 4);
        fence_mbar_init();
    }
    if (warp == 4) tmem_alloc(tmem_slot, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_trigger();

    if (warp == 4) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = attn_idesc(128, 128, 0), idesc_o = attn_idesc(128, 64, 1);
            bool ok = true;
            uint32_t ph = 0;
            for (int t = blockIdx.x; t < p.tiles && ok; t += gridDim.x, ph ^= 1) {
                // the previous tile's output has left the staging area (= the Q boxes)
                if (t != (int)blockIdx.x && !(ok = mbar_wait_spin(q_free, ph ^ 1, p.status, 41))) break;
                mbar_arrive_expect_tx(ld_full, 6 * SA_BOX);
                for (int u = 0; u < 2; ++u) {
                    const int raw = min(2 * t + u, p.items - 1), item = p.reverse ? p.items - 1 - raw : raw;
                    const int row0 = (item / p.heads) * p.S, col = (item % p.heads) * AT_DH;
                    tma_load_2d(sQ + u * SA_BOX, &tmQKV, ld_full, col, row0);
                    tma_load_2d(sK + u * SA_BOX, &tmQKV, ld_full, p.d + col, row0);
                    tma_load_2d(sV + u * SA_BOX, &tmQKV, ld_full, 2 * p.d + col, row0);
                }
                if (!(ok = mbar_wait_spin(ld_full, ph, p.status, 42))) break;
                tc_fence_after();
                const uint64_t dq = umma_desc_sw128(smem_u32(sQ)), dk = umma_desc_sw128(smem_u32(sK));
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tmem_base, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(s_full);
                if (!(ok = mbar_wait_spin(p_full, ph, p.status, 43))) break;
                tc_fence_after();
                const uint64_t dv = umma_desc_mn_sw128(smem_u32(sV));
#pragma unroll
                for (int k = 0; k < 8; ++k) umma_f16_ts(tmem_base + 64, tmem_base + k * 8, dv + (uint64_t)k * (2048 >> 4), idesc_o, k != 0);
                umma_commit(o_full);
            }
        }
    } else {
        const int r = warp * 32 + lane, hsel = r >> 6, qrow = r & 63;
        const float sl2 = 0.125f * 1.4426950408889634f;
        const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
        bool ok = true;
        uint32_t ph = 0;
        for (int t = blockIdx.x; t < p.tiles && ok; t += gridDim.x, ph ^= 1) {
            if (!(ok = mbar_wait_spin(s_full, ph, p.status, 44))) break;
            tc_fence_after();
            // this row's scores against its own problem's keys: columns [64 hsel, 64 hsel + 64)
            uint32_t a[32], b[32];
            tmem_ld_32x32(t_lane + hsel * 64, a);
            tmem_ld_32x32(t_lane + hsel * 64 + 32, b);
            tmem_ld_wait();
            mask_tail(a, p.S);
            mask_tail(b, p.S - 32);
            float m = __uint_as_float(a[0]);
#pragma unroll
            for (int j = 1; j < 32; ++j) m = fmaxf(m, __uint_as_float(a[j]));
#pragma unroll
            for (int j = 0; j < 32; ++j) m = fmaxf(m, __uint_as_float(b[j]));
            m *= sl2;
            float l0 = 0.f, l1 = 0.f;
            uint32_t w[16];
            // P columns: keys (2c, 2c + 1) of the stacked key tile in column c; own problem = columns [32 hsel, 32 hsel + 32)
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float e0 = ex2_approx(fmaf(__uint_as_float(a[2 * i]), sl2, -m)), e1 = ex2_approx(fmaf(__uint_as_float(a[2 * i + 1]), sl2, -m));
                l0 += e0; l1 += e1;
                w[i] = pack_bf16x2(e0, e1);
            }
            tmem_st_32x16(t_lane + hsel * 32, w);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float e0 = ex2_approx(fmaf(__uint_as_float(b[2 * i]), sl2, -m)), e1 = ex2_approx(fmaf(__uint_as_float(b[2 * i + 1]), sl2, -m));
                l0 += e0; l1 += e1;
                w[i] = pack_bf16x2(e0, e1);
            }
            tmem_st_32x16(t_lane + hsel * 32 + 16, w);
#pragma unroll
            for (int i = 0; i < 16; ++i) w[i] = 0u;                     // the other problem's keys
            tmem_st_32x16(t_lane + (hsel ^ 1) * 32, w);
            tmem_st_32x16(t_lane + (hsel ^ 1) * 32 + 16, w);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
            const float inv = 1.f / (l0 + l1);
            if (!(ok = mbar_wait_spin(o_full, ph, p.status, 45))) break;
            tc_fence_after();
            tmem_ld_32x32(t_lane + 64, a);
            tmem_ld_32x32(t_lane + 96, b);
            tmem_ld_wait();
            // normalised row -> one swizzled 128-byte row of the staging area (the Q boxes: both MMAs that read them are done)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t (&v)[32] = i < 4 ? a : b;
                const int o = (i & 3) * 8;
                uint4 x;
                x.x = pack_bf16x2(__uint_as_float(v[o]) * inv, __uint_as_float(v[o + 1]) * inv);
                x.y = pack_bf16x2(__uint_as_float(v[o + 2]) * inv, __uint_as_float(v[o + 3]) * inv);
                x.z = pack_bf16x2(__uint_as_float(v[o + 4]) * inv, __uint_as_float(v[o + 5]) * inv);
                x.w = pack_bf16x2(__uint_as_float(v[o + 6]) * inv, __uint_as_float(v[o + 7]) * inv);
                *reinterpret_cast<uint4*>(sQ + r * 128 + ((i ^ (r & 7)) << 4)) = x;
            }
            __syncwarp();
            // the warp's 32 rows, four full 128-byte rows per instruction
            const int raw = 2 * t + hsel;                               // warps 0,1 -> first problem, warps 2,3 -> second
            const int item = p.reverse ? p.items - 1 - raw : raw;
            if (raw < p.items) {
                const int b_img = item / p.heads, h = item % p.heads;
#pragma unroll
                for (int tt = 0; tt < 8; ++tt) {
                    const int rr = warp * 32 + tt * 4 + (lane >> 3), ch = lane & 7, q = rr & 63;
                    if (q < p.S) {
                        const uint4 x = *reinterpret_cast<const uint4*>(sQ + rr * 128 + ((ch ^ (rr & 7)) << 4));
                        *reinterpret_cast<uint4*>(p.out + (size_t)(b_img * p.S + q) * p.d + h * AT_DH + ch * 8) = x;
                    }
                }
            }
            (void)qrow;
            fence_proxy_async();                                        // the Q boxes go back to the TMA engine
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_free);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) { __syncwarp(); tmem_dealloc(tmem_base, 128); }
}

int launch_attention_short(const void* qkv, void* out, int B, int S, int d, int heads, DeviceStatus* status, cudaStream_t st, int reverse) {
    if (B <= 0) return B200_OK;
    B200_REQUIRE(S > 0 && S <= 64 && heads > 0 && d == heads * AT_DH, "attention_short: needs S <= 64 and head_dim 64");
    B200_REQUIRE((long long)B * heads < (1LL << 30), "attention_short: too many (image, head) problems");
    ShortAttnParams p;
    p.S = S; p.d = d; p.heads = heads; p.items = B * heads; p.tiles = (p.items + 1) / 2;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.status = status;
    p.reverse = reverse;
    CUtensorMap tm;
    int rc = make_tmap_2d(&tm, qkv, (uint64_t)B * S, (uint64_t)3 * d, (uint64_t)3 * d * 2, 64, false);
    if (rc) return rc;
    static bool attr_set[64] = {false};
    int dev = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev]) {
        B200_CUDA_OK(cudaFuncSetAttribute(attention_short_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SA_SMEM));
        if (dev < 64) attr_set[dev] = true;
    }
    const int grid = p.tiles < 4 * kNumSMs ? p.tiles : 4 * kNumSMs;
    B200_CUDA_OK(launch_pdl(attention_short_kernel, dim3(grid), dim3(SA_THREADS), SA_SMEM, st, tm, p));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

// Picks the kernel by shape: S <= 64 -> attention_short_kernel, longer sequences (and the causal text tower) -> attention_tc_kernel.
int launch_attention_auto(const void* qkv, void* out, int B, int S, int d, int heads, int causal, DeviceStatus* status, cudaStream_t st, int reverse) {
    B200_REQUIRE(heads > 0 && S > 0 && d == heads * AT_DH, "attention: head_dim must be 64");
    if (!causal && S <= AT_KEYS) return launch_attention_short(qkv, out, B, S, d, heads, status, st, reverse);
    return launch_attention_tc(qkv, out, B, S, d, heads, causal, status, st);
}

int launch_attention_tc(const void* qkv, void* out, int B, int S, int d, int heads, int causal, DeviceStatus* status, cudaStream_t st) {
    if (B <= 0) return B200_OK;
    B200_REQUIRE(!causal || S > AT_KEYS, "attention_tc: the causal mask is implemented for S > 64 only");
    AttnParams p;
    p.S = S; p.d = d; p.heads = heads; p.B = B;
    p.n_qt = (S + AT_ROWS - 1) / AT_ROWS;
    p.n_kt = (S + AT_KEYS - 1) / AT_KEYS;
    p.items = B * heads * p.n_qt;
    p.causal = causal ? 1 : 0;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.status = status;
    CUtensorMap tm;
    int rc = make_tmap_2d(&tm, qkv, (uint64_t)B * S, (uint64_t)3 * d, (uint64_t)3 * d * 2, 64, false);
    if (rc) return rc;
    static bool attr_set[64] = {false};
    int dev = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev]) {
        B200_CUDA_OK(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        if (dev < 64) attr_set[dev] = true;
    }
    const int n_pairs = (p.items + 1) / 2;
    const int grid = n_pairs < kNumSMs ? n_pairs : kNumSMs;
    B200_CUDA_OK(launch_pdl(attention_tc_kernel, dim3(grid), dim3(AT_THREADS), AT_SMEM, st, tm, p));
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

}  // namespace b200
