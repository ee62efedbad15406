// K10 cosine top-k fused into a tcgen05 GEMM, K12 candidate merge.
//
// Replaces, batched and without ever materialising the Q x N score matrix,
//   similarities = np.dot(img_embedding, text_embeddings.T); top = np.argsort(similarities)[::-1][:k]
//   packages/lumen-clip/src/lumen_clip/expert_bioclip/bioclip_model.py:313-316
// with the total order fixed by BASELINE.json: score descending, index ascending among equal scores.
//
// Grid = (query tiles of 128) x (gallery splits).  A CTA streams its gallery range through the same
// TMA -> tcgen05.mma -> TMEM pipeline as gemm_tc.cu (fp16 operands, fp32 accumulate).  The four epilogue
// warps own one query row per thread: each score is compared with that row's running threshold tau (its
// current k-th best) and survivors are appended, as 64-bit keys (order-preserving score bits | ~index), to a
// per-(split, query) candidate list in global scratch.  When a list is about to overflow, the warp shrinks it
// cooperatively (radix select of the k-th best key, no sort), keeps the best k and raises tau.  Because a split is scanned in
// index order, "score > tau" is exact under the index-ordered tie rule.  A second kernel merges the
// per-split (and, after the NCCL all-gather, per-GPU) lists with the same order.
#include "common.cuh"
#include "kernels.h"

#include <algorithm>
#include <cstdlib>
#include <mutex>

namespace b200 {

constexpr int TK_BM = 128, TK_BN = 256, TK_BK = 64, TK_STAGES = 4;
constexpr int TK_A_BYTES = TK_BM * TK_BK * 2, TK_B_BYTES = TK_BN * TK_BK * 2, TK_STAGE_BYTES = TK_A_BYTES + TK_B_BYTES;
constexpr int TK_SMEM = TK_STAGES * TK_STAGE_BYTES + 256 + 1024;
constexpr int TK_THREADS = 64 + 8 * 32;      // TMA warp, MMA warp, eight epilogue warps

struct TopkParams {
    int Q, D, k;
    long long N;                 // gallery rows on this GPU
    long long rows_per_split;    // multiple of 256
    int splits, qtiles;
    unsigned long long* cand;    // [splits][qtiles*128][32*CPL]
    float* part_scores;          // [Q][splits * k]: every split appends its survivors to the query's list (pcount = entries so far)
    long long* part_idx;         // [Q][splits * k]
    long long index_base;
    int scatter;                 // queries were spread over the four TMEM lane quarters (see scatter_queries_kernel)
    DeviceStatus* status;
    int stages;                  // gallery ring depth when the query tile is resident (QRES)
    unsigned* gtau;              // [Q] best lower bound any split has found for the query's k-th best score (flip_f32 key, 0 = none)
    unsigned* pubmax;            // [Q][splits] best score each split has seen so far for the query (flip_f32 key, 0 = nothing yet)
    int share_max;               // splits >= k: the k-th largest of the published maxima bounds the global k-th best (see kth_published_max)
    int* pcount;                 // [Q] candidates the splits have written for the query
    int room;                    // slots every (split, half) may use in a query's candidate slice (>= k)
    unsigned* chunkmax;          // != nullptr: seeding pass - no candidate lists, only the best score of every 32-row gallery chunk
                                 // per query, [Q][chunkmax_ld] (flip_f32 keys); see seed_tau_kernel
    int chunkmax_ld;
    int halves;                  // 2: eight epilogue warps, each lane quarter served by two warps that scan one half of a tile's columns
                                 // each and keep their own candidate lists (two independent scanners per row); 1: four warps
};

__device__ __forceinline__ uint32_t flip_f32(float s) {
    const uint32_t u = __float_as_uint(s + 0.0f);               // -0 -> +0 so equal scores have equal keys
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unflip_f32(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// Whole warp: shrink one row's candidate list (n keys, 32*CPL at most) and return the new count.
// No sort: a most-significant-bit-first radix select finds the k-th largest key (one warp-wide vote per key
// register and bit, branch-free inside), then the survivors are compacted to the front in arrival order.
//   exact == false (mid-scan): only the top 16 bits of the score are resolved.  The resulting threshold is a
//     valid lower bound of the k-th best score, a few extra candidates survive, and the select costs a quarter.
//     If too many candidates tie at that resolution to leave room for the next tile, it falls through to exact.
//   exact == true (end of the split): full 32-bit score, then -- only if candidates tie at the k-th score --
//     the low word picks the smallest indices.  With part_s/part_i the survivors are also written out as
//     (score, global index), padded with (-inf, -1); the merge kernel sorts, so no order is needed here.
// `tau_out` is the new threshold (-inf while fewer than k candidates exist).
// The select + write-back part works on keys already in registers (lane l holds entries l, 32 + l, ...; 0 = no entry), so that the
// mid-scan loop can have the next row's loads in flight while this row is processed.  `thr` != 0 first drops what scores below the
// shared bound (entries EQUAL to it stay: the index-ordered tie rule may still prefer them); the select then only runs if the row
// is still longer than `room`.
template <int CPL>
__device__ __forceinline__ int compact_regs(unsigned long long* rowptr, uint32_t (&hi)[CPL], uint32_t (&lo)[CPL], int n, int k, int lane,
                                            bool exact, bool force, int room, uint32_t thr, float* tau_out, float* part_s, long long* part_i,
                                            long long index_base, bool pad) {
    int keep = n;
    float tau = -INFINITY;
    bool changed = false;
    if (thr != 0u) {
        int c = 0;
#pragma unroll
        for (int i = 0; i < CPL; ++i) {
            if (hi[i] < thr) { hi[i] = 0u; lo[i] = 0u; }
            c += __popc(__ballot_sync(0xffffffffu, hi[i] != 0u));
        }
        changed = c != n;
        n = c;
    }
    const bool select = n >= k && (force || n > room);
    if (select || changed) {
        uint32_t p = 0, pl = 0; int kk = k;
        if (select) {
            auto select_hi = [&](int b_from, int b_to) {
#pragma unroll 1
                for (int b = b_from; b >= b_to; --b) {
                    const uint32_t want = p | (1u << b), m = 0xffffffffu << b;
                    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
                    for (int i = 0; i < CPL; i += 4) {
                        c0 += __popc(__ballot_sync(0xffffffffu, (hi[i] & m) == want));
                        c1 += __popc(__ballot_sync(0xffffffffu, (hi[i + 1] & m) == want));
                        c2 += __popc(__ballot_sync(0xffffffffu, (hi[i + 2] & m) == want));
                        c3 += __popc(__ballot_sync(0xffffffffu, (hi[i + 3] & m) == want));
                    }
                    const int c = (c0 + c1) + (c2 + c3);
                    if (c >= kk) p = want; else kk -= c;
                }
            };
            auto count_ge = [&]() {
                int c = 0;
#pragma unroll
                for (int i = 0; i < CPL; ++i) c += __popc(__ballot_sync(0xffffffffu, hi[i] >= p));
                return c;
            };
            select_hi(31, 16);
            bool coarse = !exact;
            if (coarse && count_ge() > room) coarse = false;   // degenerate: too many ties at 16-bit resolution
            if (!coarse) {
                select_hi(15, 0);
                int ceq = 0;
#pragma unroll
                for (int i = 0; i < CPL; ++i) ceq += __popc(__ballot_sync(0xffffffffu, hi[i] == p));
                if (ceq > kk) {                                // ties at the k-th score: kk-th largest low word = kk-th smallest index
#pragma unroll 1
                    for (int b = 31; b >= 0; --b) {
                        const uint32_t want = pl | (1u << b), m = 0xffffffffu << b;
                        int c = 0;
#pragma unroll
                        for (int i = 0; i < CPL; ++i) c += __popc(__ballot_sync(0xffffffffu, hi[i] == p && (lo[i] & m) == want));
                        if (c >= kk) pl = want; else kk -= c;
                    }
                }
            }
            tau = unflip_f32(p);
        } else {
            p = 1u;                                            // keep every entry that is still there
        }
        int base = 0;
        const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
        for (int i = 0; i < CPL; ++i) {
            const bool pred = hi[i] > p || (hi[i] == p && lo[i] >= pl);
            const uint32_t bal = __ballot_sync(0xffffffffu, pred);
            const int pos = base + __popc(bal & lt);
            if (pred) {
                __stcg(rowptr + pos, ((unsigned long long)hi[i] << 32) | lo[i]);
                if (part_s != nullptr && pos < k) { part_s[pos] = unflip_f32(hi[i]); part_i[pos] = index_base + (long long)(0xffffffffu - lo[i]); }
            }
            base += __popc(bal);
        }
        keep = base;
    } else if (part_s != nullptr) {
#pragma unroll
        for (int i = 0; i < CPL; ++i) {
            const int e = i * 32 + lane;
            if (e < k) {
                const bool valid = e < n;
                if (valid || pad) {
                    part_s[e] = valid ? unflip_f32(hi[i]) : -INFINITY;
                    part_i[e] = valid ? index_base + (long long)(0xffffffffu - lo[i]) : -1;
                }
            }
        }
    }
    *tau_out = tau;
    return keep;
}
template <int CPL>
__device__ __forceinline__ void load_row_keys(const unsigned long long* rowptr, int n, int lane, uint32_t (&hi)[CPL], uint32_t (&lo)[CPL]) {
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        const int e = i * 32 + lane;
        const unsigned long long key = e < n ? __ldcg(rowptr + e) : 0ull;     // 0 sorts below every real key
        hi[i] = (uint32_t)(key >> 32); lo[i] = (uint32_t)key;
    }
}
template <int CPL>
__device__ __noinline__ int compact_row(unsigned long long* rowptr, int n, int k, int lane, bool exact, int room,
                                        float* tau_out, float* part_s, long long* part_i, long long index_base, bool pad = true) {
    uint32_t hi[CPL], lo[CPL];
    __syncwarp();
    load_row_keys<CPL>(rowptr, n, lane, hi, lo);
    const int keep = compact_regs<CPL>(rowptr, hi, lo, n, k, lane, exact, /*force*/ true, room, 0u, tau_out, part_s, part_i, index_base, pad);
    __syncwarp();
    return keep;
}

// Whole warp: drop the entries of one row's list that score below `thr` (an order-preserving key).  Entries EQUAL to it stay: the
// index-ordered tie rule may still prefer them.  One pass, no select; returns the new count.
template <int CPL>
__device__ __forceinline__ int filter_row(unsigned long long* rowptr, int n, uint32_t thr, int lane) {
    unsigned long long key[CPL];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        const int e = i * 32 + lane;
        key[i] = e < n ? __ldcg(rowptr + e) : 0ull;
    }
    __syncwarp();                                     // every lane holds its entries before any slot is overwritten
    int base = 0;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        const bool pred = key[i] != 0ull && (uint32_t)(key[i] >> 32) >= thr;
        const uint32_t bal = __ballot_sync(0xffffffffu, pred);
        if (pred) __stcg(rowptr + base + __popc(bal & lt), key[i]);
        base += __popc(bal);
    }
    __syncwarp();
    return base;
}

// Small query batches run one CTA per gallery split (148 or 296 of them), each seeing only 1/148 of the rows: a split's own k-th best
// is a poor bound, and warming it up costs list compactions.  But every split can cheaply publish the BEST score it has seen for a
// query, and the k-th largest of those maxima is a lower bound of the global k-th best (k distinct rows, one per split, score at
// least that) - after one tile per split already near the 99.6th percentile, at the end within a hair of the true threshold.
// Whole warp: that k-th largest for one query, resolved to the top 16 bits of the key (rounded down: still a valid bound); 0 while
// fewer than k splits have published.
template <int MAXV = 10>                              // S <= 32 * MAXV (10: one value per split, at most 2 * 148 of them)
__device__ __forceinline__ uint32_t kth_published_max(const unsigned* __restrict__ pub, int S, int k, int lane) {
    uint32_t v[MAXV];
    int have = 0;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int idx = i * 32 + lane;
        v[i] = (i * 32 < S && idx < S) ? __ldcg(pub + idx) : 0u;
        have += __popc(__ballot_sync(0xffffffffu, v[i] != 0u));
    }
    if (have < k) return 0u;
    uint32_t p = 0; int kk = k;
#pragma unroll 1
    for (int b = 31; b >= 16; --b) {
        const uint32_t want = p | (1u << b), m = 0xffffffffu << b;
        int c = 0;
#pragma unroll
        for (int i = 0; i < MAXV; ++i) if (i * 32 < S) c += __popc(__ballot_sync(0xffffffffu, (v[i] & m) == want));
        if (c >= kk) p = want; else kk -= c;
    }
    return p;
}

// Warm start for batches of several query tiles.  A CTA that starts with no bound appends whole tiles and compacts its lists every
// few tiles until its threshold has crept up - ~0.6 ms of a 4.4 ms search (Q = 4096, 1.25 M rows), paid by every CTA of the first
// wave at once.  A bound is much cheaper to get than a top-k: the best score of each of n >= k disjoint gallery chunks are n
// distinct rows, so the k-th largest of them is a lower bound of the k-th best score.  So the scan kernel first runs over the
// first TK_SEED_ROWS rows in a mode that only records the maximum of every 32-row chunk (the fast path: MMA-bound, ~50 us at
// Q = 4096), and this kernel turns each query's 512 chunk maxima into its starting bound (one warp per query).  With k = 100 that
// bound already rejects ~99.3 % of all scores, where a cold CTA needs ~60 tiles of appending and compacting to get.  Measured on
// a 1.25 M-row shard: Q = 256 0.70 -> 0.36 ms, Q = 512 1.02 -> 0.67 ms, Q = 4096 4.34 -> 4.0 ms (4096, 8192, 32768 and 65536
// sample rows: 4.14, 4.05, 4.08 and 4.40 ms).
constexpr int TK_SEED_ROWS = 16384, TK_SEED_CHUNKS = TK_SEED_ROWS / 32;
__global__ void __launch_bounds__(256) seed_tau_kernel(const unsigned* __restrict__ chunkmax, int ld, int chunks, int Q, int k, unsigned* __restrict__ gtau) {
    const int qi = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (qi >= Q) return;
    const uint32_t kth = kth_published_max<TK_SEED_CHUNKS / 32>(chunkmax + (size_t)qi * ld, chunks, k, lane);
    if (kth != 0u && lane == 0) atomicMax(gtau + qi, kth);
}

// QRES: the CTA's query tile (128 x D fp16, D <= 512: at most 128 KB) is loaded ONCE and stays in shared memory; only gallery
// tiles go through the ring (3 stages of 32 KB at D = 512).  Without it every gallery tile re-fetches the query tile from L2,
// a third of the kernel's L2 -> SM traffic, and at Q = 4096 that traffic, not the tensor core, is what bounds the kernel.
template <int CPL, bool QRES>
__global__ void __launch_bounds__(TK_THREADS, 1)
topk_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmG, TopkParams p) {
    constexpr int C = 32 * CPL;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int kblocks = p.D / TK_BK;
    const int n_stages = QRES ? p.stages : TK_STAGES;
    const int stage_bytes = QRES ? TK_B_BYTES : TK_STAGE_BYTES;
    uint8_t* ring = smem + (QRES ? kblocks * TK_A_BYTES : 0);
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + n_stages * stage_bytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + TK_STAGES;
    uint64_t* acc_full = bars + 2 * TK_STAGES;
    uint64_t* acc_empty = acc_full + 2;
    uint64_t* q_full = acc_empty + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x, sp = blockIdx.y;
    const long long n_lo = (long long)sp * p.rows_per_split;
    const long long n_hi = min(p.N, n_lo + p.rows_per_split);
    const int tiles = n_hi > n_lo ? (int)((n_hi - n_lo + TK_BN - 1) / TK_BN) : 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmG);
        for (int s = 0; s < TK_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(q_full, 1);
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 8); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 2 * TK_BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int s = 0; uint32_t ph = 0; bool ok = true;
            if (QRES && tiles > 0) {
                mbar_arrive_expect_tx(q_full, kblocks * TK_A_BYTES);
                for (int kb = 0; kb < kblocks; ++kb) tma_load_2d(smem + kb * TK_A_BYTES, &tmQ, q_full, kb * TK_BK, qt * TK_BM);
            }
            for (int t = 0; t < tiles && ok; ++t) {
                const int n0 = (int)(n_lo + (long long)t * TK_BN);
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (!(ok = mbar_wait(&empty[s], ph ^ 1, p.status, 11))) break;
                    uint8_t* sa = ring + s * stage_bytes;
                    mbar_arrive_expect_tx(&full[s], stage_bytes);
                    if (!QRES) tma_load_2d(sa, &tmQ, &full[s], kb * TK_BK, qt * TK_BM);
                    tma_load_2d(sa + (QRES ? 0 : TK_A_BYTES), &tmG, &full[s], kb * TK_BK, n0);
                    if (++s == n_stages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc(TK_BM, TK_BN, /*f16*/ 0);
            int s = 0; uint32_t ph = 0; bool ok = true;
            if (QRES && tiles > 0) ok = mbar_wait(q_full, 0, p.status, 14);
            for (int t = 0; t < tiles && ok; ++t) {
                const int as = t & 1; const uint32_t aph = (t >> 1) & 1;
                if (!(ok = mbar_wait(&acc_empty[as], aph ^ 1, p.status, 12))) break;
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * TK_BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (!(ok = mbar_wait(&full[s], ph, p.status, 13))) break;
                    tc_fence_after();
                    const uint32_t sb = smem_u32(ring + s * stage_bytes);
                    const uint64_t da = umma_desc_sw128(QRES ? smem_u32(smem + kb * TK_A_BYTES) : sb);
                    const uint64_t db = umma_desc_sw128(QRES ? sb : sb + TK_A_BYTES);
#pragma unroll
                    for (int k = 0; k < TK_BK / 16; ++k) umma_f16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit(&empty[s]);
                    if (++s == n_stages) { s = 0; ph ^= 1; }
                }
                if (ok) umma_commit(&acc_full[as]);
            }
        }
    } else {
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;                           // which half of a tile's columns this warp scans
        const int row = q * 32 + lane;                              // query row inside the tile = TMEM lane
        // query handled by this thread: tile row order, or (small batches) interleaved so all four warps get rows
        const long long qrow = (long long)qt * TK_BM + (p.scatter ? (lane * 4 + q) : row);
        // lists are indexed by (scanner, tile row): [splits][qtiles][halves][128][C]
        unsigned long long* warp_rows = p.cand + ((((size_t)sp * p.qtiles + qt) * p.halves + half) * TK_BM + q * 32) * C;
        unsigned long long* rowptr = warp_rows + (size_t)lane * C;
        if (half >= p.halves) {                                     // four-warp mode: the second set only keeps the barrier protocol going
            for (int t = 0; t < tiles; ++t) {
                const int as = t & 1; const uint32_t aph = (t >> 1) & 1;
                if (!mbar_wait(&acc_full[as], aph, p.status, 15)) break;
                tc_fence_after();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[as]);
            }
        } else if (p.chunkmax != nullptr) {
            // seeding pass over the first rows of the gallery: the fast path only (a maximum per 32 scores), nothing is appended
            for (int t = 0; t < tiles; ++t) {
                const int as = t & 1; const uint32_t aph = (t >> 1) & 1;
                if (!mbar_wait(&acc_full[as], aph, p.status, 16)) break;
                tc_fence_after();
                const long long n0 = n_lo + (long long)t * TK_BN;
                const uint32_t tsrc = tmem_base + ((uint32_t)(q * 32) << 16) + as * TK_BN;
                uint32_t v[32];
#pragma unroll 1
                for (int c = 0; c < TK_BN / 32; ++c) {
                    tmem_ld_32x32(tsrc + c * 32, v);
                    tmem_ld_wait();
                    const long long left = n_hi - (n0 + c * 32);
                    float mx = -INFINITY;
#pragma unroll
                    for (int j = 0; j < 32; ++j) if (j < left) mx = fmaxf(mx, __uint_as_float(v[j]));
                    if (qrow < p.Q && left > 0) p.chunkmax[(size_t)qrow * p.chunkmax_ld + (n0 >> 5) + c] = flip_f32(mx);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[as]);
            }
        } else {
        const int c_lo = p.halves == 2 ? half * (TK_BN / 64) : 0, c_hi = p.halves == 2 ? c_lo + TK_BN / 64 : TK_BN / 32;   // this warp's 32-column chunks
        const int span = (c_hi - c_lo) * 32;                        // columns (= possible appends) per tile
        float tau = qrow < p.Q ? -INFINITY : INFINITY;      // padding rows of the last query tile never collect candidates
        int cnt = 0;
        bool ok = true;
        float rowmax = -INFINITY;                            // best score this split has seen for the row (share_max)
        for (int t = 0; t < tiles && ok; ++t) {
            const int as = t & 1; const uint32_t aph = (t >> 1) & 1;
            const unsigned g_pre = qrow < p.Q ? __ldcg(p.gtau + qrow) : 0u;     // issued before the wait: the latency hides behind it
            if (!(ok = mbar_wait(&acc_full[as], aph, p.status, 14))) break;
            tc_fence_after();
            const long long n0 = n_lo + (long long)t * TK_BN;
            const uint32_t tsrc = tmem_base + ((uint32_t)(q * 32) << 16) + as * TK_BN;
            if (p.share_max && t == 0) {
                // No bound exists yet, and appending the whole first tile (256 entries per row, one scattered 8-byte store each) is
                // what small batches used to pay most for.  So the first tile is read twice from TMEM: once for the row maxima only,
                // which are published at once; then the CTA waits - bounded, a few microseconds - until k splits have reported for
                // its query and the bounds of its own rows have arrived from the CTAs that own them; then the normal scan runs
                // against a bound that already sits near the 99.6th percentile.  If the wait times out the scan just appends more.
                uint32_t v0[32];
#pragma unroll 1
                for (int c = c_lo; c < c_hi; ++c) {
                    tmem_ld_32x32(tsrc + c * 32, v0);
                    tmem_ld_wait();
                    const long long lf = n_hi - (n0 + c * 32);
#pragma unroll
                    for (int j = 0; j < 32; ++j) if (j < lf) rowmax = fmaxf(rowmax, __uint_as_float(v0[j]));
                }
                if (qrow < p.Q && rowmax > -INFINITY) atomicMax(p.pubmax + (size_t)qrow * p.splits + sp, flip_f32(rowmax));
                if (warp == 2) {
                    const int myq = sp % p.Q;
                    uint32_t kth = 0u;
#pragma unroll 1
                    for (int tries = 0; tries < 32 && kth == 0u; ++tries) {
                        kth = kth_published_max(p.pubmax + (size_t)myq * p.splits, p.splits, p.k, lane);
                        if (kth == 0u) __nanosleep(400);
                    }
                    if (kth != 0u && lane == 0) atomicMax(p.gtau + myq, kth);
                }
#pragma unroll 1
                for (int tries = 0; tries < 96; ++tries) {
                    const unsigned g0 = qrow < p.Q ? __ldcg(p.gtau + qrow) : 2u;
                    if (!__any_sync(0xffffffffu, g0 <= 1u)) break;
                    __nanosleep(400);
                }
            }
            // Threshold sharing: the k-th best score another gallery split has already established for this query bounds the global
            // k-th best from below, so nothing below it can reach the final top-k.  Scores EQUAL to it are kept (key - 1 = the next
            // float below): the index-ordered tie rule may still prefer them.  Without this every split warms up on its own.
            if (qrow < p.Q) {
                unsigned g = (p.share_max && t == 0) ? __ldcg(p.gtau + qrow) : g_pre;    // tile 0 has just waited for its bound
                if (g > 1u) {
                    // the float just below the shared bound.  Below +0.0 that is the negative denormal, not -0.0 (key - 1), which
                    // compares EQUAL to +0.0 and would drop exact-zero scores (zero or padding rows) the tie rule may still prefer.
                    g = g == 0x80000000u ? 0x7ffffffeu : g - 1u;
                    tau = fmaxf(tau, unflip_f32(g));
                }
            }
            // Scan 8 chunks of 32 scores; the TMEM load of chunk c+1 is in flight while chunk c is examined.
            // A thread first reduces its 32 scores to their maximum: only if that beats tau (rare once the
            // threshold has settled) does it walk the chunk and append the survivors.
            const bool ragged = n_hi - n0 < TK_BN;                               // last tile of the gallery: columns past the end hold zeros
            // Scan this warp's chunks of 32 scores; the TMEM load of chunk c+1 is in flight while chunk c is examined.
            // A thread first reduces its 32 scores to their maximum: only if that beats tau (rare once the
            // threshold has settled) does it walk the chunk and append the survivors.
            auto scan = [&](const uint32_t (&v)[32], int c) {
                float mx = __uint_as_float(v[0]);
#pragma unroll
                for (int j = 1; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(v[j]));
                const long long nb = n0 + c * 32;
                const long long left = n_hi - nb;                                 // < 32 only in the last tile of the gallery
                if (p.share_max) {
                    if (!ragged) rowmax = fmaxf(rowmax, mx);
                    else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) if (j < left) rowmax = fmaxf(rowmax, __uint_as_float(v[j]));
                    }
                }
                if (!__any_sync(0xffffffffu, mx > tau)) return;                   // nothing in this 32x32 block beats any row's threshold
                // branch-free bit mask of this row's survivors, then one loop iteration per survivor (usually 0 or 1):
                // the score is picked out of the 32 registers with a 5-level select tree instead of 32 branches
                uint32_t mask = 0;
#pragma unroll
                for (int j = 0; j < 32; ++j) mask |= (__uint_as_float(v[j]) > tau ? 1u : 0u) << j;
                if (left < 32) mask = left <= 0 ? 0u : (mask & ((1u << (int)left) - 1u));
                const uint32_t inv = 0xffffffffu - (uint32_t)nb;
                while (mask) {
                    const int j = __ffs(mask) - 1;
                    mask &= mask - 1;
                    uint32_t a16[16], a8[8], a4[4], a2[2];
#pragma unroll
                    for (int i = 0; i < 16; ++i) a16[i] = (j & 1) ? v[2 * i + 1] : v[2 * i];
#pragma unroll
                    for (int i = 0; i < 8; ++i) a8[i] = (j & 2) ? a16[2 * i + 1] : a16[2 * i];
#pragma unroll
                    for (int i = 0; i < 4; ++i) a4[i] = (j & 4) ? a8[2 * i + 1] : a8[2 * i];
#pragma unroll
                    for (int i = 0; i < 2; ++i) a2[i] = (j & 8) ? a4[2 * i + 1] : a4[2 * i];
                    const float s = __uint_as_float((j & 16) ? a2[1] : a2[0]);
                    __stcg(rowptr + cnt, ((unsigned long long)flip_f32(s) << 32) | (inv - (uint32_t)j));
                    ++cnt;
                }
            };
            uint32_t va[32], vb[32];
            tmem_ld_32x32(tsrc + c_lo * 32, va);
#pragma unroll 1
            for (int c = c_lo; c < c_hi; c += 2) {
                tmem_ld_wait();
                tmem_ld_32x32(tsrc + (c + 1) * 32, vb);
                scan(va, c);
                tmem_ld_wait();
                if (c + 2 < c_hi) tmem_ld_32x32(tsrc + (c + 2) * 32, va);
                scan(vb, c + 1);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[as]);
            if (p.share_max) {
                // publish this split's best score per query, and turn everybody's maxima into a bound for ONE query per CTA and tile
                // (query (sp + t) mod Q: with at least as many CTAs as queries every query is refreshed by somebody each round;
                // the result reaches all CTAs through gtau).  One row's worth of work (~1 us) for the first epilogue warp.
                if (qrow < p.Q && rowmax > -INFINITY) atomicMax(p.pubmax + (size_t)qrow * p.splits + sp, flip_f32(rowmax));
                if (warp == 2 && (t < 4 || (t & 3) == 0)) {
                    const int myq = (sp + t) % p.Q;
                    const uint32_t kth = kth_published_max(p.pubmax + (size_t)myq * p.splits, p.splits, p.k, lane);
                    if (kth != 0u && lane == 0) atomicMax(p.gtau + myq, kth);
                }
            }
            // make room for the next tile (256 possible appends) while the MMAs of tile t+1 run: drop what the shared bound already
            // rules out, and only if the list is still too long select its k best.  One row at a time for the whole warp, so the
            // cost is L2 round trips: every lane fetches its own row's bound up front, a row's keys are loaded once for both steps,
            // and the next row's loads are in flight while this one is processed (a cold start spends most of its time here).
            unsigned need = __ballot_sync(0xffffffffu, cnt + span > C);
            if (need) {
                const unsigned g_mine = qrow < p.Q ? __ldcg(p.gtau + qrow) : 0u;
                uint32_t hi[CPL], lo[CPL];
                int src = __ffs(need) - 1;
                __syncwarp();                                // the row's appends are visible to the whole warp
                load_row_keys<CPL>(warp_rows + (size_t)src * C, __shfl_sync(0xffffffffu, cnt, src), lane, hi, lo);
                while (need) {
                    need &= need - 1;
                    const int nxt = need ? __ffs(need) - 1 : -1;
                    constexpr int CP2 = CPL <= 16 ? CPL : 1;      // (1024-entry lists: two rows of keys do not fit the register file)
                    uint32_t hi2[CP2], lo2[CP2];
                    if (CPL <= 16 && nxt >= 0) load_row_keys<CP2>(warp_rows + (size_t)nxt * C, __shfl_sync(0xffffffffu, cnt, nxt), lane, hi2, lo2);
                    const int n0r = __shfl_sync(0xffffffffu, cnt, src);
                    const unsigned g = __shfl_sync(0xffffffffu, g_mine, src);
                    float tnew = -INFINITY;
                    const int n = compact_regs<CPL>(warp_rows + (size_t)src * C, hi, lo, n0r, p.k, lane, false, false, C - span, g > 1u ? g : 0u,
                                                    &tnew, nullptr, nullptr, 0, false);
                    if (lane == src) {
                        if (tnew > -INFINITY && qrow < p.Q) atomicMax(p.gtau + qrow, flip_f32(tnew));   // k rows of THIS split score >= tnew
                        cnt = n; tau = fmaxf(tau, tnew);
                    }
                    if (nxt >= 0) {
                        if (CPL <= 16) {
#pragma unroll
                            for (int i = 0; i < CP2; ++i) { hi[i] = hi2[i]; lo[i] = lo2[i]; }
                        } else {
                            __syncwarp();
                            load_row_keys<CPL>(warp_rows + (size_t)nxt * C, __shfl_sync(0xffffffffu, cnt, nxt), lane, hi, lo);
                        }
                    }
                    src = nxt;
                }
                __syncwarp();
            }
        }
        // final: what is left of every row's list goes to the query's candidate slice for the merge kernel, as (score, global index).
        // One LANE per row: the shared bound usually leaves fewer than k entries, which need no selecting - and done one row at a
        // time by the whole warp this stage was a chain of 32 x (load, atomic, load, store) round trips to L2, ~8 us each, ~250 us
        // per CTA during which its SM issued no MMA (a tenth of a Q = 4096 search over 1.25 M rows).
        const size_t cap = (size_t)p.splits * p.halves * p.room;
        const bool live = qrow < p.Q;
        int m = 0;
        if (p.room > p.k) {
            // several query tiles: the lists go to the merge kernel whole, no filtering (the shared bound is one split's own k-th
            // best here and removes little).  Warp-wide copy, two rows per step and up to sixteen loads in flight per lane.
            m = cnt;
            const bool direct = live && cnt > 0 && cnt <= p.room;
            int base = 0;
            if (direct) base = atomicAdd(p.pcount + qrow, cnt);                       // this split's part of the query's candidate slice
            const unsigned todo = __ballot_sync(0xffffffffu, direct);
#pragma unroll 1
            for (int r0 = 0; r0 < 32; r0 += 2) {
                int n[2]; long long qr[2]; int bs[2];
#pragma unroll
                for (int x = 0; x < 2; ++x) {
                    n[x] = ((todo >> (r0 + x)) & 1u) ? __shfl_sync(0xffffffffu, cnt, r0 + x) : 0;
                    qr[x] = __shfl_sync(0xffffffffu, qrow, r0 + x);
                    bs[x] = __shfl_sync(0xffffffffu, base, r0 + x);
                }
#pragma unroll 1
                for (int e0 = 0; e0 < max(n[0], n[1]); e0 += 256) {
                    unsigned long long key[2][8];
#pragma unroll
                    for (int x = 0; x < 2; ++x)
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int e = e0 + i * 32 + lane;
                            key[x][i] = e < n[x] ? __ldcg(warp_rows + (size_t)(r0 + x) * C + e) : 0ull;
                        }
#pragma unroll
                    for (int x = 0; x < 2; ++x)
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int e = e0 + i * 32 + lane;
                            if (e < n[x]) {
                                p.part_scores[(size_t)qr[x] * cap + bs[x] + e] = unflip_f32((uint32_t)(key[x][i] >> 32));
                                p.part_idx[(size_t)qr[x] * cap + bs[x] + e] = p.index_base + (long long)(0xffffffffu - (uint32_t)key[x][i]);
                            }
                        }
                }
            }
        } else if (live) {
            const unsigned g = __ldcg(p.gtau + qrow);
            const uint32_t thr = g > 1u ? g : 0u;
#pragma unroll 1
            for (int e = 0; e < cnt; e += 8) {
                unsigned long long key[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) key[j] = e + j < cnt ? __ldcg(rowptr + e + j) : 0ull;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (key[j] != 0ull && (uint32_t)(key[j] >> 32) >= thr) __stcg(rowptr + m++, key[j]);     // m <= e + j: never ahead of the reads
            }
            if (m > 0 && m <= p.room) {
                const int base = atomicAdd(p.pcount + qrow, m);                       // this split's slice of the query's candidate list
                float* ps = p.part_scores + (size_t)qrow * cap + base;
                long long* pi = p.part_idx + (size_t)qrow * cap + base;
#pragma unroll 1
                for (int e = 0; e < m; e += 4) {
                    unsigned long long key[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) key[j] = e + j < m ? __ldcg(rowptr + e + j) : 0ull;
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (e + j < m) { ps[e + j] = unflip_f32((uint32_t)(key[j] >> 32)); pi[e + j] = p.index_base + (long long)(0xffffffffu - (uint32_t)key[j]); }
                }
            }
        }
        __syncwarp();
        // rows longer than their share of the slice (rare: the list was close to full when the scan ended): the warp selects the k
        // best, one row at a time.  Everything shorter goes to the merge kernel as it is - selecting k out of ~200 here, row after
        // row, only to have the merge select again cost every CTA ~200 us at its end.
        unsigned big = __ballot_sync(0xffffffffu, live && m > p.room);
        while (big) {
            const int src = __ffs(big) - 1;
            big &= big - 1;
            const int n = __shfl_sync(0xffffffffu, m, src);
            const long long qr = __shfl_sync(0xffffffffu, qrow, src);
            int base = 0;
            if (lane == 0) base = atomicAdd(p.pcount + qr, p.k);
            base = __shfl_sync(0xffffffffu, base, 0);
            float tnew;
            compact_row<CPL>(warp_rows + (size_t)src * C, n, p.k, lane, true, C, &tnew, p.part_scores + (size_t)qr * cap + base,
                             p.part_idx + (size_t)qr * cap + base, p.index_base, /*pad*/ false);
        }
        }   // scanning warps
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { __syncwarp(); tmem_dealloc(tmem_base, 2 * TK_BN); }
}

// ------------------------------------------------------------------------------------------ merge (K12)
// parts: scores [P][Q][k] fp32, idx [P][Q][k] int64 (idx < 0 = empty slot), in any order.  One CTA per query:
//   1. block radix select (8 bits per pass, shared-memory histogram) of the k-th best score over the P*k candidates,
//   2. if more candidates tie with that score than fit, a second select keeps the smallest indices among them,
//   3. the <= k survivors are gathered into shared memory and sorted there (score descending, index ascending).
// Work is O(P*k) per query instead of a full sort of P*k entries.
__device__ __forceinline__ bool better(float sa, long long ia, float sb, long long ib) {
    return sa > sb || (sa == sb && ia < ib);
}
// one CTA per query: 1024 threads when the queries are few (the kernel is a latency chain then and gets every warp an SM has), 256 otherwise
constexpr int MERGE_KMAX = 1024;

// hist[digit] += 1 for every participating lane, one shared-memory atomic per distinct digit in the warp
// (candidate scores cluster in a few high-order digits, so plain per-thread atomics would serialise).
__device__ __forceinline__ void hist_add(int* hist, bool active, int digit) {
    const unsigned act = __ballot_sync(0xffffffffu, active);
    if (!active) return;
    const unsigned peers = __match_any_sync(act, digit);
    if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[digit], __popc(peers));
}

// Finds the digit d (scanning from `hi_first ? 255 : 0`) where the running count reaches kk; returns it and the count before
// it, plus the histogram total.  Called by all threads after the histogram is complete; warp 0 scans (8 bins per lane + a
// shuffle prefix sum: a single thread walking 256 shared-memory words was a fifth of this kernel's time at Q = 1).
__device__ __forceinline__ void scan_hist(const int* hist, int kk, bool hi_first, int* sh_digit, int* sh_before, int* sh_total) {
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int loc[8], sum = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { const int s = lane * 8 + i; loc[i] = hist[hi_first ? 255 - s : s]; sum += loc[i]; }
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
        const unsigned hit = __ballot_sync(0xffffffffu, inc >= kk);
        const int src = hit ? __ffs(hit) - 1 : 31;
        const int total = __shfl_sync(0xffffffffu, inc, 31);
        if (lane == src) {
            int acc = inc - sum, i = 0;
            for (; i < 7; ++i) { if (acc + loc[i] >= kk) break; acc += loc[i]; }
            if (!hit) acc = total;                            // never reached (callers exclude it): same answer as a full walk
            const int s = lane * 8 + i;
            *sh_digit = hi_first ? 255 - s : s; *sh_before = acc; *sh_total = total;
        }
    }
}

// Candidate exchange format between GPUs: one 64-bit key per candidate = order-preserving score bits (high word) | ~(row index inside
// the owning shard) (low word), 0 = empty slot.  8 bytes instead of 4 + 8, ONE all-gather instead of two; the receiving merge adds each
// part's index_base back (bases[part]).  KEYS_IN selects that input form; out_keys != nullptr makes the kernel emit it (indices
// relative to key_base) instead of (out_s, out_i).
template <int NT, bool KEYS_IN>
__global__ void __launch_bounds__(NT) merge_topk_kernel(const float* __restrict__ ps, const long long* __restrict__ pi,
                                                                   const unsigned long long* __restrict__ pkeys, const long long* __restrict__ bases, int P,
                                                                   int Q, int k, int cache_cap, float* __restrict__ out_s,
                                                                   long long* __restrict__ out_i, unsigned long long* __restrict__ out_keys, long long key_base,
                                                                   const int* __restrict__ counts, int slot) {
    extern __shared__ __align__(16) uint32_t key_cache[];     // [T] order-preserving score keys (0 = empty slot) when cache_keys
    __shared__ int hist[256];
    __shared__ int sh_digit, sh_before, sh_total, sh_count, sh_gt, sh_eq;
    __shared__ float ss[MERGE_KMAX];
    __shared__ long long si[MERGE_KMAX];
    // counts != nullptr: the query's candidates are one list of counts[qi] entries at [qi][slot] (what topk_tc_kernel writes);
    // otherwise P parts of k slots each, [P][Q][k], empty slots marked by a negative index (the public merge entry point)
    const int qi = blockIdx.x, tid = threadIdx.x;
    const int T = counts != nullptr ? min(counts[qi], slot) : P * k;
    const bool cache_keys = T <= cache_cap;                    // the keys fit in the shared memory this launch was given
    auto offset = [&](int e) -> size_t {
        if (counts != nullptr) return (size_t)qi * slot + e;
        const int part = e / k;
        return ((size_t)part * Q + qi) * k + (e - part * k);
    };
    // key of candidate e: from shared memory when cached, else recomputed from global memory
    auto key_of = [&](int e) -> uint32_t {
        if (cache_keys) return key_cache[e];
        const size_t o = offset(e);
        if (KEYS_IN) return (uint32_t)(pkeys[o] >> 32);
        return pi[o] < 0 ? 0u : flip_f32(ps[o] + 0.0f);
    };
    auto id_of = [&](int e) -> long long {       // global index of a valid candidate
        const size_t o = offset(e);
        if (KEYS_IN) return bases[e / k] + (long long)(0xffffffffu - (uint32_t)pkeys[o]);
        return pi[o];
    };
    if (tid == 0) { sh_count = 0; sh_gt = 0; sh_eq = 0; }
    if (cache_keys) {                                          // one pass over global memory, four loads in flight per thread
        for (int e0 = 0; e0 < T; e0 += 4 * NT) {
            float s[4]; long long id[4]; unsigned long long kk8[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int e = e0 + u * NT + tid;
                id[u] = -1; s[u] = 0.f; kk8[u] = 0ull;
                if (e < T) {
                    const size_t o = offset(e);
                    if (KEYS_IN) kk8[u] = pkeys[o]; else { id[u] = pi[o]; s[u] = ps[o]; }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int e = e0 + u * NT + tid;
                if (e < T) key_cache[e] = KEYS_IN ? (uint32_t)(kk8[u] >> 32) : (id[u] < 0 ? 0u : flip_f32(s[u] + 0.0f));
            }
        }
    }
    __syncthreads();
    // ---- k-th best score: 4 passes of 8 bits over the order-preserving 32-bit image of the score
    // (real keys are never 0: flip_f32 maps -inf to 0x007fffff and NaNs are not produced by finite inputs)
    uint32_t prefix = 0; int kk = k; bool all = false;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        if (tid < 256) hist[tid] = 0;
        __syncthreads();
        for (int e0 = 0; e0 < T; e0 += NT) {          // whole warps iterate together (hist_add votes)
            const int e = e0 + tid;
            const uint32_t key = e < T ? key_of(e) : 0u;
            const bool act = key != 0u && (pass == 0 || (key >> (shift + 8)) == (prefix >> (shift + 8)));
            hist_add(hist, act, (key >> shift) & 255);
        }
        __syncthreads();
        scan_hist(hist, kk, true, &sh_digit, &sh_before, &sh_total);
        __syncthreads();
        if (pass == 0 && sh_total <= k) { all = true; break; }   // fewer valid candidates than k: keep them all
        prefix |= (uint32_t)sh_digit << shift;
        kk -= sh_before;
        __syncthreads();
    }
    // ---- count strictly better / tied candidates
    long long idx_cut = 0x7fffffffffffffffLL;
    if (!all) {
        int gt = 0, eq = 0;
        for (int e = tid; e < T; e += NT) {
            const uint32_t key = key_of(e);
            gt += (key != 0u && key > prefix); eq += key == prefix;
        }
        atomicAdd(&sh_gt, gt); atomicAdd(&sh_eq, eq);
        __syncthreads();
        const int need = k - sh_gt;                        // ties to keep (1 <= need <= sh_eq)
        if (sh_eq > need) {                                // rare: pick the `need` smallest indices among the ties
            unsigned long long ip = 0; int kn = need;
            for (int pass = 0; pass < 8; ++pass) {
                const int shift = 56 - 8 * pass;
                __syncthreads();
                if (tid < 256) hist[tid] = 0;
                __syncthreads();
                for (int e0 = 0; e0 < T; e0 += NT) {
                    const int e = e0 + tid;
                    bool act = e < T && key_of(e) == prefix;
                    unsigned long long u = 0;
                    if (act) { u = (unsigned long long)id_of(e); act = pass == 0 || (u >> (shift + 8)) == (ip >> (shift + 8)); }
                    hist_add(hist, act, (int)((u >> shift) & 255));
                }
                __syncthreads();
                scan_hist(hist, kn, false, &sh_digit, &sh_before, &sh_total);
                __syncthreads();
                ip |= (unsigned long long)sh_digit << shift;
                kn -= sh_before;
            }
            idx_cut = (long long)ip;                       // the need-th smallest tied index
        }
    }
    __syncthreads();
    // ---- gather survivors, sort them, write
    for (int e = tid; e < T; e += NT) {
        const uint32_t key = key_of(e);
        if (key == 0u) continue;
        if (all || key > prefix || key == prefix) {
            const long long id = id_of(e);
            if (all || key > prefix || id <= idx_cut) {
                const int pos = atomicAdd(&sh_count, 1);
                if (pos < MERGE_KMAX) { ss[pos] = KEYS_IN ? unflip_f32(key) : ps[offset(e)] + 0.0f; si[pos] = id; }
            }
        }
    }
    __syncthreads();
    const int cnt = min(sh_count, k);
    int np2 = 1;
    while (np2 < cnt) np2 <<= 1;
    for (int e = cnt + tid; e < np2; e += NT) { ss[e] = -INFINITY; si[e] = 0x7fffffffffffffffLL; }
    __syncthreads();
    for (int kk2 = 2; kk2 <= np2; kk2 <<= 1)
        for (int j = kk2 >> 1; j > 0; j >>= 1) {
            for (int e = tid; e < np2; e += NT) {
                const int o = e ^ j;
                if (o > e) {
                    const bool desc = (e & kk2) == 0;
                    const float sa = ss[e], sb = ss[o]; const long long ia = si[e], ib = si[o];
                    const bool sw = desc ? better(sb, ib, sa, ia) : better(sa, ia, sb, ib);
                    if (sw) { ss[e] = sb; ss[o] = sa; si[e] = ib; si[o] = ia; }
                }
            }
            __syncthreads();
        }
    for (int e = tid; e < k; e += NT) {
        const bool valid = e < cnt;
        if (out_keys != nullptr) {
            out_keys[(size_t)qi * k + e] = valid ? (((unsigned long long)flip_f32(ss[e]) << 32) | (0xffffffffu - (uint32_t)(si[e] - key_base))) : 0ull;
        } else {
            out_s[(size_t)qi * k + e] = valid ? ss[e] : -INFINITY;
            out_i[(size_t)qi * k + e] = valid ? si[e] : -1;
        }
    }
}

template <bool KEYS_IN>
static int merge_launch(const float* ps, const long long* pi, const unsigned long long* pkeys, const long long* bases, int P, int Q, int k,
                        float* out_s, long long* out_i, unsigned long long* out_keys, long long key_base, cudaStream_t st, const int* counts = nullptr,
                        int slot = 0) {
    B200_REQUIRE(P > 0 && k > 0, "merge_topk: need at least one part and k > 0");
    B200_REQUIRE(k <= MERGE_KMAX, "merge_topk: k larger than 1024 is not supported");
    B200_REQUIRE((long long)P * k < (1LL << 30), "merge_topk: too many candidates per query");
    if (Q == 0) return B200_OK;
    const size_t T = counts != nullptr ? (size_t)slot : (size_t)P * k;     // most candidates a query can have
    B200_REQUIRE(T < (1u << 30), "merge_topk: too many candidates per query");
    // keys are cached in shared memory (next to the sort buffers) when a query's candidates fit; a list of counts[] entries is
    // usually far shorter than its slice, so the kernel decides per query
    const int cache = (int)std::min<size_t>(T, 40 * 1024);
    const size_t smem = (size_t)cache * 4;
    static const int mid_nt = [] { const char* e = getenv("B200_MERGE_MID"); return e ? atoi(e) : 512; }();
    if (Q <= 2 * kNumSMs) {
        if (smem > 32 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(merge_topk_kernel<1024, KEYS_IN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        merge_topk_kernel<1024, KEYS_IN><<<Q, 1024, smem, st>>>(ps, pi, pkeys, bases, P, Q, k, cache, out_s, out_i, out_keys, key_base, counts, slot);
    } else if (Q <= 10 * kNumSMs && mid_nt == 512) {
        // a few CTAs per SM (the key cache bounds the occupancy, not the threads): longer CTAs shorten every pass of the latency chain
        if (smem > 32 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(merge_topk_kernel<512, KEYS_IN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        merge_topk_kernel<512, KEYS_IN><<<Q, 512, smem, st>>>(ps, pi, pkeys, bases, P, Q, k, cache, out_s, out_i, out_keys, key_base, counts, slot);
    } else {
        if (smem > 32 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(merge_topk_kernel<256, KEYS_IN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        merge_topk_kernel<256, KEYS_IN><<<Q, 256, smem, st>>>(ps, pi, pkeys, bases, P, Q, k, cache, out_s, out_i, out_keys, key_base, counts, slot);
    }
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}
int merge_topk_run(const float* ps, const long long* pi, int P, int Q, int k, float* out_s, long long* out_i, cudaStream_t st) {
    return merge_launch<false>(ps, pi, nullptr, nullptr, P, Q, k, out_s, out_i, nullptr, 0, st);
}
// parts as exchange keys [P][Q][k] + the index_base of every part (device array [P]) -> (scores, global indices)
int merge_topk_keys_run(const unsigned long long* pkeys, const long long* bases_dev, int P, int Q, int k, float* out_s, long long* out_i, cudaStream_t st) {
    return merge_launch<true>(nullptr, nullptr, pkeys, bases_dev, P, Q, k, out_s, out_i, nullptr, 0, st);
}

// ------------------------------------------------------------------------------------------ host side
static size_t topk_gtau_bytes(int Q) { return ((size_t)Q * 4 + 255) & ~(size_t)255; }
// gtau [Q], pcount [Q], pubmax [Q][splits]: all start every call at zero (one memset)
static size_t topk_shared_bytes(int Q, int splits) { return 2 * topk_gtau_bytes(Q) + (((size_t)Q * splits * 4 + 255) & ~(size_t)255); }

int topk_plan(int Q, long long N, int k, TopkPlan* plan) {
    B200_REQUIRE(Q > 0 && N > 0 && k > 0, "topk: Q, N and k must be positive");
    B200_REQUIRE(k <= 768, "topk: k larger than 768 is not supported");
    plan->qtiles = (Q + TK_BM - 1) / TK_BM;
    const long long ntiles = (N + TK_BN - 1) / TK_BN;
    // qtiles * splits CTAs: fill one wave of 148 SMs when the query tiles are few, two waves otherwise
    long long splits = kNumSMs / plan->qtiles;
    if (splits * plan->qtiles * 10 < 9LL * kNumSMs) splits = (2LL * kNumSMs) / plan->qtiles;   // one wave must fill >= 90 % of the SMs
    if (splits < 1) splits = 1;
    if (splits > ntiles) splits = ntiles;
    if (splits * k > 16384) splits = 16384 / k;                 // merge kernel capacity
    const long long tiles_per_split = (ntiles + splits - 1) / splits;
    plan->rows_per_split = tiles_per_split * TK_BN;
    plan->splits = (int)((ntiles + tiles_per_split - 1) / tiles_per_split);
    // lists of 512 entries, or 1024 for k > 256 and in the maxima-sharing mode (one query tile, at least k splits): there the first
    // tiles are appended whole (no bound exists yet) and must fit until the shared bound arrives, two or three tiles in
    plan->cpl = (k <= 256 && !(plan->qtiles == 1 && plan->splits >= k)) ? 16 : 32;
    // two scanners per row (eight epilogue warps) pay for batches of one query tile, where per-row bookkeeping dominates
    // (Q = 64: 0.312 -> 0.299 ms on a 1.25 M-row shard); with several query tiles they measured neutral to slightly worse
    static const int halves_env = [] { const char* e = getenv("B200_TOPK_HALVES"); return e ? atoi(e) : 0; }();
    plan->halves = halves_env ? (halves_env == 1 ? 1 : 2) : (plan->qtiles == 1 ? 2 : 1);
    const size_t parts = (size_t)plan->splits * plan->halves;
    // a scanner hands its whole list to the merge kernel when it holds at most `room` entries (a list is compacted as soon as it
    // cannot take another tile, so it ends a scan with at most C - span + span entries, and nearly always far fewer)
    const int C = 32 * plan->cpl, span = TK_BN / plan->halves;
    // Only with several query tiles: a single tile runs with the maxima-sharing bound, which leaves a scanner a handful of entries.
    plan->room = plan->qtiles > 1 ? std::max(k, std::min(C, C - span + 64)) : k;
    plan->cand_bytes = parts * plan->qtiles * TK_BM * 32 * plan->cpl * 8;
    plan->part_bytes = ((parts * Q * plan->room * 4 + 255) & ~(size_t)255) + parts * Q * plan->room * 8;
    plan->part_bytes = ((plan->part_bytes + 255) & ~(size_t)255) + topk_shared_bytes(Q, plan->splits);     // the shared thresholds sit at the end
    plan->qstage_bytes = Q < TK_BM ? (size_t)TK_BM * 1024 * 2 : 0;      // staged, interleaved copy of a small query batch (D <= 1024)
    return B200_OK;
}

// Query i of a batch smaller than one tile goes to tile row (i % 4) * 32 + i / 4, i.e. TMEM lane quarter i % 4, so the
// per-row candidate bookkeeping is shared by all four epilogue warps instead of piling up in the first one.
__global__ void scatter_queries_kernel(const __half* __restrict__ src, __half* __restrict__ dst, int Q, int D) {
    const int i = blockIdx.x;
    const int r = (i & 3) * 32 + (i >> 2);
    for (int c = threadIdx.x * 8; c < D; c += blockDim.x * 8)
        *reinterpret_cast<uint4*>(dst + (size_t)r * D + c) = *reinterpret_cast<const uint4*>(src + (size_t)i * D + c);
}

template <int CPL> static int launch_topk(const CUtensorMap& tq, const CUtensorMap& tg, TopkParams p, cudaStream_t st) {
    static const int qres_on = [] { const char* e = getenv("B200_TOPK_QRES"); return e ? atoi(e) : 1; }();
    const int q_bytes = (p.D / TK_BK) * TK_A_BYTES;
    if (qres_on && p.D <= 512) {
        p.stages = std::min(TK_STAGES, (227 * 1024 - 256 - 1024 - q_bytes) / TK_B_BYTES);
        const int smem = q_bytes + p.stages * TK_B_BYTES + 256 + 1024;
        B200_CUDA_OK(cudaFuncSetAttribute(topk_tc_kernel<CPL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        topk_tc_kernel<CPL, true><<<dim3(p.qtiles, p.splits), TK_THREADS, smem, st>>>(tq, tg, p);
    } else {
        B200_CUDA_OK(cudaFuncSetAttribute(topk_tc_kernel<CPL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM));
        topk_tc_kernel<CPL, false><<<dim3(p.qtiles, p.splits), TK_THREADS, TK_SMEM, st>>>(tq, tg, p);
    }
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
}

// queries fp16 [Q, D], gallery fp16 [N, D] (both row-major, D % 64 == 0).  scratch must hold
// plan.cand_bytes + plan.part_bytes bytes (256-byte aligned).  Outputs: scores fp32 [Q,k], idx int64 [Q,k].
int topk_run(const void* queries, const void* gallery, int Q, long long N, int D, int k, long long index_base,
             void* scratch, const TopkPlan& plan, float* out_s, long long* out_i, DeviceStatus* status, cudaStream_t st,
             unsigned long long* out_keys) {
    B200_REQUIRE(D > 0 && D % 64 == 0, "topk: embedding dim must be a multiple of 64");
    B200_REQUIRE(plan.qstage_bytes == 0 || D <= 1024, "topk: embedding dim above 1024 is not supported for batches below 128 queries");
    B200_REQUIRE(N < (1LL << 31), "topk: at most 2^31-1 gallery rows per GPU");
    CUtensorMap tq, tg;
    uint8_t* sc = reinterpret_cast<uint8_t*>(scratch);
    const int scatter = plan.qstage_bytes > 0 ? 1 : 0;
    int rc;
    if (scatter) {
        __half* staged = reinterpret_cast<__half*>(sc + plan.cand_bytes + plan.part_bytes);
        B200_CUDA_OK(cudaMemsetAsync(staged, 0, plan.qstage_bytes, st));
        scatter_queries_kernel<<<Q, 64, 0, st>>>(reinterpret_cast<const __half*>(queries), staged, Q, D);
        B200_CUDA_OK(cudaGetLastError());
        rc = make_tmap_2d(&tq, staged, TK_BM, D, (uint64_t)D * 2, TK_BM, true);
    } else {
        rc = make_tmap_2d(&tq, queries, Q, D, (uint64_t)D * 2, TK_BM, true);
    }
    if (rc) return rc;
    rc = make_tmap_2d(&tg, gallery, (uint64_t)N, D, (uint64_t)D * 2, TK_BN, true);
    if (rc) return rc;
    float* part_s = reinterpret_cast<float*>(sc + plan.cand_bytes);
    const int parts = plan.splits * plan.halves;
    long long* part_i = reinterpret_cast<long long*>(sc + plan.cand_bytes + (((size_t)parts * Q * plan.room * 4 + 255) & ~(size_t)255));
    unsigned* gtau = reinterpret_cast<unsigned*>(sc + plan.cand_bytes + plan.part_bytes - topk_shared_bytes(Q, plan.splits));
    int* pcount = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(gtau) + topk_gtau_bytes(Q));
    unsigned* pubmax = reinterpret_cast<unsigned*>(reinterpret_cast<uint8_t*>(gtau) + 2 * topk_gtau_bytes(Q));
    B200_CUDA_OK(cudaMemsetAsync(gtau, 0, topk_shared_bytes(Q, plan.splits), st));
    static const int share_on = [] { const char* e = getenv("B200_TOPK_SHAREMAX"); return e ? atoi(e) : 1; }();
    const int share_max = (share_on && plan.qtiles == 1 && plan.splits >= k && plan.splits <= 320) ? 1 : 0;
    TopkParams p{Q, D, k, N, plan.rows_per_split, plan.splits, plan.qtiles, reinterpret_cast<unsigned long long*>(sc),
                 part_s, part_i, index_base, scatter, status, 0, gtau, pubmax, share_max, pcount, plan.room, nullptr, 0, plan.halves};
    static const int seed_on = [] { const char* e = getenv("B200_TOPK_SEED"); return e ? atoi(e) : 1; }();
    if (seed_on && plan.qtiles > 1 && k <= TK_SEED_CHUNKS && N >= 16LL * TK_SEED_ROWS && (size_t)parts * plan.room >= (size_t)TK_SEED_CHUNKS) {
        // (the chunk maxima borrow the front of the per-query result slices, which nothing touches before the scan proper ends)
        TopkParams ps = p;
        ps.N = TK_SEED_ROWS;
        const int seed_tiles = TK_SEED_ROWS / TK_BN;
        int ss = std::max(1, std::min(seed_tiles, kNumSMs / plan.qtiles));
        while (seed_tiles % ss) --ss;
        ps.splits = ss; ps.rows_per_split = (long long)(seed_tiles / ss) * TK_BN;
        ps.share_max = 0;
        ps.chunkmax = reinterpret_cast<unsigned*>(part_s); ps.chunkmax_ld = TK_SEED_CHUNKS;
        rc = plan.cpl == 16 ? launch_topk<16>(tq, tg, ps, st) : launch_topk<32>(tq, tg, ps, st);
        if (rc) return rc;
        seed_tau_kernel<<<(Q + 7) / 8, 256, 0, st>>>(ps.chunkmax, TK_SEED_CHUNKS, TK_SEED_CHUNKS, Q, k, gtau);
        B200_CUDA_OK(cudaGetLastError());
    }
    rc = plan.cpl == 16 ? launch_topk<16>(tq, tg, p, st) : launch_topk<32>(tq, tg, p, st);
    if (rc) return rc;
    // out_keys: exchange form for the sharded search, indices relative to this shard
    return merge_launch<false>(part_s, part_i, nullptr, nullptr, parts, Q, k, out_s, out_i, out_keys, index_base, st, pcount, parts * plan.room);
}

}  // namespace b200

// ------------------------------------------------------------------------------------------ allocation-free search
// Scratch for b200clip_topk_static: lives in the library image, so every GPU that loads libb200clip holds these
// bytes until the process exits.  First 256 bytes: watchdog status words; the rest: candidate lists + per-split results.
namespace {
constexpr size_t kStaticScratchBytes = 32u << 20;
__device__ __align__(256) unsigned char g_topk_static_scratch[kStaticScratchBytes];
std::mutex g_static_mu;
void* g_static_ptr[64] = {nullptr};          // device address of the array, resolved once per device
}  // namespace

#include "../../include/b200clip.h"

extern "C" int b200clip_topk_static(const void* queries_f16_dev, int32_t Q, const void* gallery_f16_dev, int64_t N, int32_t D,
                                    int32_t k, int64_t index_base, float* out_scores_dev, int64_t* out_index_dev, void* stream) {
    using namespace b200;
    B200_REQUIRE(queries_f16_dev && gallery_f16_dev && out_scores_dev && out_index_dev, "topk_static: null buffer");
    B200_REQUIRE(Q > 0 && N > 0 && k > 0, "topk_static: Q, N and k must be positive");
    B200_REQUIRE(k <= 256, "topk_static: k is limited to 256 (the handle-based b200clip_topk goes to 768)");
    int dev = -1;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < 64, "topk_static: unsupported device ordinal");
    void* base = nullptr;
    {
        // One-time set-up per device, with that device current.  In host code the array's name is only a stand-in
        // for the symbol; the address a kernel can use has to be asked for.
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (g_static_ptr[dev] == nullptr) {
            void* ptr = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&ptr, g_topk_static_scratch);
            if (e != cudaSuccess) return fail(B200_ECUDA, std::string("topk_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            g_static_ptr[dev] = ptr;
        }
        base = g_static_ptr[dev];
    }
    // Same work split as b200clip_topk, shrunk until candidate lists + per-split results fit the fixed array.
    TopkPlan plan;
    plan.cpl = 16;
    plan.room = k;                            // per-split results of k slots, as sized below
    plan.halves = 1;                          // four scanning warps: half the list memory of the handle path, so the array holds twice the queries
    plan.qtiles = (Q + TK_BM - 1) / TK_BM;
    const long long ntiles = (N + TK_BN - 1) / TK_BN;
    const size_t budget = kStaticScratchBytes - 256;
    auto need = [&](long long splits) {
        return (size_t)splits * plan.qtiles * TK_BM * 32 * 16 * 8 + (((size_t)splits * Q * k * 4 + 255) & ~(size_t)255) +
               (((size_t)splits * Q * k * 8 + 255) & ~(size_t)255) + topk_shared_bytes(Q, (int)splits) + 256;
    };
    long long splits = (2LL * kNumSMs) / plan.qtiles;
    if (splits < 1) splits = 1;
    if (splits > ntiles) splits = ntiles;
    if (splits * k > 16384) splits = 16384 / k;
    while (splits > 1 && need(splits) > budget) --splits;
    if (need(splits) > budget)
        return fail(B200_EINVAL, "topk_static: Q=" + std::to_string(Q) + ", k=" + std::to_string(k) + " needs " + std::to_string(need(splits)) +
                                 " bytes of scratch, the built-in array holds " + std::to_string(budget) +
                                 " (about 6200 queries at k=100); use b200clip_topk with a handle instead");
    const long long tiles_per_split = (ntiles + splits - 1) / splits;
    plan.rows_per_split = tiles_per_split * TK_BN;
    plan.splits = (int)((ntiles + tiles_per_split - 1) / tiles_per_split);
    plan.cand_bytes = (size_t)plan.splits * plan.qtiles * TK_BM * 32 * 16 * 8;
    plan.part_bytes = (((size_t)plan.splits * Q * k * 4 + 255) & ~(size_t)255) + (((size_t)plan.splits * Q * k * 8 + 255) & ~(size_t)255) +
                      topk_shared_bytes(Q, plan.splits);
    plan.qstage_bytes = 0;
    cudaStream_t st = (cudaStream_t)stream;
    // The array keeps whatever the previous call left.  Only the watchdog words have to start at zero: candidate
    // counts live in registers and every result slot that is read is written first.
    B200_CUDA_OK(cudaMemsetAsync(base, 0, 256, st));
    return topk_run(queries_f16_dev, gallery_f16_dev, Q, N, D, k, index_base, reinterpret_cast<unsigned char*>(base) + 256, plan,
                    out_scores_dev, reinterpret_cast<long long*>(out_index_dev), reinterpret_cast<DeviceStatus*>(base), st, nullptr);
}

extern "C" int b200clip_topk_static_status(void* stream) {
    using namespace b200;
    int dev = -1;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < 64, "topk_static_status: unsupported device ordinal");
    void* base = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        base = g_static_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    B200_CUDA_OK(cudaStreamSynchronize(st));
    if (base == nullptr) return B200_OK;             // b200clip_topk_static has not run on this device yet
    DeviceStatus ds;
    B200_CUDA_OK(cudaMemcpyAsync(&ds, base, sizeof(ds), cudaMemcpyDeviceToHost, st));
    B200_CUDA_OK(cudaStreamSynchronize(st));
    if (ds.code != 0) {
        B200_CUDA_OK(cudaMemsetAsync(base, 0, 256, st));
        return fail(-5, "device watchdog: a pipeline barrier of b200clip_topk_static never completed (site " + std::to_string(ds.where) + ")");
    }
    return B200_OK;
}
