// N2 (SURVEY.md §8 "next"): baseline JPEG decode in front of K1.  Replaces `Image.open(io.BytesIO(b)).convert("RGB")`
// (reference: backends/onnxrt_backend.py:478,412; backends/torch_backend.py:570,635), i.e. Pillow -> libjpeg-turbo with its
// defaults (JDCT_ISLOW, fancy upsampling), bit for bit.
//
// Split: the Huffman bit stream is inherently serial per image, so entropy decoding runs on host threads (one image per
// thread, coefficients written straight into pinned memory); everything with data parallelism runs on the device:
//   jpeg_idct_kernel   dequantise + 8x8 integer IDCT (jidctint.c islow arithmetic) -> component planes, one thread per block
//   jpeg_rgb_kernel    fancy h2v1 / h2v2 chroma upsampling (jdsample.c) + YCbCr -> RGB (jdcolor.c fixed point) -> packed
//                      ragged HWC uint8, exactly the layout b200clip_preprocess (K1) consumes
// Supported: SOF0/SOF1 8-bit Huffman, 1 or 3 components in one scan, 4:4:4 / 4:2:2 / 4:2:0, restart intervals, YCbCr or
// grayscale.  Anything else (progressive, CMYK, 4:1:1, ...) is reported per image so the caller can hand those to its
// host decoder; there is no silent fallback in here.
#include "common.cuh"
#include "kernels.h"

#include <atomic>
#include <cstring>
#include <thread>
#include <vector>

namespace b200 {

namespace {

constexpr uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                 41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// Canonical Huffman table with a 9-bit first-level lookup (most codes of real tables are shorter than that).
struct HuffTable {
    bool present = false;
    uint8_t counts[17] = {0}, symbols[256] = {0};
    int total = 0;                       // number of symbols
    int32_t max_code[18];
    int32_t first_index[17];
    uint32_t first_code[17];
    uint16_t fast[512];                  // (length << 8) | symbol, 0 = not a short code
    int16_t fast_ac[512];                // AC tables: (value << 8) | (run << 4) | (code + value bits) when both fit in 9 bits, else 0

    // Builds the lookup tables; false when the code-length counts do not describe a prefix code (libjpeg's
    // jpeg_make_d_derived_tbl refuses those too): a length that is over-subscribed would index past `fast` / `symbols`.
    bool finish() {
        uint32_t code = 0;
        int k = 0;
        present = false;
        for (int len = 1; len <= 16; ++len) {
            if (code + counts[len] > (1u << len)) return false;
            code = (code + counts[len]) << 1;
            k += counts[len];
        }
        if (k > 256) return false;
        code = 0; k = 0;
        std::memset(fast, 0, sizeof fast);
        for (int len = 1; len <= 16; ++len) {
            first_index[len] = k;
            first_code[len] = code;
            for (int i = 0; i < counts[len]; ++i, ++k, ++code)
                if (len <= 9)
                    for (uint32_t f = code << (9 - len), e = f + (1u << (9 - len)); f < e; ++f) fast[f] = (uint16_t)((len << 8) | symbols[k]);
            max_code[len] = counts[len] ? (int32_t)code - 1 : -1;
            code <<= 1;
        }
        max_code[17] = INT32_MAX;
        // run/size symbol and the coefficient's value bits resolved by ONE 9-bit lookup when they are short enough
        // (most AC coefficients of a photograph are: small magnitudes have short codes)
        for (int i = 0; i < 512; ++i) {
            fast_ac[i] = 0;
            const uint16_t f = fast[i];
            if (!f) continue;
            const int len = f >> 8, rs = f & 0xFF, run = rs >> 4, mag = rs & 15;
            if (mag == 0 || len + mag > 9) continue;
            int v = ((i << len) & 511) >> (9 - mag);
            if (v < (1 << (mag - 1))) v += (int)((~0u) << mag) + 1;
            if (v >= -128 && v <= 127) fast_ac[i] = (int16_t)(v * 256 + run * 16 + len + mag);
        }
        total = k;
        present = true;
        return true;
    }
};

// MSB-first bit reader over an entropy-coded segment: removes FF00 stuffing, stops at any marker and feeds zeros after it.
class BitStream {
public:
    BitStream(const uint8_t* p, const uint8_t* end) : p_(p), end_(end) {}
    inline void need(int n) { if (have_ < n) refill(); }
    inline uint32_t peek(int n) const { return (uint32_t)(acc_ >> (64 - n)); }
    inline void drop(int n) { acc_ <<= n; have_ -= n; }
    inline int bits(int n) { const uint32_t v = peek(n); drop(n); return (int)v; }
    bool ran_off_the_end() const { return off_end_; }       // the data stopped without any marker (no EOI): truncated file
    // byte-align and step over the expected RSTn marker
    bool restart(int expect) {
        acc_ = 0; have_ = 0; at_marker_ = false;
        while (p_ + 1 < end_ && !(p_[0] == 0xFF && p_[1] >= 0xD0 && p_[1] <= 0xD7)) ++p_;
        if (p_ + 1 >= end_ || p_[1] != 0xD0 + expect) return false;
        p_ += 2;
        return true;
    }
private:
    void refill() {
        // fast path: eight plain bytes ahead (no 0xFF anywhere: neither stuffing nor a marker) -> take as many as fit at once
        if (!at_marker_ && end_ - p_ >= 8) {
            uint64_t w;
            std::memcpy(&w, p_, 8);
            const uint64_t inv = ~w;
            if (((inv - 0x0101010101010101ull) & ~inv & 0x8080808080808080ull) == 0) {      // no byte of w equals 0xFF
                const int take = (64 - have_) >> 3;                                          // whole bytes that fit (1..8)
                w = __builtin_bswap64(w);
                if (take == 8) acc_ = w;                                                     // have_ == 0
                else acc_ |= (w >> (64 - 8 * take)) << (64 - have_ - 8 * take);
                p_ += take;
                have_ += 8 * take;
                return;
            }
        }
        while (have_ <= 56) {
            uint64_t byte = 0;
            if (!at_marker_ && p_ < end_) {
                byte = *p_;
                if (byte == 0xFF) {
                    if (p_ + 1 < end_ && p_[1] == 0) p_ += 2;
                    else { at_marker_ = true; byte = 0; }
                } else {
                    ++p_;
                }
            } else {
                if (!at_marker_) off_end_ = true;
                at_marker_ = true;
            }
            acc_ |= byte << (56 - have_);
            have_ += 8;
        }
    }
    const uint8_t *p_, *end_;
    uint64_t acc_ = 0;
    int have_ = 0;
    bool at_marker_ = false, off_end_ = false;
};

inline int decode_symbol(BitStream& bs, const HuffTable& t) {
    bs.need(32);
    const uint16_t f = t.fast[bs.peek(9)];
    if (f) { bs.drop(f >> 8); return f & 0xFF; }
    const uint32_t w = bs.peek(16);
    for (int len = 10; len <= 16; ++len) {
        const int32_t c = (int32_t)(w >> (16 - len));
        if (c <= t.max_code[len]) {
            const int at = t.first_index[len] + (c - (int32_t)t.first_code[len]);
            if (at < 0 || at >= t.total) return -1;
            bs.drop(len);
            return t.symbols[at];
        }
    }
    return -1;
}
inline int sign_extend(int v, int n) { return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v; }

}  // namespace

// What the device kernels need to know about one image.
struct JpegDesc {
    int32_t width, height, ncomp, hmax, vmax;
    int32_t blocks_w[3], blocks_h[3];    // per component, padded to whole MCUs
    int32_t ds_w[3], ds_h[3];            // real (downsampled) component size
    int32_t block_begin[4];              // running block count over the components (block_begin[ncomp] = all blocks of the image)
    int64_t coef_off[3];                 // int16 elements into the coefficient arena
    int64_t plane_off[3];                // bytes into the plane arena
    int64_t dst_off;                     // bytes into the packed RGB output
    uint16_t quant[3][64];               // natural order
};

namespace {

struct Parsed {
    JpegDesc d{};
    int comp_id[3], hs[3], vs[3], tq[3], td[3], ta[3];
    uint16_t qt[4][64];
    bool qt_present[4] = {false, false, false, false};
    HuffTable dc[4], ac[4];
    int restart_interval = 0;
    const uint8_t* scan = nullptr;
    const uint8_t* end = nullptr;
    std::string error;
    int code = 0;                        // 0 ok, 1 unsupported, 2 corrupt
};

int div_up(int a, int b) { return (a + b - 1) / b; }

bool fail_parse(Parsed& P, int code, const char* why) { P.code = code; P.error = why; return false; }

// Marker segments up to and including SOS; fills geometry, tables and the position of the entropy-coded data.
bool parse_headers(const uint8_t* data, size_t len, Parsed& P) {
    const uint8_t *p = data, *end = data + len;
    P.end = end;
    if (len < 4 || p[0] != 0xFF || p[1] != 0xD8) return fail_parse(P, 1, "not a JPEG stream");
    p += 2;
    bool have_frame = false, jfif = false, adobe = false;
    int adobe_transform = 0, ncomp = 0;
    for (;;) {
        while (p < end && *p != 0xFF) ++p;
        while (p < end && *p == 0xFF) ++p;
        if (p >= end) return fail_parse(P, 2, "truncated before the scan");
        const int m = *p++;
        if (m == 0xD9) return fail_parse(P, 2, "EOI before any scan");
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (p + 2 > end) return fail_parse(P, 2, "truncated segment");
        const int L = (p[0] << 8) | p[1];
        if (L < 2 || p + L > end) return fail_parse(P, 2, "truncated segment");
        const uint8_t *s = p + 2, *se = p + L;
        switch (m) {
        case 0xDB:
            while (s < se) {
                const int wide = s[0] >> 4, id = s[0] & 15;
                ++s;
                if (id > 3 || s + (wide ? 128 : 64) > se) return fail_parse(P, 2, "bad DQT");
                for (int i = 0; i < 64; ++i, s += wide ? 2 : 1) P.qt[id][kZigzag[i]] = wide ? (uint16_t)((s[0] << 8) | s[1]) : s[0];
                P.qt_present[id] = true;
            }
            break;
        case 0xC4:
            while (s < se) {
                const int cls = s[0] >> 4, id = s[0] & 15;
                ++s;
                if (cls > 1 || id > 3 || s + 16 > se) return fail_parse(P, 2, "bad DHT");
                HuffTable& t = cls ? P.ac[id] : P.dc[id];
                int total = 0;
                for (int i = 1; i <= 16; ++i) { t.counts[i] = s[i - 1]; total += s[i - 1]; }
                s += 16;
                if (total > 256 || s + total > se) return fail_parse(P, 2, "bad DHT");
                std::memcpy(t.symbols, s, total);
                s += total;
                if (!t.finish()) return fail_parse(P, 2, "bad DHT (code lengths are not a prefix code)");
            }
            break;
        case 0xC0: case 0xC1:
            if (L < 8 || s[0] != 8) return fail_parse(P, 1, "only 8-bit samples");
            P.d.height = (s[1] << 8) | s[2];
            P.d.width = (s[3] << 8) | s[4];
            ncomp = s[5];
            if (P.d.width <= 0 || P.d.height <= 0) return fail_parse(P, 1, "zero-sized frame");
            // Pillow warns above Image.MAX_IMAGE_PIXELS (89,478,485) and raises DecompressionBombError above twice that; a header
            // may claim 65535 x 65535.  Such frames are left to the caller's PIL call, which applies exactly that policy.
            if ((int64_t)P.d.width * P.d.height > 89478485LL) return fail_parse(P, 1, "frame larger than PIL's MAX_IMAGE_PIXELS");
            if (ncomp != 1 && ncomp != 3) return fail_parse(P, 1, "only 1 or 3 components");
            if (L < 8 + 3 * ncomp) return fail_parse(P, 2, "bad SOF");
            for (int c = 0; c < ncomp; ++c) {
                P.comp_id[c] = s[6 + 3 * c]; P.hs[c] = s[7 + 3 * c] >> 4; P.vs[c] = s[7 + 3 * c] & 15; P.tq[c] = s[8 + 3 * c];
                if (P.tq[c] > 3) return fail_parse(P, 2, "bad SOF");
            }
            P.d.ncomp = ncomp;
            have_frame = true;
            break;
        case 0xDD:
            if (L < 4) return fail_parse(P, 2, "bad DRI");
            P.restart_interval = (s[0] << 8) | s[1];
            break;
        case 0xE0: if (L >= 7 && !std::memcmp(s, "JFIF", 5)) jfif = true; break;
        case 0xEE: if (L >= 14 && !std::memcmp(s, "Adobe", 5)) { adobe = true; adobe_transform = s[11]; } break;
        case 0xDA: {
            if (!have_frame) return fail_parse(P, 2, "SOS before SOF");
            if (L < 3) return fail_parse(P, 2, "bad SOS");
            const int ns = s[0];
            if (ns != ncomp || L < 6 + 2 * ns) return fail_parse(P, 1, "components split over several scans");
            for (int i = 0; i < ns; ++i) {
                if (s[1 + 2 * i] != P.comp_id[i]) return fail_parse(P, 1, "scan component order differs from the frame");
                P.td[i] = s[2 + 2 * i] >> 4; P.ta[i] = s[2 + 2 * i] & 15;
                if (P.td[i] > 3 || P.ta[i] > 3 || !P.dc[P.td[i]].present || !P.ac[P.ta[i]].present || !P.qt_present[P.tq[i]])
                    return fail_parse(P, 2, "scan refers to a missing table");
            }
            P.scan = p + L;
            goto headers_done;
        }
        default:
            if (m >= 0xC2 && m <= 0xCF && m != 0xC8 && m != 0xCC) return fail_parse(P, 1, "progressive / lossless / arithmetic-coded JPEG");
            break;
        }
        p += L;
    }
headers_done:
    JpegDesc& d = P.d;
    if (d.ncomp == 1) {
        P.hs[0] = P.vs[0] = 1;
        d.hmax = d.vmax = 1;
    } else {
        if (P.hs[1] != 1 || P.vs[1] != 1 || P.hs[2] != 1 || P.vs[2] != 1) return fail_parse(P, 1, "chroma sampling other than 1x1");
        const bool ok = (P.hs[0] == 1 && P.vs[0] == 1) || (P.hs[0] == 2 && P.vs[0] == 1) || (P.hs[0] == 2 && P.vs[0] == 2);
        if (!ok) return fail_parse(P, 1, "luma sampling other than 1x1, 2x1, 2x2");
        d.hmax = P.hs[0]; d.vmax = P.vs[0];
        if (!jfif) {                                          // colour space as libjpeg's default_decompress_parms decides
            if (adobe) { if (adobe_transform != 1) return fail_parse(P, 1, "Adobe RGB / unknown transform"); }
            else if (P.comp_id[0] == 'R' && P.comp_id[1] == 'G' && P.comp_id[2] == 'B') return fail_parse(P, 1, "RGB-coded JPEG");
        }
    }
    const int mcux = div_up(d.width, 8 * d.hmax), mcuy = div_up(d.height, 8 * d.vmax);
    d.block_begin[0] = 0;
    for (int c = 0; c < d.ncomp; ++c) {
        d.blocks_w[c] = mcux * P.hs[c];
        d.blocks_h[c] = mcuy * P.vs[c];
        d.ds_w[c] = div_up(d.width * P.hs[c], d.hmax);
        d.ds_h[c] = div_up(d.height * P.vs[c], d.vmax);
        d.block_begin[c + 1] = d.block_begin[c] + d.blocks_w[c] * d.blocks_h[c];
        std::memcpy(d.quant[c], P.qt[P.tq[c]], sizeof d.quant[c]);
    }
    return true;
}

// Entropy decoding of the single interleaved scan into `coef` (zero-initialised by the caller), natural coefficient order.
bool decode_entropy(Parsed& P, int16_t* coef) {
    const JpegDesc& d = P.d;
    BitStream bs(P.scan, P.end);
    const int mcux = div_up(d.width, 8 * d.hmax), mcuy = div_up(d.height, 8 * d.vmax);
    int pred[3] = {0, 0, 0}, left = P.restart_interval, rst = 0;
    for (int my = 0; my < mcuy; ++my)
        for (int mx = 0; mx < mcux; ++mx) {
            if (P.restart_interval && left == 0) {
                if (!bs.restart(rst)) return fail_parse(P, 2, "restart marker missing or out of order");
                rst = (rst + 1) & 7;
                pred[0] = pred[1] = pred[2] = 0;
                left = P.restart_interval;
            }
            for (int c = 0; c < d.ncomp; ++c) {
                const HuffTable &dc = P.dc[P.td[c]], &ac = P.ac[P.ta[c]];
                for (int by = 0; by < P.vs[c]; ++by)
                    for (int bx = 0; bx < P.hs[c]; ++bx) {
                        int16_t* blk = coef + d.coef_off[c] + ((int64_t)(my * P.vs[c] + by) * d.blocks_w[c] + (mx * P.hs[c] + bx)) * 64;
                        const int s = decode_symbol(bs, dc);
                        if (s < 0 || s > 15) return fail_parse(P, 2, "bad DC code");
                        bs.need(32);
                        if (s) pred[c] += sign_extend(bs.bits(s), s);
                        blk[0] = (int16_t)pred[c];
                        for (int k = 1; k < 64;) {
                            bs.need(32);
                            const int fa = ac.fast_ac[bs.peek(9)];
                            if (fa) {                         // symbol and value bits in one lookup
                                k += (fa >> 4) & 15;
                                if (k > 63) return fail_parse(P, 2, "AC run past the block");
                                bs.drop(fa & 15);
                                blk[kZigzag[k++]] = (int16_t)(fa >> 8);
                                continue;
                            }
                            const int rs = decode_symbol(bs, ac);
                            if (rs < 0) return fail_parse(P, 2, "bad AC code");
                            const int run = rs >> 4, size = rs & 15;
                            if (size == 0) {
                                if (run != 15) break;         // EOB
                                k += 16;                      // ZRL
                                continue;
                            }
                            k += run;
                            if (k > 63) return fail_parse(P, 2, "AC run past the block");
                            bs.need(32);
                            blk[kZigzag[k]] = (int16_t)sign_extend(bs.bits(size), size);
                            ++k;
                        }
                    }
            }
            if (P.restart_interval) --left;
        }
    if (bs.ran_off_the_end()) return fail_parse(P, 2, "entropy-coded data is truncated (no EOI)");
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ device kernels
namespace {

// jidctint.c constants (CONST_BITS = 13, PASS1_BITS = 2)
constexpr int kC0298 = 2446, kC0390 = 3196, kC0541 = 4433, kC0765 = 6270, kC0899 = 7373, kC1175 = 9633, kC1501 = 12299,
              kC1847 = 15137, kC1961 = 16069, kC2053 = 16819, kC2562 = 20995, kC3072 = 25172;

template <int SHIFT>
__device__ __forceinline__ void idct8(const int (&in)[8], int (&out)[8]) {
    int z1 = (in[2] + in[6]) * kC0541;
    const int t2 = z1 - in[6] * kC1847, t3 = z1 + in[2] * kC0765;
    const int t0 = (in[0] + in[4]) * 8192, t1 = (in[0] - in[4]) * 8192;
    const int e0 = t0 + t3, e3 = t0 - t3, e1 = t1 + t2, e2 = t1 - t2;
    int o0 = in[7], o1 = in[5], o2 = in[3], o3 = in[1];
    z1 = o0 + o3;
    int z2 = o1 + o2, z3 = o0 + o2, z4 = o1 + o3;
    const int z5 = (z3 + z4) * kC1175;
    o0 *= kC0298; o1 *= kC2053; o2 *= kC3072; o3 *= kC1501;
    z1 *= -kC0899; z2 *= -kC2562;
    z3 = z3 * -kC1961 + z5;
    z4 = z4 * -kC0390 + z5;
    o0 += z1 + z3; o1 += z2 + z4; o2 += z2 + z3; o3 += z1 + z4;
    constexpr int R = 1 << (SHIFT - 1);
    out[0] = (e0 + o3 + R) >> SHIFT; out[7] = (e0 - o3 + R) >> SHIFT;
    out[1] = (e1 + o2 + R) >> SHIFT; out[6] = (e1 - o2 + R) >> SHIFT;
    out[2] = (e2 + o1 + R) >> SHIFT; out[5] = (e2 - o1 + R) >> SHIFT;
    out[3] = (e3 + o0 + R) >> SHIFT; out[4] = (e3 - o0 + R) >> SHIFT;
}
// libjpeg's post-IDCT range limit: a 1024-entry table indexed with (x & 1023), centred on +128
__device__ __forceinline__ uint32_t idct_limit(int x) {
    const int i = x & 1023;
    return i < 128 ? i + 128 : i < 512 ? 255 : i < 896 ? 0 : i - 896;
}

// grid = (ceil(max blocks per image / 128), images); one thread per 8x8 block: 128 B of coefficients in, 8 x 8 B out
__global__ void __launch_bounds__(128) jpeg_idct_kernel(const JpegDesc* __restrict__ descs, const int16_t* __restrict__ coef, uint8_t* __restrict__ planes) {
    const JpegDesc& d = descs[blockIdx.y];
    const int b = blockIdx.x * 128 + threadIdx.x;
    if (b >= d.block_begin[d.ncomp]) return;
    const int c = b >= d.block_begin[2] && d.ncomp == 3 ? 2 : b >= d.block_begin[1] && d.ncomp == 3 ? 1 : 0;
    const int lb = b - d.block_begin[c], bx = lb % d.blocks_w[c], by = lb / d.blocks_w[c];
    const uint4* src = reinterpret_cast<const uint4*>(coef + d.coef_off[c] + (int64_t)lb * 64);
    const uint16_t* q = d.quant[c];
    int ws[8][8];
    {
        int rows[8][8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const uint4 v = __ldg(src + r);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                rows[r][2 * i] = (int)(int16_t)(w[i] & 0xFFFF) * (int)q[r * 8 + 2 * i];
                rows[r][2 * i + 1] = (int)(int16_t)(w[i] >> 16) * (int)q[r * 8 + 2 * i + 1];
            }
        }
#pragma unroll
        for (int col = 0; col < 8; ++col) {
            int in[8], out[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) in[r] = rows[r][col];
            idct8<11>(in, out);
#pragma unroll
            for (int r = 0; r < 8; ++r) ws[r][col] = out[r];
        }
    }
    uint8_t* dst = planes + d.plane_off[c] + (int64_t)by * 8 * (d.blocks_w[c] * 8) + bx * 8;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int out[8];
        idct8<18>(ws[r], out);
        uint2 px;
        px.x = idct_limit(out[0]) | (idct_limit(out[1]) << 8) | (idct_limit(out[2]) << 16) | (idct_limit(out[3]) << 24);
        px.y = idct_limit(out[4]) | (idct_limit(out[5]) << 8) | (idct_limit(out[6]) << 16) | (idct_limit(out[7]) << 24);
        *reinterpret_cast<uint2*>(dst + (int64_t)r * (d.blocks_w[c] * 8)) = px;
    }
}

// jdsample.c: value of the fancy-upsampled chroma plane at full-resolution (x, y)
__device__ __forceinline__ int chroma_sample(const uint8_t* __restrict__ pl, int stride, int ds_w, int ds_h, int hmax, int vmax, int x, int y) {
    if (hmax == 1) return pl[(int64_t)y * stride + x];
    const int cx = x >> 1;
    if (vmax == 1) {                                          // h2v1
        const uint8_t* row = pl + (int64_t)y * stride;
        if (ds_w <= 2) return row[cx];
        if (x & 1) return cx == ds_w - 1 ? row[cx] : (3 * row[cx] + row[cx + 1] + 2) >> 2;
        return cx == 0 ? row[0] : (3 * row[cx] + row[cx - 1] + 1) >> 2;
    }
    const int cy = y >> 1;                                    // h2v2
    if (ds_w <= 2) return pl[(int64_t)cy * stride + cx];
    const int ny = min(max((y & 1) ? cy + 1 : cy - 1, 0), ds_h - 1);   // the nearer neighbour row, replicated at the image edge
    const uint8_t *r0 = pl + (int64_t)cy * stride, *r1 = pl + (int64_t)ny * stride;
    const int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) return cx == ds_w - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    return cx == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}
__device__ __forceinline__ uint32_t clamp255(int v) { return (uint32_t)min(max(v, 0), 255); }

// grid = (ceil(max pixels per image / (256 * 4)), images); one thread = 4 consecutive pixels of a row (12 output bytes)
__global__ void __launch_bounds__(256) jpeg_rgb_kernel(const JpegDesc* __restrict__ descs, const uint8_t* __restrict__ planes, uint8_t* __restrict__ dst) {
    const JpegDesc& d = descs[blockIdx.y];
    const int groups_per_row = (d.width + 3) >> 2;
    const int gidx = blockIdx.x * 256 + threadIdx.x;
    if (gidx >= groups_per_row * d.height) return;
    const int y = gidx / groups_per_row, x0 = (gidx - y * groups_per_row) * 4;
    const uint8_t* py = planes + d.plane_off[0] + (int64_t)y * (d.blocks_w[0] * 8);
    uint8_t* out = dst + d.dst_off + ((int64_t)y * d.width + x0) * 3;
    const int n = min(4, d.width - x0);
    uint8_t px[12];
    for (int i = 0; i < n; ++i) {
        const int x = x0 + i, Y = py[x];
        if (d.ncomp == 1) { px[3 * i] = px[3 * i + 1] = px[3 * i + 2] = (uint8_t)Y; continue; }
        const int cb = chroma_sample(planes + d.plane_off[1], d.blocks_w[1] * 8, d.ds_w[1], d.ds_h[1], d.hmax, d.vmax, x, y) - 128;
        const int cr = chroma_sample(planes + d.plane_off[2], d.blocks_w[2] * 8, d.ds_w[2], d.ds_h[2], d.hmax, d.vmax, x, y) - 128;
        // jdcolor.c build_ycc_rgb_table: FIX(1.40200) = 91881, FIX(1.77200) = 116130, FIX(0.71414) = 46802, FIX(0.34414) = 22554
        px[3 * i] = (uint8_t)clamp255(Y + ((91881 * cr + 32768) >> 16));
        px[3 * i + 1] = (uint8_t)clamp255(Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
        px[3 * i + 2] = (uint8_t)clamp255(Y + ((116130 * cb + 32768) >> 16));
    }
    if (n == 4 && ((reinterpret_cast<uintptr_t>(out) & 3) == 0)) {
        uint32_t* o32 = reinterpret_cast<uint32_t*>(out);
        o32[0] = px[0] | (px[1] << 8) | (px[2] << 16) | ((uint32_t)px[3] << 24);
        o32[1] = px[4] | (px[5] << 8) | (px[6] << 16) | ((uint32_t)px[7] << 24);
        o32[2] = px[8] | (px[9] << 8) | (px[10] << 16) | ((uint32_t)px[11] << 24);
    } else {
        for (int i = 0; i < 3 * n; ++i) out[i] = px[i];
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host driver
struct JpegState {
    int16_t* coef_host = nullptr; size_t coef_host_cap = 0;      // pinned
    JpegDesc* desc_host = nullptr; size_t desc_host_cap = 0;     // pinned
    int16_t* coef_dev = nullptr; size_t coef_dev_cap = 0;
    uint8_t* plane_dev = nullptr; size_t plane_dev_cap = 0;
    JpegDesc* desc_dev = nullptr; size_t desc_dev_cap = 0;
    cudaEvent_t uploaded = nullptr;                              // the pinned arenas are free again once this has fired
};

JpegState* jpeg_state_create(int* rc) {
    auto* s = new JpegState();
    if (cudaEventCreateWithFlags(&s->uploaded, cudaEventDisableTiming) != cudaSuccess) { *rc = fail(B200_ECUDA, "jpeg: cudaEventCreate failed"); delete s; return nullptr; }
    *rc = B200_OK;
    return s;
}
void jpeg_state_destroy(JpegState* s) {
    if (!s) return;
    cudaFreeHost(s->coef_host); cudaFreeHost(s->desc_host);
    cudaFree(s->coef_dev); cudaFree(s->plane_dev); cudaFree(s->desc_dev);
    if (s->uploaded) cudaEventDestroy(s->uploaded);
    delete s;
}

int jpeg_probe(const void* buf, size_t len, int32_t* hw, int32_t* ncomp) {
    Parsed P;
    if (!parse_headers(static_cast<const uint8_t*>(buf), len, P)) return fail(P.code == 1 ? B200_EINVAL : B200_ESTATE, "jpeg: " + P.error);
    hw[0] = P.d.height; hw[1] = P.d.width; *ncomp = P.d.ncomp;
    return B200_OK;
}

template <class T> static int grow_dev(T** p, size_t* cap, size_t need) {
    if (need <= *cap) return B200_OK;
    if (*p) { B200_CUDA_OK(cudaDeviceSynchronize()); B200_CUDA_OK(cudaFree(*p)); *p = nullptr; *cap = 0; }
    const size_t n = need + need / 4 + 256;
    B200_CUDA_OK(cudaMalloc(reinterpret_cast<void**>(p), n * sizeof(T)));
    *cap = n;
    return B200_OK;
}
template <class T> static int grow_host(T** p, size_t* cap, size_t need) {
    if (need <= *cap) return B200_OK;
    if (*p) { B200_CUDA_OK(cudaFreeHost(*p)); *p = nullptr; *cap = 0; }
    const size_t n = need + need / 4 + 256;
    B200_CUDA_OK(cudaHostAlloc(reinterpret_cast<void**>(p), n * sizeof(T), cudaHostAllocDefault));
    *cap = n;
    return B200_OK;
}

// status_out[i]: 0 decoded; 1 = left for the caller's host decoder (hw_out = 0): a flavour this decoder does not cover, or a stream it
// found damaged (bad tables, bad codes, missing restart markers, no EOI).  Pillow / libjpeg decode many slightly damaged files with a
// warning and refuse others, so that verdict stays with the reference's own PIL call, per image: one bad file never fails the batch.
int jpeg_decode_batch(JpegState* s, const void* const* bufs, const size_t* lens, int n, void* dst, size_t dst_capacity,
                      int64_t* offsets_out, int32_t* hw_out, int32_t* status_out, int64_t* total_out, int threads, cudaStream_t st, int* launches) {
    *launches = 0;
    *total_out = 0;
    if (n <= 0) return B200_OK;
    std::vector<Parsed> parsed(n);
    size_t coef_total = 0, plane_total = 0, dst_total = 0;
    int max_blocks = 0, max_groups = 0, ok_count = 0;
    for (int i = 0; i < n; ++i) {
        Parsed& P = parsed[i];
        if (bufs[i] == nullptr || !parse_headers(static_cast<const uint8_t*>(bufs[i]), lens[i], P)) {
            status_out[i] = 1; offsets_out[i] = 0; hw_out[2 * i] = hw_out[2 * i + 1] = 0;
            continue;
        }
        JpegDesc& d = P.d;
        for (int c = 0; c < d.ncomp; ++c) {
            const size_t blocks = (size_t)d.blocks_w[c] * d.blocks_h[c];
            d.coef_off[c] = (int64_t)coef_total; coef_total += blocks * 64;
            d.plane_off[c] = (int64_t)plane_total; plane_total += (blocks * 64 + 15) & ~(size_t)15;
        }
        d.dst_off = (int64_t)dst_total;
        offsets_out[i] = d.dst_off; hw_out[2 * i] = d.height; hw_out[2 * i + 1] = d.width; status_out[i] = 0;
        dst_total += ((size_t)d.width * d.height * 3 + 15) & ~(size_t)15;
        max_blocks = std::max(max_blocks, d.block_begin[d.ncomp]);
        max_groups = std::max(max_groups, ((d.width + 3) / 4) * d.height);
        ++ok_count;
    }
    *total_out = (int64_t)dst_total;
    if (ok_count == 0) return B200_OK;
    B200_REQUIRE(dst != nullptr && dst_total <= dst_capacity, "jpeg: destination buffer too small for the decoded batch");
    // the previous call's uploads must have left the pinned arenas
    B200_CUDA_OK(cudaEventSynchronize(s->uploaded));
    int rc;
    if ((rc = grow_host(&s->coef_host, &s->coef_host_cap, coef_total))) return rc;
    if ((rc = grow_host(&s->desc_host, &s->desc_host_cap, (size_t)n))) return rc;
    if ((rc = grow_dev(&s->coef_dev, &s->coef_dev_cap, coef_total))) return rc;
    if ((rc = grow_dev(&s->plane_dev, &s->plane_dev_cap, plane_total))) return rc;
    if ((rc = grow_dev(&s->desc_dev, &s->desc_dev_cap, (size_t)n))) return rc;

    // entropy decoding: one image per task, tasks pulled in index order by a small pool of host threads.  The calling thread
    // follows behind them and uploads the coefficients of finished images in runs of >= 8 MB, so the H2D copy (2 bytes per
    // padded sample: as large as the decoded image) overlaps the decoding instead of following it.
    std::atomic<int> next{0};
    std::vector<unsigned char> damaged(n, 0);                     // written by the worker that owns image i, read after the join
    std::vector<std::atomic<unsigned char>> done(n);
    for (int i = 0; i < n; ++i) done[i].store(status_out[i] != 0 ? 1 : 0, std::memory_order_relaxed);
    auto worker = [&] {
        for (int i; (i = next.fetch_add(1)) < n;) {
            Parsed& P = parsed[i];
            if (status_out[i] != 0) continue;
            const JpegDesc& d = P.d;
            std::memset(s->coef_host + d.coef_off[0], 0, (size_t)d.block_begin[d.ncomp] * 64 * sizeof(int16_t));
            if (!decode_entropy(P, s->coef_host)) damaged[i] = 1;
            done[i].store(1, std::memory_order_release);
        }
    };
    // coefficient range [elements) of image i inside the arena (images the decoder does not cover own none)
    auto coef_end = [&](int i) { const JpegDesc& d = parsed[i].d; return (size_t)d.coef_off[0] + (size_t)d.block_begin[d.ncomp] * 64; };
    size_t sent = 0;                                              // arena elements already queued for upload
    cudaError_t copy_err = cudaSuccess;
    auto upload_upto = [&](size_t upto) {
        if (upto > sent && copy_err == cudaSuccess)
            copy_err = cudaMemcpyAsync(s->coef_dev + sent, s->coef_host + sent, (upto - sent) * sizeof(int16_t), cudaMemcpyHostToDevice, st);
        if (upto > sent) sent = upto;
    };
    int T = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    T = std::max(1, std::min(std::min(T, 32), ok_count));
    if (T == 1) {
        worker();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < T; ++t) pool.emplace_back(worker);
        size_t ready = 0;
        for (int i = 0; i < n; ++i) {
            while (!done[i].load(std::memory_order_acquire)) std::this_thread::yield();
            if (status_out[i] == 0) ready = coef_end(i);
            if ((ready - sent) * sizeof(int16_t) >= (8u << 20)) upload_upto(ready);
        }
        for (auto& th : pool) th.join();
    }
    for (int i = 0; i < n; ++i)
        if (damaged[i]) { status_out[i] = 1; offsets_out[i] = 0; hw_out[2 * i] = hw_out[2 * i + 1] = 0; }   // its slot in dst stays unused
    for (int i = 0; i < n; ++i) {
        s->desc_host[i] = parsed[i].d;
        if (status_out[i] != 0) { s->desc_host[i].ncomp = 0; s->desc_host[i].block_begin[0] = 0; s->desc_host[i].width = s->desc_host[i].height = 0; }
    }
    upload_upto(coef_total);
    B200_CUDA_OK(copy_err);
    B200_CUDA_OK(cudaMemcpyAsync(s->desc_dev, s->desc_host, (size_t)n * sizeof(JpegDesc), cudaMemcpyHostToDevice, st));
    B200_CUDA_OK(cudaEventRecord(s->uploaded, st));
    jpeg_idct_kernel<<<dim3((max_blocks + 127) / 128, n), 128, 0, st>>>(s->desc_dev, s->coef_dev, s->plane_dev);
    B200_CUDA_OK(cudaGetLastError());
    jpeg_rgb_kernel<<<dim3((max_groups + 255) / 256, n), 256, 0, st>>>(s->desc_dev, s->plane_dev, static_cast<uint8_t*>(dst));
    B200_CUDA_OK(cudaGetLastError());
    *launches = 2;
    return B200_OK;
}

}  // namespace b200
