/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing in the product path may link, import or call this.
 *
 * CPU restatement (plain C) of the JPEG decode the reference performs with
 *   `Image.open(io.BytesIO(image_bytes)).convert("RGB")`
 *   reference call sites: packages/lumen-clip/src/lumen_clip/backends/onnxrt_backend.py:478,412 and
 *   backends/torch_backend.py:570,635  (SURVEY.md §8 row P0 / "next" row N2).
 *
 * The algorithm lives in third-party dependencies that are NOT under /root/reference: Pillow (pinned 12.0.0 in the
 * reference's uv.lock; 12.2.0 installed here) calling libjpeg-turbo with its defaults (dct_method = JDCT_ISLOW,
 * do_fancy_upsampling = TRUE, no draft mode).  Restated here from the published libjpeg(-turbo) algorithm for
 * baseline sequential 8-bit JPEGs:
 *   - jdhuff.c    Huffman entropy decoding (DC differences, AC run/size, EOB, ZRL, restart intervals)
 *   - jidctint.c  jpeg_idct_islow: dequantise, 8x8 integer IDCT, CONST_BITS = 13, PASS1_BITS = 2, range limit with the
 *                 10-bit wrapped table (sample_range_limit + CENTERJSAMPLE, index & 1023)
 *   - jdsample.c  h2v1_fancy_upsample / h2v2_fancy_upsample (triangle filter, the 1,2 / 8,7 rounding pattern, first and
 *                 last column special cases, top/bottom rows replicated as jdmainct.c's context pointers do); components
 *                 with downsampled_width <= 2 use plain replication as jinit_upsampler decides
 *   - jdcolor.c   ycc_rgb_convert with the 16-bit fixed-point tables (FIX(1.40200), FIX(1.77200), FIX(0.71414),
 *                 FIX(0.34414), ONE_HALF folded as in build_ycc_rgb_table); grayscale replicated to RGB
 * Supported: SOF0/SOF1 (baseline / extended sequential Huffman, 8 bit), 1 or 3 components in one scan, sampling 1x1,
 * 2x1 (4:2:2) and 2x2 (4:2:0) with 1x1 chroma, JFIF / Adobe-transform-1 / id-(1,2,3) YCbCr.  Everything else returns
 * a negative code (the product raises for those too).
 * Parity of this restatement is PINNED against Pillow itself: tests/golden/jpeg.npz (made by tests/golden/make_golden_jpeg.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { J_OK = 0, J_ETRUNC = -1, J_EUNSUPPORTED = -2, J_EBAD = -3, J_ECAP = -4 };

static const uint8_t ZZ[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                               41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                               30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

typedef struct {
    int present;
    uint8_t bits[17], vals[256];
    int32_t maxcode[18], valptr[17];
    uint16_t mincode[17];
    uint8_t look_n[512], look_v[512];   /* 9-bit lookahead: code length (0 = longer) and symbol */
} HuffTab;

static void huff_build(HuffTab *h) {
    int code = 0, k = 0;
    memset(h->look_n, 0, sizeof h->look_n);
    for (int l = 1; l <= 16; ++l) {
        h->valptr[l] = k;
        h->mincode[l] = (uint16_t)code;
        for (int i = 0; i < h->bits[l]; ++i, ++k, ++code) {
            if (l <= 9) {
                int lo = code << (9 - l), n = 1 << (9 - l);
                for (int j = 0; j < n; ++j) { h->look_n[lo + j] = (uint8_t)l; h->look_v[lo + j] = h->vals[k]; }
            }
        }
        h->maxcode[l] = h->bits[l] ? code - 1 : -1;
        code <<= 1;
    }
    h->maxcode[17] = 0x7fffffff;
}

typedef struct {
    const uint8_t *p, *end;
    uint64_t acc;
    int nbits;
    int marker;                          /* a marker was hit: feed zeros */
    int off_end;                         /* the data stopped without any marker (no EOI): truncated file */
} BitReader;

static void br_fill(BitReader *b) {
    while (b->nbits <= 56) {
        uint32_t byte = 0;
        if (!b->marker && b->p < b->end) {
            byte = *b->p;
            if (byte == 0xFF) {
                if (b->p + 1 < b->end && b->p[1] == 0x00) b->p += 2;
                else { b->marker = 1; byte = 0; }
            } else {
                b->p += 1;
            }
        } else {
            if (!b->marker) b->off_end = 1;
            b->marker = 1;
        }
        b->acc |= (uint64_t)byte << (56 - b->nbits);
        b->nbits += 8;
    }
}
static inline uint32_t br_peek(BitReader *b, int n) { return (uint32_t)(b->acc >> (64 - n)); }
static inline void br_skip(BitReader *b, int n) { b->acc <<= n; b->nbits -= n; }
static inline int br_get(BitReader *b, int n) {
    if (n == 0) return 0;
    uint32_t v = br_peek(b, n);
    br_skip(b, n);
    return (int)v;
}
static inline int huff_extend(int v, int n) { return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v; }

static int huff_decode(BitReader *b, const HuffTab *h) {
    if (b->nbits < 32) br_fill(b);
    uint32_t look = br_peek(b, 9);
    int n = h->look_n[look];
    if (n) { br_skip(b, n); return h->look_v[look]; }
    uint32_t code = br_peek(b, 16);
    for (int l = 10; l <= 16; ++l) {
        int32_t c = (int32_t)(code >> (16 - l));
        if (c <= h->maxcode[l]) { br_skip(b, l); return h->vals[h->valptr[l] + c - h->mincode[l]]; }
    }
    return -1;
}

typedef struct {
    int width, height, ncomp;
    int id[3], hs[3], vs[3], tq[3], td[3], ta[3];
    int hmax, vmax;
    int blocks_w[3], blocks_h[3];        /* padded to whole MCUs */
    int ds_w[3], ds_h[3];                /* downsampled (real) component size */
    uint16_t quant[4][64];               /* natural order */
    int quant_present[4];
    HuffTab dc[4], ac[4];
    int restart_interval;
    int saw_jfif, saw_adobe, adobe_transform;
    int16_t *coef[3];                    /* [blocks_h][blocks_w][64] natural order */
} Jpeg;

static int div_up(int a, int b) { return (a + b - 1) / b; }

static int decode_scan(Jpeg *j, const uint8_t *p, const uint8_t *end) {
    BitReader b = {p, end, 0, 0, 0, 0};
    const int mcux = div_up(j->width, 8 * j->hmax), mcuy = div_up(j->height, 8 * j->vmax);
    int pred[3] = {0, 0, 0}, todo = j->restart_interval, next_rst = 0;
    for (int my = 0; my < mcuy; ++my)
        for (int mx = 0; mx < mcux; ++mx) {
            if (j->restart_interval && todo == 0) {
                /* byte-align, expect RSTn */
                b.acc = 0; b.nbits = 0; b.marker = 0;
                while (b.p + 1 < b.end && !(b.p[0] == 0xFF && b.p[1] >= 0xD0 && b.p[1] <= 0xD7)) b.p++;
                if (b.p + 1 >= b.end) return J_ETRUNC;
                if (b.p[1] != 0xD0 + next_rst) return J_EBAD;
                b.p += 2;
                next_rst = (next_rst + 1) & 7;
                pred[0] = pred[1] = pred[2] = 0;
                todo = j->restart_interval;
            }
            for (int c = 0; c < j->ncomp; ++c) {
                const int hs = j->ncomp == 1 ? 1 : j->hs[c], vs = j->ncomp == 1 ? 1 : j->vs[c];
                for (int by = 0; by < vs; ++by)
                    for (int bx = 0; bx < hs; ++bx) {
                        int16_t *blk = j->coef[c] + ((size_t)(my * vs + by) * j->blocks_w[c] + (mx * hs + bx)) * 64;
                        int s = huff_decode(&b, &j->dc[j->td[c]]);
                        if (s < 0 || s > 15) return J_EBAD;
                        if (b.nbits < 32) br_fill(&b);
                        int diff = s ? huff_extend(br_get(&b, s), s) : 0;
                        pred[c] += diff;
                        blk[0] = (int16_t)pred[c];
                        for (int k = 1; k < 64;) {
                            int rs = huff_decode(&b, &j->ac[j->ta[c]]);
                            if (rs < 0) return J_EBAD;
                            int r = rs >> 4, sz = rs & 15;
                            if (sz == 0) {
                                if (r != 15) break;
                                k += 16;
                                continue;
                            }
                            k += r;
                            if (k > 63) return J_EBAD;
                            if (b.nbits < 32) br_fill(&b);
                            blk[ZZ[k]] = (int16_t)huff_extend(br_get(&b, sz), sz);
                            ++k;
                        }
                    }
            }
            if (j->restart_interval) --todo;
        }
    return b.off_end ? J_ETRUNC : J_OK;
}

static int parse(Jpeg *j, const uint8_t *d, size_t len) {
    const uint8_t *p = d, *end = d + len;
    if (len < 4 || p[0] != 0xFF || p[1] != 0xD8) return J_EUNSUPPORTED;   /* not a JPEG stream */
    p += 2;
    int have_sof = 0;
    for (;;) {
        while (p < end && *p != 0xFF) ++p;
        while (p < end && *p == 0xFF) ++p;
        if (p >= end) return J_ETRUNC;
        int m = *p++;
        if (m == 0xD9) return J_EBAD;                         /* EOI before SOS */
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (p + 2 > end) return J_ETRUNC;
        int L = (p[0] << 8) | p[1];
        if (L < 2 || p + L > end) return J_ETRUNC;
        const uint8_t *s = p + 2, *se = p + L;
        if (m == 0xDB) {                                      /* DQT */
            while (s < se) {
                int pq = s[0] >> 4, tq = s[0] & 15;
                ++s;
                if (tq > 3 || s + (pq ? 128 : 64) > se) return J_EBAD;
                for (int i = 0; i < 64; ++i) {
                    j->quant[tq][ZZ[i]] = pq ? (uint16_t)((s[0] << 8) | s[1]) : s[0];
                    s += pq ? 2 : 1;
                }
                j->quant_present[tq] = 1;
            }
        } else if (m == 0xC4) {                               /* DHT */
            while (s < se) {
                int tc = s[0] >> 4, th = s[0] & 15;
                ++s;
                if (tc > 1 || th > 3 || s + 16 > se) return J_EBAD;
                HuffTab *h = tc ? &j->ac[th] : &j->dc[th];
                int n = 0;
                h->bits[0] = 0;
                for (int i = 1; i <= 16; ++i) { h->bits[i] = s[i - 1]; n += s[i - 1]; }
                s += 16;
                if (n > 256 || s + n > se) return J_EBAD;
                memcpy(h->vals, s, n);
                s += n;
                huff_build(h);
                h->present = 1;
            }
        } else if (m == 0xC0 || m == 0xC1) {                  /* SOF0 / SOF1 */
            if (L < 8 || s[0] != 8) return J_EUNSUPPORTED;
            j->height = (s[1] << 8) | s[2];
            j->width = (s[3] << 8) | s[4];
            j->ncomp = s[5];
            if (j->width <= 0 || j->height <= 0) return J_EUNSUPPORTED;
            if (j->ncomp != 1 && j->ncomp != 3) return J_EUNSUPPORTED;
            if (L < 8 + 3 * j->ncomp) return J_EBAD;
            for (int c = 0; c < j->ncomp; ++c) {
                j->id[c] = s[6 + 3 * c];
                j->hs[c] = s[7 + 3 * c] >> 4;
                j->vs[c] = s[7 + 3 * c] & 15;
                j->tq[c] = s[8 + 3 * c];
                if (j->tq[c] > 3) return J_EBAD;
            }
            have_sof = 1;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            return J_EUNSUPPORTED;                            /* progressive, lossless, arithmetic */
        } else if (m == 0xDD) {
            if (L < 4) return J_EBAD;
            j->restart_interval = (s[0] << 8) | s[1];
        } else if (m == 0xE0) {
            if (L >= 7 && !memcmp(s, "JFIF", 5)) j->saw_jfif = 1;
        } else if (m == 0xEE) {
            if (L >= 14 && !memcmp(s, "Adobe", 5)) { j->saw_adobe = 1; j->adobe_transform = s[11]; }
        } else if (m == 0xDA) {                               /* SOS */
            if (!have_sof) return J_EBAD;
            int ns = s[0];
            if (ns != j->ncomp || L < 6 + 2 * ns) return J_EUNSUPPORTED;
            for (int i = 0; i < ns; ++i) {
                if (s[1 + 2 * i] != j->id[i]) return J_EUNSUPPORTED;
                j->td[i] = s[2 + 2 * i] >> 4;
                j->ta[i] = s[2 + 2 * i] & 15;
                if (j->td[i] > 3 || j->ta[i] > 3 || !j->dc[j->td[i]].present || !j->ac[j->ta[i]].present) return J_EBAD;
                if (!j->quant_present[j->tq[i]]) return J_EBAD;
            }
            p += L;
            break;
        }
        p += L;
    }
    /* geometry */
    if (j->ncomp == 1) {
        j->hs[0] = j->vs[0] = 1;
        j->hmax = j->vmax = 1;
    } else {
        if (j->hs[1] != 1 || j->vs[1] != 1 || j->hs[2] != 1 || j->vs[2] != 1) return J_EUNSUPPORTED;
        if (!((j->hs[0] == 1 && j->vs[0] == 1) || (j->hs[0] == 2 && j->vs[0] == 1) || (j->hs[0] == 2 && j->vs[0] == 2)))
            return J_EUNSUPPORTED;
        j->hmax = j->hs[0];
        j->vmax = j->vs[0];
        /* colour space as jdapimin.c default_decompress_parms decides; only YCbCr is restated */
        if (!j->saw_jfif) {
            if (j->saw_adobe) { if (j->adobe_transform != 1) return J_EUNSUPPORTED; }
            else if (j->id[0] == 'R' && j->id[1] == 'G' && j->id[2] == 'B') return J_EUNSUPPORTED;
        }
    }
    const int mcux = div_up(j->width, 8 * j->hmax), mcuy = div_up(j->height, 8 * j->vmax);
    for (int c = 0; c < j->ncomp; ++c) {
        j->blocks_w[c] = mcux * j->hs[c];
        j->blocks_h[c] = mcuy * j->vs[c];
        j->ds_w[c] = div_up(j->width * j->hs[c], j->hmax);
        j->ds_h[c] = div_up(j->height * j->vs[c], j->vmax);
        j->coef[c] = (int16_t *)calloc((size_t)j->blocks_w[c] * j->blocks_h[c] * 64, sizeof(int16_t));
        if (!j->coef[c]) return J_ECAP;
    }
    return decode_scan(j, p, end);
}

/* ---- jidctint.c: jpeg_idct_islow ------------------------------------------------------------------------------ */
#define CONST_BITS 13
#define PASS1_BITS 2
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
#define FIX_0_298631336 2446
#define FIX_0_390180644 3196
#define FIX_0_541196100 4433
#define FIX_0_765366865 6270
#define FIX_0_899976223 7373
#define FIX_1_175875602 9633
#define FIX_1_501321110 12299
#define FIX_1_847759065 15137
#define FIX_1_961570560 16069
#define FIX_2_053119869 16819
#define FIX_2_562915447 20995
#define FIX_3_072711026 25172

/* sample_range_limit + CENTERJSAMPLE, indexed with (x & 1023) */
static inline uint8_t range_limit_idct(int x) {
    int i = x & 1023;
    return (uint8_t)(i < 128 ? i + 128 : i < 512 ? 255 : i < 896 ? 0 : i - 896);
}

static void idct_1d(const int32_t in[8], int32_t out[8], int shift) {
    int32_t z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
    z2 = in[2]; z3 = in[6];
    z1 = (z2 + z3) * FIX_0_541196100;
    tmp2 = z1 + z3 * (-FIX_1_847759065);
    tmp3 = z1 + z2 * FIX_0_765366865;
    z2 = in[0]; z3 = in[4];
    tmp0 = (z2 + z3) * (1 << CONST_BITS);
    tmp1 = (z2 - z3) * (1 << CONST_BITS);
    tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
    tmp0 = in[7]; tmp1 = in[5]; tmp2 = in[3]; tmp3 = in[1];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
    z5 = (z3 + z4) * FIX_1_175875602;
    tmp0 *= FIX_0_298631336; tmp1 *= FIX_2_053119869; tmp2 *= FIX_3_072711026; tmp3 *= FIX_1_501321110;
    z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447; z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    out[0] = DESCALE(tmp10 + tmp3, shift); out[7] = DESCALE(tmp10 - tmp3, shift);
    out[1] = DESCALE(tmp11 + tmp2, shift); out[6] = DESCALE(tmp11 - tmp2, shift);
    out[2] = DESCALE(tmp12 + tmp1, shift); out[5] = DESCALE(tmp12 - tmp1, shift);
    out[3] = DESCALE(tmp13 + tmp0, shift); out[4] = DESCALE(tmp13 - tmp0, shift);
}

static void idct_block(const int16_t *coef, const uint16_t *q, uint8_t *dst, int stride) {
    int32_t ws[64], col[8], res[8];
    for (int c = 0; c < 8; ++c) {
        for (int r = 0; r < 8; ++r) col[r] = (int32_t)coef[r * 8 + c] * (int32_t)q[r * 8 + c];
        idct_1d(col, res, CONST_BITS - PASS1_BITS);
        for (int r = 0; r < 8; ++r) ws[r * 8 + c] = res[r];
    }
    for (int r = 0; r < 8; ++r) {
        idct_1d(ws + r * 8, res, CONST_BITS + PASS1_BITS + 3);
        for (int c = 0; c < 8; ++c) dst[r * stride + c] = range_limit_idct(res[c]);
    }
}

/* ---- jdsample.c + jdcolor.c ------------------------------------------------------------------------------------- */
static inline uint8_t clamp8(int x) { return (uint8_t)(x < 0 ? 0 : x > 255 ? 255 : x); }

/* value of the fancy-upsampled chroma plane at full-resolution (x, y); plane holds ds_w x ds_h real samples */
static int chroma_at(const uint8_t *pl, int stride, int ds_w, int ds_h, int hs, int vs, int x, int y) {
    if (hs == 1 && vs == 1) return pl[y * stride + x];
    const int cx = x >> 1;
    if (hs == 2 && vs == 1) {
        const uint8_t *row = pl + y * stride;
        if (ds_w <= 2) return row[cx];                       /* h2v1_upsample: replication */
        if (x & 1) return cx == ds_w - 1 ? row[cx] : (3 * row[cx] + row[cx + 1] + 2) >> 2;
        return cx == 0 ? row[0] : (3 * row[cx] + row[cx - 1] + 1) >> 2;
    }
    /* h2v2 */
    const int cy = y >> 1;
    if (ds_w <= 2) return pl[cy * stride + cx];              /* h2v2_upsample: replication */
    int ny = (y & 1) ? cy + 1 : cy - 1;                      /* nearer neighbour row: above for even y, below for odd y */
    if (ny < 0) ny = 0;
    if (ny > ds_h - 1) ny = ds_h - 1;
    const uint8_t *r0 = pl + cy * stride, *r1 = pl + ny * stride;
    const int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) {
        if (cx == ds_w - 1) return (cur * 4 + 7) >> 4;
        return (cur * 3 + (3 * r0[cx + 1] + r1[cx + 1]) + 7) >> 4;
    }
    if (cx == 0) return (cur * 4 + 8) >> 4;
    return (cur * 3 + (3 * r0[cx - 1] + r1[cx - 1]) + 8) >> 4;
}

#define SCALEBITS 16
#define ONE_HALF ((int32_t)1 << (SCALEBITS - 1))
#define FIXC(x) ((int32_t)((x) * (1L << SCALEBITS) + 0.5))

int jpeg_oracle_info(const uint8_t *data, size_t len, int *w, int *h, int *ncomp) {
    Jpeg j;
    memset(&j, 0, sizeof j);
    /* header only: run the parser but stop before the scan by giving it no room to allocate is not possible; decode fully */
    int rc = parse(&j, data, len);
    for (int c = 0; c < 3; ++c) free(j.coef[c]);
    if (rc == J_OK) { *w = j.width; *h = j.height; *ncomp = j.ncomp; }
    return rc;
}

/* Decodes to RGB HWC uint8.  rgb must hold width*height*3 bytes (cap). */
int jpeg_oracle_decode(const uint8_t *data, size_t len, uint8_t *rgb, size_t cap, int *w, int *h) {
    Jpeg j;
    memset(&j, 0, sizeof j);
    int rc = parse(&j, data, len);
    uint8_t *plane[3] = {0, 0, 0};
    if (rc == J_OK && (size_t)j.width * j.height * 3 > cap) rc = J_ECAP;
    if (rc == J_OK) {
        for (int c = 0; c < j.ncomp; ++c) {
            const int stride = j.blocks_w[c] * 8;
            plane[c] = (uint8_t *)malloc((size_t)stride * j.blocks_h[c] * 8);
            for (int by = 0; by < j.blocks_h[c]; ++by)
                for (int bx = 0; bx < j.blocks_w[c]; ++bx)
                    idct_block(j.coef[c] + ((size_t)by * j.blocks_w[c] + bx) * 64, j.quant[j.tq[c]],
                               plane[c] + (size_t)by * 8 * stride + bx * 8, stride);
        }
        *w = j.width; *h = j.height;
        for (int y = 0; y < j.height; ++y)
            for (int x = 0; x < j.width; ++x) {
                uint8_t *o = rgb + ((size_t)y * j.width + x) * 3;
                const int Y = plane[0][(size_t)y * j.blocks_w[0] * 8 + x];
                if (j.ncomp == 1) { o[0] = o[1] = o[2] = (uint8_t)Y; continue; }
                const int cb = chroma_at(plane[1], j.blocks_w[1] * 8, j.ds_w[1], j.ds_h[1], j.hmax, j.vmax, x, y) - 128;
                const int cr = chroma_at(plane[2], j.blocks_w[2] * 8, j.ds_w[2], j.ds_h[2], j.hmax, j.vmax, x, y) - 128;
                o[0] = clamp8(Y + (int)((FIXC(1.40200) * cr + ONE_HALF) >> SCALEBITS));
                o[1] = clamp8(Y + (int)((-FIXC(0.34414) * cb + ONE_HALF - FIXC(0.71414) * cr) >> SCALEBITS));
                o[2] = clamp8(Y + (int)((FIXC(1.77200) * cb + ONE_HALF) >> SCALEBITS));
            }
    }
    for (int c = 0; c < 3; ++c) { free(j.coef[c]); free(plane[c]); }
    return rc;
}
