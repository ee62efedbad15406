/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing in the product path may link, import or call this.
 *
 * CPU restatement (plain C) of the integer resampler that the reference's preprocessing closure
 * reaches through `image.resize(image_size, PILImage.Resampling.BICUBIC)`
 *   reference call site: packages/lumen-clip/src/lumen_clip/backends/onnxrt_backend.py:410-431
 *   (and the open_clip transform built at backends/torch_backend.py:201-204).
 *
 * The arithmetic itself lives in a third-party dependency that is NOT under /root/reference:
 * Pillow (pinned 12.0.0 in the reference's uv.lock:1698-1699; 12.2.0 is installed in this image),
 * `src/libImaging/Resample.c`.  Its published algorithm, restated here for 8-bit RGB + BICUBIC:
 *   - per axis: scale = in/out; filterscale = max(scale, 1); support = 2.0 * filterscale
 *     (Keys cubic, a = -0.5, support 2); ksize = ceil(support) * 2 + 1
 *   - per output index xx: center = (xx + 0.5) * scale; xmin = max((int)(center - support + 0.5), 0);
 *     xmax = min((int)(center + support + 0.5), in) - xmin;
 *     w_j = cubic((j + xmin - center + 0.5) / filterscale), normalised to sum 1 (all in double)
 *   - 8bpc path: coefficients quantised to 22 fractional bits, round half away from zero
 *   - horizontal pass first, each sample = clip8((2^21 + sum w_j * px_j) >> 22) stored as uint8,
 *     then the vertical pass with the same rule on the uint8 intermediate.
 * Parity of this restatement is PINNED against Pillow itself (tests/golden/*.npz, produced by
 * tests/golden/make_golden.py with PIL.Image.resize and the reference closure).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PRECISION_BITS (32 - 8 - 2)

static double cubic_a05(double x) {
    const double a = -0.5;
    if (x < 0.0) x = -x;
    if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
    if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
    return 0.0;
}

/* Returns ksize; fills bounds[2*out] = {xmin, count} and kk[out*ksize] (22-bit fixed point). */
int oracle_resample_coeffs(int in_size, int out_size, int32_t **bounds_out, int32_t **kk_out) {
    double scale = (double)in_size / out_size;
    double filterscale = scale < 1.0 ? 1.0 : scale;
    double support = 2.0 * filterscale;
    int ksize = (int)ceil(support) * 2 + 1;
    int32_t *bounds = (int32_t *)malloc(sizeof(int32_t) * 2 * out_size);
    int32_t *kk = (int32_t *)calloc((size_t)out_size * ksize, sizeof(int32_t));
    double *k = (double *)malloc(sizeof(double) * ksize);
    for (int xx = 0; xx < out_size; xx++) {
        double center = (xx + 0.5) * scale;
        double ww = 0.0, ss = 1.0 / filterscale;
        int xmin = (int)(center - support + 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = (int)(center + support + 0.5);
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        for (int x = 0; x < xmax; x++) {
            double w = cubic_a05((x + xmin - center + 0.5) * ss);
            k[x] = w;
            ww += w;
        }
        for (int x = 0; x < xmax; x++) {
            if (ww != 0.0) k[x] /= ww;
            double v = k[x];
            kk[(size_t)xx * ksize + x] =
                v < 0 ? (int32_t)(-0.5 + v * (1 << PRECISION_BITS)) : (int32_t)(0.5 + v * (1 << PRECISION_BITS));
        }
        bounds[xx * 2 + 0] = xmin;
        bounds[xx * 2 + 1] = xmax;
    }
    free(k);
    *bounds_out = bounds;
    *kk_out = kk;
    return ksize;
}

static inline uint8_t clip8(int32_t v) {
    v >>= PRECISION_BITS; /* arithmetic shift, as Pillow's lookup table index */
    return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

/* src: h_in x w_in x 3 uint8 (HWC, packed).  dst: h_out x w_out x 3 uint8. */
void oracle_resize_bicubic_rgb(const uint8_t *src, int h_in, int w_in, uint8_t *dst, int h_out, int w_out) {
    int32_t *bh, *kh, *bv, *kv;
    int ksh = oracle_resample_coeffs(w_in, w_out, &bh, &kh);
    int ksv = oracle_resample_coeffs(h_in, h_out, &bv, &kv);
    /* PIL.Image.Image.resize (Pillow >= 12, the reference's pin: packages/lumen-clip/pyproject.toml "pillow>=12.0.0"):
     *     if self.size[1] > self.size[0] * 100 and size[1] < self.size[1]:   # very tall and narrow, shrinking vertically
     *         im = self.im.resize((self.size[0], size[1]), ...); im = im.resize(size, ...)
     * i.e. two complete resizes, VERTICAL FIRST, each rounding to 8 bits - the results differ from horizontal-first by up to
     * ~12 grey levels (checked against Pillow 12.2.0 on 300 random sizes: this rule explains every difference). */
    if ((long long)h_in > (long long)w_in * 100 && h_out < h_in) {
        uint8_t *tv = (uint8_t *)malloc((size_t)h_out * w_in * 3);
        for (int yy = 0; yy < h_out; yy++) {
            int ymin = bv[yy * 2], cnt = bv[yy * 2 + 1];
            const int32_t *k = kv + (size_t)yy * ksv;
            for (int i = 0; i < w_in * 3; i++) {
                int32_t ss = 1 << (PRECISION_BITS - 1);
                for (int y = 0; y < cnt; y++) ss += (int32_t)src[(size_t)(y + ymin) * w_in * 3 + i] * k[y];
                tv[(size_t)yy * w_in * 3 + i] = clip8(ss);
            }
        }
        for (int y = 0; y < h_out; y++) {
            const uint8_t *row = tv + (size_t)y * w_in * 3;
            for (int xx = 0; xx < w_out; xx++) {
                int xmin = bh[xx * 2], cnt = bh[xx * 2 + 1];
                const int32_t *k = kh + (size_t)xx * ksh;
                for (int c = 0; c < 3; c++) {
                    int32_t ss = 1 << (PRECISION_BITS - 1);
                    for (int x = 0; x < cnt; x++) ss += (int32_t)row[(x + xmin) * 3 + c] * k[x];
                    dst[((size_t)y * w_out + xx) * 3 + c] = clip8(ss);
                }
            }
        }
        free(tv); free(bh); free(kh); free(bv); free(kv);
        return;
    }
    uint8_t *tmp = (uint8_t *)malloc((size_t)h_in * w_out * 3);
    for (int y = 0; y < h_in; y++) {
        const uint8_t *row = src + (size_t)y * w_in * 3;
        for (int xx = 0; xx < w_out; xx++) {
            int xmin = bh[xx * 2], cnt = bh[xx * 2 + 1];
            const int32_t *k = kh + (size_t)xx * ksh;
            for (int c = 0; c < 3; c++) {
                int32_t ss = 1 << (PRECISION_BITS - 1);
                for (int x = 0; x < cnt; x++) ss += (int32_t)row[(x + xmin) * 3 + c] * k[x];
                tmp[((size_t)y * w_out + xx) * 3 + c] = clip8(ss);
            }
        }
    }
    for (int yy = 0; yy < h_out; yy++) {
        int ymin = bv[yy * 2], cnt = bv[yy * 2 + 1];
        const int32_t *k = kv + (size_t)yy * ksv;
        for (int i = 0; i < w_out * 3; i++) {
            int32_t ss = 1 << (PRECISION_BITS - 1);
            for (int y = 0; y < cnt; y++) ss += (int32_t)tmp[(size_t)(y + ymin) * w_out * 3 + i] * k[y];
            dst[(size_t)yy * w_out * 3 + i] = clip8(ss);
        }
    }
    free(tmp); free(bh); free(kh); free(bv); free(kv);
}

/*
 * Full preprocessing of one image, restating onnxrt_backend.py:410-431 (mode 0: squash to SxS, no crop)
 * and the open_clip / torchvision transform of torch_backend.py:201-204 (mode 1: shortest side -> S with
 * int(S*long/short) on the long side, then a centre crop whose offset is round-half-even((dim-S)/2)).
 * out: float32 [3][S][S] = (px/255 - mean[c]) / std[c], computed in float32 in that order.
 * u8_out (optional): the S x S x 3 resized+cropped bytes before normalisation.
 */
void oracle_preprocess_rgb(const uint8_t *src, int h_in, int w_in, int S, int mode,
                           const float *mean, const float *std, float *out, uint8_t *u8_out) {
    int new_h = S, new_w = S, top = 0, left = 0;
    if (mode == 1) {
        if (w_in <= h_in) { new_w = S; new_h = (int)((long long)S * h_in / w_in); }
        else              { new_h = S; new_w = (int)((long long)S * w_in / h_in); }
        top  = (int)nearbyint((new_h - S) / 2.0);   /* default rounding mode: half to even, as Python round() */
        left = (int)nearbyint((new_w - S) / 2.0);
    }
    uint8_t *rs = (uint8_t *)malloc((size_t)new_h * new_w * 3);
    if (new_h == h_in && new_w == w_in) memcpy(rs, src, (size_t)new_h * new_w * 3);   /* Image.resize returns a copy */
    else oracle_resize_bicubic_rgb(src, h_in, w_in, rs, new_h, new_w);
    for (int y = 0; y < S; y++)
        for (int x = 0; x < S; x++)
            for (int c = 0; c < 3; c++) {
                uint8_t p = rs[((size_t)(y + top) * new_w + (x + left)) * 3 + c];
                if (u8_out) u8_out[((size_t)y * S + x) * 3 + c] = p;
                float v = (float)p / 255.0f;
                out[((size_t)c * S + y) * S + x] = (v - mean[c]) / std[c];
            }
    free(rs);
}
