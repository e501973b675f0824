/*
 * norm_oracle.c -- CPU restatement of friendlymatthew/norm's image-effect hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see norm_oracle.h).  Build: oracle/Makefile
 *   gcc -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp -shared -fPIC ... -lz -lm
 * -ffp-contract=off matters: every fused multiply-add below is an explicit fmaf(); every
 * plain `a*b + c` is two roundings, as in the Rust / non-fusing shader it restates.
 *
 * Parity status: decode + expansion PINNED by the reference's 13 golden blobs and Adam7
 * vectors; effect shaders and SSIM are PARITY UNPINNED (no live reference test, fixture
 * or runnable build exists for them) -- each function cites the source lines it follows.
 *
 * Two forms of every filter:
 *   R-form  literal loop / operation order of the WGSL, unit-scale (c/255) floats.
 *   S-form  the fixed-order definition the CUDA kernels must match BIT-EXACTLY: raw-scale
 *           (0..255) floats, separable blur with a host-computed normalised tap table,
 *           explicit fmaf chains.  S vs R differs by at most 1 LSB (tests/test_oracle.py).
 * Texel convention for the blur's half-texel taps through the NEAREST sampler: floor.
 */
#include "norm_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------------------ */
/* helpers                                                                              */
/* ------------------------------------------------------------------------------------ */

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

static inline uint32_t be32(const uint8_t *p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* float -> unorm8 store of a unit-scale value: clamp, scale, round-half-even
 * (storage textures are rgba8unorm, gpu_state.rs:608). */
static inline uint8_t unorm8(float v) {
    if (!(v > 0.0f)) return 0; /* also NaN -> 0 */
    if (v > 1.0f) v = 1.0f;
    return (uint8_t)rintf(v * 255.0f);
}

/* raw-scale (0..255) float -> u8, round-half-even, saturating. */
static inline uint8_t quant_raw(float v) {
    if (!(v > 0.0f)) return 0;
    if (v > 255.0f) v = 255.0f;
    return (uint8_t)rintf(v);
}

static int channels_of(int ct) {
    /* src/image/grammar.rs:19-29 */
    switch (ct) {
    case 0: return 1;
    case 2: return 3;
    case 3: return 1;
    case 4: return 2;
    case 6: return 4;
    default: return 0;
    }
}

size_t oracle_row_bytes(uint32_t w, int ct, int bd) {
    return ((size_t)w * (size_t)channels_of(ct) * (size_t)bd + 7) / 8;
}

/* ------------------------------------------------------------------------------------ */
/* PNG decode front-end                                                                 */
/* ------------------------------------------------------------------------------------ */

static const int A7_X0[7] = {0, 4, 0, 2, 0, 1, 0};
static const int A7_Y0[7] = {0, 0, 4, 0, 2, 0, 1};
static const int A7_DX[7] = {8, 8, 4, 4, 2, 2, 1};
static const int A7_DY[7] = {8, 8, 8, 4, 4, 2, 2};

/* interlace.rs:35-83 -- same pass geometry, but safe for w,h < 5 (the reference computes
 * `(width - 4).div_ceil(8)` on usize and underflows, interlace.rs:47). */
void oracle_adam7_pass_sizes(uint32_t w, uint32_t h, uint32_t pw[7], uint32_t ph[7]) {
    for (int p = 0; p < 7; ++p) {
        pw[p] = w > (uint32_t)A7_X0[p] ? (w - A7_X0[p] + A7_DX[p] - 1) / A7_DX[p] : 0;
        ph[p] = h > (uint32_t)A7_Y0[p] ? (h - A7_Y0[p] + A7_DY[p] - 1) / A7_DY[p] : 0;
    }
}

/* scanline_reader.rs:149-167 */
static inline uint8_t paeth(uint8_t left, uint8_t up, uint8_t up_left) {
    int a = left, b = up, c = up_left;
    int p = a + b - c;
    int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    if (pa <= pb && pa <= pc) return left;
    if (pb <= pc) return up;
    return up_left;
}

/* scanline_reader.rs:40-146: undo one row's filter.  `prev` is the previous unfiltered row
 * of the same (sub-)image or NULL.  bpp = max(1, channels*bit_depth/8) per the PNG spec
 * (the reference's num_bytes_per_pixel rounds up too, grammar.rs:32-36). */
static int unfilter_row(int ft, const uint8_t *row, const uint8_t *prev, uint8_t *dst, size_t n,
                        size_t bpp) {
    switch (ft) {
    case 0:
        memcpy(dst, row, n);
        return 0;
    case 1:
        for (size_t j = 0; j < n; ++j) dst[j] = (uint8_t)(row[j] + (j >= bpp ? dst[j - bpp] : 0));
        return 0;
    case 2:
        for (size_t j = 0; j < n; ++j) dst[j] = (uint8_t)(row[j] + (prev ? prev[j] : 0));
        return 0;
    case 3:
        for (size_t j = 0; j < n; ++j) {
            unsigned a = j >= bpp ? dst[j - bpp] : 0;
            unsigned b = prev ? prev[j] : 0;
            dst[j] = (uint8_t)(row[j] + ((a + b) >> 1));
        }
        return 0;
    case 4:
        for (size_t j = 0; j < n; ++j) {
            uint8_t a = j >= bpp ? dst[j - bpp] : 0;
            uint8_t b = prev ? prev[j] : 0;
            uint8_t c = (prev && j >= bpp) ? prev[j - bpp] : 0;
            dst[j] = (uint8_t)(row[j] + paeth(a, b, c));
        }
        return 0;
    default:
        return -1; /* grammar.rs:48-62: "Unrecognized filter method" */
    }
}

static int valid_depth(int ct, int bd) {
    switch (ct) {
    case 0: return bd == 1 || bd == 2 || bd == 4 || bd == 8 || bd == 16;
    case 3: return bd == 1 || bd == 2 || bd == 4 || bd == 8;
    case 2:
    case 4:
    case 6: return bd == 8 || bd == 16;
    default: return 0;
    }
}

void oracle_png_free(oracle_png *p) {
    if (p && p->pixels) {
        free(p->pixels);
        p->pixels = NULL;
    }
}

int oracle_png_decode(const uint8_t *data, size_t n, oracle_png *out) {
    static const uint8_t SIG[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    memset(out, 0, sizeof(*out));
    /* decoder.rs:30-33 */
    if (n < 8 || memcmp(data, SIG, 8) != 0) return ORACLE_E_CORRUPT;

    size_t cur = 8;
    int have_ihdr = 0, have_iend = 0;
    uint8_t *idat = NULL;
    size_t idat_len = 0, idat_cap = 0;
    int rc = ORACLE_E_CORRUPT;

    /* decoder.rs:119-222 */
    while (!have_iend) {
        if (cur + 12 > n) goto done;
        uint32_t length = be32(data + cur);
        const uint8_t *type = data + cur + 4;
        if ((size_t)length > n - cur - 12) goto done; /* decoder.rs:127-130 */
        const uint8_t *body = data + cur + 8;
        uint32_t expect = be32(body + length);
        /* decoder.rs:132-140, crc32.rs:20-26: CRC over type + data */
        uint32_t crc = (uint32_t)crc32(crc32(0L, Z_NULL, 0), type, 4 + length);
        if (crc != expect) goto done;

        if (!memcmp(type, "IHDR", 4)) {
            if (have_ihdr || length != 13) goto done; /* decoder.rs:144 */
            out->width = be32(body);
            out->height = be32(body + 4);
            out->bit_depth = body[8];
            out->color_type = body[9];
            out->compression = body[10];
            out->filter = body[11];
            out->interlace = body[12];
            if (!channels_of(out->color_type)) goto done; /* image/grammar.rs:31-46 */
            if (!valid_depth(out->color_type, out->bit_depth)) goto done;
            if (out->compression != 0 || out->filter != 0) goto done; /* decoder.rs:50-53,92-95 */
            if (out->interlace > 1 || out->width == 0 || out->height == 0) goto done;
            have_ihdr = 1;
        } else {
            if (!have_ihdr) goto done; /* decoder.rs:46-48 */
            if (!memcmp(type, "PLTE", 4)) {
                if (length % 3 != 0 || length > 768) goto done;             /* decoder.rs:158 */
                if (out->color_type == 0 || out->color_type == 4) goto done; /* decoder.rs:170-173 */
                memcpy(out->plte, body, length);
                out->n_plte = (int32_t)(length / 3);
            } else if (!memcmp(type, "tRNS", 4)) {
                /* not parsed by the reference (falls into `_foreign`, decoder.rs:204-209) */
                if (length > 256) goto done;
                memcpy(out->trns, body, length);
                out->n_trns = (int32_t)length;
            } else if (!memcmp(type, "gAMA", 4)) {
                if (length != 4) goto done;
                out->gamma = be32(body); /* decoder.rs:186 */
            } else if (!memcmp(type, "IDAT", 4)) {
                /* decoder.rs:55-76: concatenate every IDAT */
                if (idat_len + length > idat_cap) {
                    idat_cap = (idat_len + length) * 2 + 64;
                    uint8_t *t = (uint8_t *)realloc(idat, idat_cap);
                    if (!t) goto done;
                    idat = t;
                }
                memcpy(idat + idat_len, body, length);
                idat_len += length;
            } else if (!memcmp(type, "IEND", 4)) {
                have_iend = 1;
            }
        }
        cur += 12 + (size_t)length;
    }
    if (!have_ihdr || idat_len == 0) goto done; /* decoder.rs:97 */
    if (out->color_type == 3 && out->n_plte == 0) goto done;

    {
        const int ct = out->color_type, bd = out->bit_depth;
        const uint32_t w = out->width, h = out->height;
        const size_t rowb = oracle_row_bytes(w, ct, bd);
        size_t bpp = (size_t)channels_of(ct) * bd / 8;
        if (bpp == 0) bpp = 1;
        uint32_t pw[7], ph[7];
        size_t expect_len = 0;
        if (out->interlace) {
            oracle_adam7_pass_sizes(w, h, pw, ph);
            for (int p = 0; p < 7; ++p)
                if (pw[p] && ph[p]) expect_len += (size_t)ph[p] * (1 + oracle_row_bytes(pw[p], ct, bd));
        } else {
            expect_len = (size_t)h * (1 + rowb);
        }
        /* decoder.rs:84-86 (flate2 -> miniz_oxide; system zlib is equivalent) */
        uint8_t *raw = (uint8_t *)malloc(expect_len + 1);
        if (!raw) goto done;
        uLongf got = (uLongf)(expect_len + 1);
        int zr = uncompress(raw, &got, idat, (uLong)idat_len);
        if (zr != Z_OK || (size_t)got != expect_len) {
            free(raw);
            goto done;
        }
        out->len = rowb * h;
        out->pixels = (uint8_t *)calloc(out->len ? out->len : 1, 1);
        if (!out->pixels) {
            free(raw);
            goto done;
        }
        int bad = 0;
        if (!out->interlace) {
            /* scanline_reader.rs:40-146 */
            for (uint32_t y = 0; y < h && !bad; ++y) {
                const uint8_t *src = raw + (size_t)y * (1 + rowb);
                uint8_t *dst = out->pixels + (size_t)y * rowb;
                bad |= unfilter_row(src[0], src + 1, y ? dst - rowb : NULL, dst, rowb, bpp) != 0;
            }
        } else {
            /* scanline_reader.rs:170-230 restated per the PNG spec: each pass is a reduced
             * image, unfiltered against its own previous row, then scattered. */
            const int bits = channels_of(ct) * bd; /* bits per pixel */
            size_t off = 0;
            for (int p = 0; p < 7 && !bad; ++p) {
                if (!pw[p] || !ph[p]) continue;
                const size_t prow = oracle_row_bytes(pw[p], ct, bd);
                uint8_t *buf = (uint8_t *)malloc(2 * prow + 2);
                uint8_t *curr = buf, *prev = buf + prow;
                for (uint32_t i = 0; i < ph[p] && !bad; ++i) {
                    const uint8_t *src = raw + off + (size_t)i * (1 + prow);
                    bad |= unfilter_row(src[0], src + 1, i ? prev : NULL, curr, prow, bpp) != 0;
                    if (bad) break;
                    const uint32_t y = A7_Y0[p] + i * A7_DY[p];
                    uint8_t *drow = out->pixels + (size_t)y * rowb;
                    for (uint32_t j = 0; j < pw[p]; ++j) {
                        const uint32_t x = A7_X0[p] + j * A7_DX[p];
                        if (bits >= 8) {
                            memcpy(drow + (size_t)x * (bits / 8), curr + (size_t)j * (bits / 8),
                                   bits / 8);
                        } else {
                            const unsigned sb = (unsigned)(j * bits), db = (unsigned)(x * bits);
                            const unsigned mask = (1u << bits) - 1;
                            unsigned v = (curr[sb >> 3] >> (8 - bits - (sb & 7))) & mask;
                            drow[db >> 3] |= (uint8_t)(v << (8 - bits - (db & 7)));
                        }
                    }
                    uint8_t *t = curr;
                    curr = prev;
                    prev = t;
                }
                free(buf);
                off += (size_t)ph[p] * (1 + prow);
            }
        }
        free(raw);
        if (bad) {
            oracle_png_free(out);
            goto done;
        }
    }
    rc = ORACLE_OK;
done:
    free(idat);
    if (rc != ORACLE_OK) oracle_png_free(out);
    return rc;
}

/* src/png/scanline_writer.rs:5-61: per row the five filter candidates, smallest sum of output bytes
 * wins (first minimum in the order None, Sub, Up, Average, Paeth).  paeth_compat != 0 reproduces the
 * reference's predictor p = a + b + c (scanline_writer.rs:8); 0 is the PNG-spec p = a + b - c. */
static uint8_t paeth_enc(uint8_t oa, uint8_t ob, uint8_t oc, int compat) {
    const int a = oa, b = ob, c = oc;
    const int p = compat ? a + b + c : a + b - c;
    const int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    if (pa <= pb && pa <= pc) return oa;
    if (pb <= pc) return ob;
    return oc;
}

void oracle_filter_scanlines(const uint8_t *pixels, uint32_t rows, size_t row_bytes, size_t bpp, int paeth_compat,
                             uint8_t *stream, uint8_t *filter_types) {
    uint8_t *cand[5];
    for (int k = 0; k < 5; ++k) cand[k] = (uint8_t *)malloc(row_bytes ? row_bytes : 1);
    uint8_t *zero = (uint8_t *)calloc(row_bytes ? row_bytes : 1, 1);
    for (uint32_t y = 0; y < rows; ++y) {
        const uint8_t *cur = pixels + (size_t)y * row_bytes, *prev = y ? cur - row_bytes : zero;
        uint32_t sum[5] = {0, 0, 0, 0, 0};
        for (size_t i = 0; i < row_bytes; ++i) {
            const uint8_t a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            cand[0][i] = cur[i];
            cand[1][i] = (uint8_t)(cur[i] - a);
            cand[2][i] = (uint8_t)(cur[i] - b);
            cand[3][i] = (uint8_t)(cur[i] - (uint8_t)(((unsigned)a + b) / 2));
            cand[4][i] = (uint8_t)(cur[i] - paeth_enc(a, b, c, paeth_compat));
            for (int k = 0; k < 5; ++k) sum[k] += cand[k][i];
        }
        int best = 0;
        for (int k = 1; k < 5; ++k)
            if (sum[k] < sum[best]) best = k;
        uint8_t *out = stream + (size_t)y * (row_bytes + 1);
        out[0] = (uint8_t)best;
        memcpy(out + 1, cand[best], row_bytes);
        if (filter_types) filter_types[y] = (uint8_t)best;
    }
    for (int k = 0; k < 5; ++k) free(cand[k]);
    free(zero);
}

/* ------------------------------------------------------------------------------------ */
/* Pixel expansion                                                                      */
/* ------------------------------------------------------------------------------------ */

/* 16 -> 8 bit: (v + 128) / 257 -- the convention of the `image` crate (0.25.5, Cargo.lock)
 * that the reference's own tests cross-check against (decoder.rs:267-278). */
static inline uint8_t down16(unsigned v) { return (uint8_t)((v + 128u) / 257u); }

static inline unsigned sample_at(const uint8_t *row, size_t idx, int bd) {
    switch (bd) {
    case 8: return row[idx];
    case 16: return ((unsigned)row[2 * idx] << 8) | row[2 * idx + 1];
    default: {
        const size_t bit = idx * (size_t)bd;
        return (row[bit >> 3] >> (8 - bd - (int)(bit & 7))) & ((1u << bd) - 1u);
    }
    }
}

static inline uint8_t to8(unsigned v, int bd) {
    switch (bd) {
    case 1: return (uint8_t)(v * 255u);
    case 2: return (uint8_t)(v * 85u);
    case 4: return (uint8_t)(v * 17u);
    case 8: return (uint8_t)v;
    default: return down16(v);
    }
}

int oracle_expand_rgba8(const oracle_png *p, int mode, uint8_t *out) {
    const int ct = p->color_type, bd = p->bit_depth;
    const uint32_t w = p->width, h = p->height;
    const size_t rowb = oracle_row_bytes(w, ct, bd);
    if (!channels_of(ct) || !valid_depth(ct, bd)) return ORACLE_E_INVALID;
    if (p->len < rowb * h) return ORACLE_E_INVALID;

    if (mode == ORACLE_EXPAND_COMPAT) {
        /* src/png/grammar.rs:123-155 -- byte shuffles on 8-bit samples; opaque types get
         * alpha 0 (grammar.rs:130,139); Palette is `unimplemented!` (grammar.rs:153); the
         * reference has no correct behaviour for other depths / interlace. */
        if (bd != 8 || ct == 3) return ORACLE_E_UNSUPPORTED;
        const size_t np = (size_t)w * h;
        const uint8_t *s = p->pixels;
        for (size_t i = 0; i < np; ++i) {
            uint8_t *o = out + 4 * i;
            switch (ct) {
            case 6: memcpy(o, s + 4 * i, 4); break;
            case 2: o[0] = s[3 * i], o[1] = s[3 * i + 1], o[2] = s[3 * i + 2], o[3] = 0; break;
            case 0: o[0] = o[1] = o[2] = s[i], o[3] = 0; break;
            case 4: o[0] = o[1] = o[2] = s[2 * i], o[3] = s[2 * i + 1]; break;
            }
        }
        return ORACLE_OK;
    }
    if (mode != ORACLE_EXPAND_SPEC) return ORACLE_E_INVALID;

    /* tRNS colour key at native depth (PNG spec 11.3.2.1) */
    unsigned key[3] = {0, 0, 0};
    int has_key = 0;
    const unsigned smask = bd == 16 ? 0xFFFFu : ((1u << bd) - 1u);
    if (ct == 0 && p->n_trns >= 2) {
        key[0] = (((unsigned)p->trns[0] << 8) | p->trns[1]) & smask;
        has_key = 1;
    } else if (ct == 2 && p->n_trns >= 6) {
        for (int c = 0; c < 3; ++c)
            key[c] = (((unsigned)p->trns[2 * c] << 8) | p->trns[2 * c + 1]) & smask;
        has_key = 1;
    }

    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t *row = p->pixels + (size_t)y * rowb;
        uint8_t *o = out + (size_t)y * w * 4;
        for (uint32_t x = 0; x < w; ++x, o += 4) {
            switch (ct) {
            case 0: {
                unsigned v = sample_at(row, x, bd);
                o[0] = o[1] = o[2] = to8(v, bd);
                o[3] = (has_key && v == key[0]) ? 0 : 255;
                break;
            }
            case 2: {
                unsigned r = sample_at(row, 3 * (size_t)x, bd), g = sample_at(row, 3 * (size_t)x + 1, bd),
                         b = sample_at(row, 3 * (size_t)x + 2, bd);
                o[0] = to8(r, bd), o[1] = to8(g, bd), o[2] = to8(b, bd);
                o[3] = (has_key && r == key[0] && g == key[1] && b == key[2]) ? 0 : 255;
                break;
            }
            case 3: {
                unsigned i = sample_at(row, x, bd);
                if ((int)i < p->n_plte) {
                    o[0] = p->plte[3 * i], o[1] = p->plte[3 * i + 1], o[2] = p->plte[3 * i + 2];
                } else {
                    o[0] = o[1] = o[2] = 0;
                }
                o[3] = (int)i < p->n_trns ? p->trns[i] : 255;
                break;
            }
            case 4: {
                unsigned v = sample_at(row, 2 * (size_t)x, bd), a = sample_at(row, 2 * (size_t)x + 1, bd);
                o[0] = o[1] = o[2] = to8(v, bd);
                o[3] = to8(a, bd);
                break;
            }
            case 6: {
                for (int c = 0; c < 4; ++c) o[c] = to8(sample_at(row, 4 * (size_t)x + c, bd), bd);
                break;
            }
            }
        }
    }
    return ORACLE_OK;
}

/* src/png/grammar.rs:89-121 */
int oracle_rgb8(const oracle_png *p, uint8_t *out) {
    if (p->bit_depth != 8 || p->color_type == 3) return ORACLE_E_UNSUPPORTED;
    const size_t np = (size_t)p->width * p->height;
    const uint8_t *s = p->pixels;
    for (size_t i = 0; i < np; ++i) {
        uint8_t *o = out + 3 * i;
        switch (p->color_type) {
        case 2: memcpy(o, s + 3 * i, 3); break;
        case 6: memcpy(o, s + 4 * i, 3); break;
        case 4: o[0] = o[1] = o[2] = s[2 * i]; break;
        case 0: o[0] = o[1] = o[2] = s[i]; break;
        default: return ORACLE_E_UNSUPPORTED;
        }
    }
    return ORACLE_OK;
}

/* ------------------------------------------------------------------------------------ */
/* Front stage: sRGB decode at load + gamma pass + grayscale + invert                   */
/* ------------------------------------------------------------------------------------ */

/* IEC 61966-2-1 EOTF -- what sampling an Rgba8UnormSrgb texture does (texture.rs:36). */
static double srgb_eotf(double c) {
    return c <= 0.04045 ? c / 12.92 : pow((c + 0.055) / 1.055, 2.4);
}

/* gamma_correct_compute.wgsl:25-40: load (format decodes sRGB), optional pow, unorm8 store.
 * Every output depends on one 8-bit input only, so the pass is a 256-entry table. */
static inline float front_channel(int cs, uint32_t gamma, unsigned c) {
    const double u = (double)c / 255.0;
    float v = (float)(cs == ORACLE_CS_REFERENCE ? srgb_eotf(u) : u);
    if (gamma != 0u) {
        const float gamma_value = (float)gamma / 100000.0f; /* :29 */
        const float inv_gamma = 1.0f / gamma_value;         /* :30 */
        v = powf(v, inv_gamma);                             /* :32-37 */
    }
    return v;
}

void oracle_front_lut(int cs, uint32_t gamma, uint8_t lut[256]) {
    for (unsigned c = 0; c < 256; ++c) lut[c] = unorm8(front_channel(cs, gamma, c));
}

static const float LUMA_R = 0.29891f, LUMA_G = 0.58661f, LUMA_B = 0.11448f;

void oracle_front_s(const uint8_t *in, uint32_t w, uint32_t h, const uint8_t lut[256], int grayscale,
                    int invert, uint8_t *out) {
    const size_t np = (size_t)w * h;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < np; ++i) {
        const uint8_t *s = in + 4 * i;
        uint8_t r = lut[s[0]], g = lut[s[1]], b = lut[s[2]], a = s[3]; /* alpha is linear */
        if (grayscale) {
            /* grayscale_compute.wgsl:25, raw scale, fixed fused order */
            float y = fmaf((float)b, LUMA_B, fmaf((float)g, LUMA_G, (float)r * LUMA_R));
            r = g = b = quant_raw(y);
        }
        if (invert) { /* invert_compute.wgsl:25-30: 1 - c is exact in raw scale */
            r = (uint8_t)(255 - r), g = (uint8_t)(255 - g), b = (uint8_t)(255 - b);
        }
        uint8_t *o = out + 4 * i;
        o[0] = r, o[1] = g, o[2] = b, o[3] = a;
    }
}

void oracle_front_r(const uint8_t *in, uint32_t w, uint32_t h, int cs, uint32_t gamma, int grayscale,
                    int invert, uint8_t *out) {
    const size_t np = (size_t)w * h;
    for (size_t i = 0; i < np; ++i) {
        const uint8_t *s = in + 4 * i;
        uint8_t q[4];
        for (int c = 0; c < 3; ++c) q[c] = unorm8(front_channel(cs, gamma, s[c]));
        q[3] = unorm8((float)s[3] / 255.0f);
        if (grayscale) {
            float r = (float)q[0] / 255.0f, g = (float)q[1] / 255.0f, b = (float)q[2] / 255.0f;
            float y = r * 0.29891f + g * 0.58661f + b * 0.11448f; /* grayscale_compute.wgsl:25 */
            q[0] = q[1] = q[2] = unorm8(y);
        }
        if (invert) {
            for (int c = 0; c < 3; ++c) q[c] = unorm8(1.0f - (float)q[c] / 255.0f);
        }
        memcpy(out + 4 * i, q, 4);
    }
}

/* ------------------------------------------------------------------------------------ */
/* Gaussian blur                                                                        */
/* ------------------------------------------------------------------------------------ */

#define ORACLE_MAX_RADIUS 255

/* image_shader.wgsl:83-103.  Loop variable x runs -r/2, -r/2+1, ... < r/2  (r taps);
 * tap k has offset off = k - r/2 and, through the NEAREST sampler under the floor
 * convention, texel delta floor(0.5 + off).  The 2-D weight factorises
 * (exp(-(x^2+y^2)/2s^2) = exp(-x^2/2s^2) exp(-y^2/2s^2)) and its constant 1/(2 PI s^2)
 * cancels in color/acc, so the S-form uses normalised 1-D taps computed in double. */
int oracle_blur_taps(uint32_t radius, float *taps, int *dmin) {
    if (radius < 1 || radius > ORACLE_MAX_RADIUS) return ORACLE_E_INVALID;
    const double r = (double)radius, sigma = r * 0.25;
    double w[ORACLE_MAX_RADIUS], sum = 0.0;
    for (uint32_t k = 0; k < radius; ++k) {
        const double off = -r / 2.0 + (double)k;
        w[k] = exp(-(off * off) / (2.0 * sigma * sigma));
        sum += w[k];
    }
    for (uint32_t k = 0; k < radius; ++k) taps[k] = (float)(w[k] / sum);
    *dmin = (int)floor(0.5 - r / 2.0);
    return ORACLE_OK;
}

/* S-form: horizontal pass then vertical pass, both  acc = fmaf(tap[k], v[k], acc)  for
 * k ascending from acc = 0, on raw-scale floats; the intermediate is kept in f32 (never
 * quantised); ClampToEdge addressing (gpu_state.rs:614-616); all four channels. */
int oracle_blur_s(const uint8_t *in, uint32_t w, uint32_t h, uint32_t radius, uint8_t *out) {
    float taps[ORACLE_MAX_RADIUS];
    int dmin;
    if (oracle_blur_taps(radius, taps, &dmin) != ORACLE_OK) return ORACLE_E_INVALID;
    const int r = (int)radius;
    const size_t rowf = (size_t)w * 4;
    int nthr = oracle_num_threads();
    int band = (int)((h + nthr - 1) / nthr);
    if (band < 16) band = 16;
    if (band > 128) band = 128;
    const int nbands = (int)((h + band - 1) / band);

#pragma omp parallel
    {
        float *pad = (float *)malloc(((size_t)w + r) * 4 * sizeof(float));
        float *H = (float *)malloc((size_t)(band + r) * rowf * sizeof(float));
        float *acc = (float *)malloc(rowf * sizeof(float));
#pragma omp for schedule(dynamic, 1)
        for (int bi = 0; bi < nbands; ++bi) {
            const int y0 = bi * band, y1 = (y0 + band < (int)h) ? y0 + band : (int)h;
            /* source rows needed (unclamped): y0+dmin .. y1-1+dmin+r-1 */
            const int s0 = y0 + dmin, ns = (y1 - y0) + r - 1;
            int last_src = -1;
            for (int j = 0; j < ns; ++j) {
                const int sy = clampi(s0 + j, 0, (int)h - 1);
                float *Hrow = H + (size_t)j * rowf;
                if (sy == last_src) { /* replicated edge row */
                    memcpy(Hrow, Hrow - rowf, rowf * sizeof(float));
                    continue;
                }
                last_src = sy;
                const uint8_t *src = in + (size_t)sy * rowf;
                for (int i = 0; i < (int)w + r - 1; ++i) {
                    const uint8_t *px = src + 4 * (size_t)clampi(i + dmin, 0, (int)w - 1);
                    pad[4 * i] = px[0], pad[4 * i + 1] = px[1], pad[4 * i + 2] = px[2],
                            pad[4 * i + 3] = px[3];
                }
                for (size_t i = 0; i < rowf; ++i) Hrow[i] = 0.0f;
                for (int k = 0; k < r; ++k) {
                    const float tk = taps[k];
                    const float *p = pad + 4 * (size_t)k;
                    for (size_t i = 0; i < rowf; ++i) Hrow[i] = fmaf(tk, p[i], Hrow[i]);
                }
            }
            for (int y = y0; y < y1; ++y) {
                for (size_t i = 0; i < rowf; ++i) acc[i] = 0.0f;
                for (int k = 0; k < r; ++k) {
                    const float tk = taps[k];
                    const float *p = H + (size_t)(y - y0 + k) * rowf;
                    for (size_t i = 0; i < rowf; ++i) acc[i] = fmaf(tk, p[i], acc[i]);
                }
                uint8_t *o = out + (size_t)y * rowf;
                for (size_t i = 0; i < rowf; ++i) o[i] = quant_raw(acc[i]);
            }
        }
        free(pad);
        free(H);
        free(acc);
    }
    return ORACLE_OK;
}

/* R-form: literal image_shader.wgsl:88-103 -- x outer, y inner, f32 throughout, the
 * un-normalised 2-D weight of :83-86 with PI = 22/7 (:81), color/acc at the end.  The weight
 * depends only on (x,y) so it is tabulated (same f32 values, same order). */
int oracle_blur_r(const uint8_t *in, uint32_t w, uint32_t h, uint32_t radius, uint8_t *out) {
    if (radius < 1 || radius > ORACLE_MAX_RADIUS) return ORACLE_E_INVALID;
    const float PI = 22.0f / 7.0f;
    const float rad = (float)radius;
    const float sigma = rad * 0.25f;
    const int r = (int)radius;
    float *wt = (float *)malloc((size_t)r * r * sizeof(float));
    int *dl = (int *)malloc((size_t)r * sizeof(int));
    int n = 0;
    for (float x = -1.0f * rad / 2.0f; x < rad / 2.0f; x = x + 1.0f) {
        if (n >= r) break;
        dl[n] = (int)floorf(0.5f + x);
        int m = 0;
        for (float y = -1.0f * rad / 2.0f; y < rad / 2.0f; y = y + 1.0f) {
            if (m >= r) break;
            wt[n * r + m] = 1.0f / (2.0f * PI * sigma * sigma) *
                            expf(-((x * x + y * y) / (2.0f * sigma * sigma)));
            ++m;
        }
        ++n;
    }
#pragma omp parallel for schedule(static)
    for (int py = 0; py < (int)h; ++py) {
        for (int px = 0; px < (int)w; ++px) {
            float acc = 0.0f, col[4] = {0, 0, 0, 0};
            for (int ix = 0; ix < r; ++ix) {
                const int sx = clampi(px + dl[ix], 0, (int)w - 1);
                for (int iy = 0; iy < r; ++iy) {
                    const int sy = clampi(py + dl[iy], 0, (int)h - 1);
                    const float weight = wt[ix * r + iy];
                    const uint8_t *s = in + ((size_t)sy * w + sx) * 4;
                    for (int c = 0; c < 4; ++c) col[c] = col[c] + ((float)s[c] / 255.0f) * weight;
                    acc = acc + weight;
                }
            }
            uint8_t *o = out + ((size_t)py * w + px) * 4;
            for (int c = 0; c < 4; ++c) o[c] = unorm8(col[c] / acc);
        }
    }
    free(wt);
    free(dl);
    return ORACLE_OK;
}

/* ------------------------------------------------------------------------------------ */
/* Sobel line detection                                                                 */
/* ------------------------------------------------------------------------------------ */

/* image_shader.wgsl:132-156.  Texture v grows downward, so "top" (+y) is row py+1.
 *   gx = tl + 2 l + bl - tr - 2 r - br
 *   gy = -l - 2 t - tr + bl + 2 b + br        (note -l where a textbook Sobel has -tl, :151)
 * S-form works on raw-scale intensities: I = sqrt(r^2+g^2+b^2) with an exact integer
 * radicand; the left-to-right sums are order-exact because 2*v is exact; the magnitude is
 * sqrtf(fmaf(gx, gx, gy*gy)); output (c,c,c,255). */
void oracle_sobel_s(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out) {
    float *I = (float *)malloc((size_t)w * h * sizeof(float));
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)h; ++y)
        for (int x = 0; x < (int)w; ++x) {
            const uint8_t *s = in + ((size_t)y * w + x) * 4;
            const float r = s[0], g = s[1], b = s[2];
            I[(size_t)y * w + x] = sqrtf(fmaf(b, b, fmaf(g, g, r * r)));
        }
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)h; ++y) {
        const float *up = I + (size_t)clampi(y + 1, 0, (int)h - 1) * w; /* "top" = +v */
        const float *mid = I + (size_t)y * w;
        const float *dn = I + (size_t)clampi(y - 1, 0, (int)h - 1) * w; /* "bottom" */
        for (int x = 0; x < (int)w; ++x) {
            const int xl = clampi(x - 1, 0, (int)w - 1), xr = clampi(x + 1, 0, (int)w - 1);
            const float tl = up[xl], t = up[x], tr = up[xr];
            const float l = mid[xl], r = mid[xr];
            const float bl = dn[xl], b = dn[x], br = dn[xr];
            const float gx = tl + 2.0f * l + bl - tr - 2.0f * r - br;
            const float gy = -l - 2.0f * t - tr + bl + 2.0f * b + br;
            const uint8_t c = quant_raw(sqrtf(fmaf(gx, gx, gy * gy)));
            uint8_t *o = out + ((size_t)y * w + x) * 4;
            o[0] = o[1] = o[2] = c, o[3] = 255;
        }
    }
    free(I);
}

static inline float intensity_r(const uint8_t *s) {
    /* image_shader.wgsl:132-135 on unit-scale samples, no fusion */
    const float r = (float)s[0] / 255.0f, g = (float)s[1] / 255.0f, b = (float)s[2] / 255.0f;
    return sqrtf(r * r + g * g + b * b);
}

void oracle_sobel_r(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out) {
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)h; ++y) {
        const int yu = clampi(y + 1, 0, (int)h - 1), yd = clampi(y - 1, 0, (int)h - 1);
        for (int x = 0; x < (int)w; ++x) {
            const int xl = clampi(x - 1, 0, (int)w - 1), xr = clampi(x + 1, 0, (int)w - 1);
#define PX(xx, yy) (in + ((size_t)(yy) * w + (xx)) * 4)
            const float top_left = intensity_r(PX(xl, yu)), left = intensity_r(PX(xl, y)),
                        bottom_left = intensity_r(PX(xl, yd)), top = intensity_r(PX(x, yu)),
                        bottom = intensity_r(PX(x, yd)), top_right = intensity_r(PX(xr, yu)),
                        right = intensity_r(PX(xr, y)), bottom_right = intensity_r(PX(xr, yd));
#undef PX
            const float gx =
                    top_left + 2.0f * left + bottom_left - top_right - 2.0f * right - bottom_right;
            const float gy =
                    -left - 2.0f * top - top_right + bottom_left + 2.0f * bottom + bottom_right;
            const float color = sqrtf((gx * gx) + (gy * gy));
            uint8_t *o = out + ((size_t)y * w + x) * 4;
            o[0] = o[1] = o[2] = unorm8(color), o[3] = unorm8(1.0f);
        }
    }
}

/* ------------------------------------------------------------------------------------ */
/* Sharpen                                                                              */
/* ------------------------------------------------------------------------------------ */

/* image_shader.wgsl:120-129: (1 + 4f) c - f (up + left + right + down), all 4 channels.
 * In raw scale every term is an exact integer, so the S-form is integer arithmetic. */
void oracle_sharpen_s(const uint8_t *in, uint32_t w, uint32_t h, uint32_t factor, uint8_t *out) {
    const int f = (int)factor;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)h; ++y) {
        const int yu = clampi(y + 1, 0, (int)h - 1), yd = clampi(y - 1, 0, (int)h - 1);
        for (int x = 0; x < (int)w; ++x) {
            const int xl = clampi(x - 1, 0, (int)w - 1), xr = clampi(x + 1, 0, (int)w - 1);
            for (int c = 0; c < 4; ++c) {
#define CH(xx, yy) ((int)in[((size_t)(yy) * w + (xx)) * 4 + c])
                const int v = (1 + 4 * f) * CH(x, y) - f * (CH(x, yu) + CH(xl, y) + CH(xr, y) + CH(x, yd));
#undef CH
                out[((size_t)y * w + x) * 4 + c] = (uint8_t)clampi(v, 0, 255);
            }
        }
    }
}

void oracle_sharpen_r(const uint8_t *in, uint32_t w, uint32_t h, uint32_t factor, uint8_t *out) {
    const float f = (float)factor;
    for (int y = 0; y < (int)h; ++y) {
        const int yu = clampi(y + 1, 0, (int)h - 1), yd = clampi(y - 1, 0, (int)h - 1);
        for (int x = 0; x < (int)w; ++x) {
            const int xl = clampi(x - 1, 0, (int)w - 1), xr = clampi(x + 1, 0, (int)w - 1);
            for (int c = 0; c < 4; ++c) {
#define CH(xx, yy) ((float)in[((size_t)(yy) * w + (xx)) * 4 + c] / 255.0f)
                const float v = (1.0f + 4.0f * f) * CH(x, y) -
                                f * (CH(x, yu) + CH(xl, y) + CH(xr, y) + CH(x, yd));
#undef CH
                out[((size_t)y * w + x) * 4 + c] = unorm8(v);
            }
        }
    }
}

/* ------------------------------------------------------------------------------------ */
/* Box blur (dead code in the reference, image_shader.wgsl:105-118)                     */
/* ------------------------------------------------------------------------------------ */

/* 3x3 mean of all four channels, ClampToEdge.  S-form: the nine raw bytes sum exactly in
 * integers, one f32 division by 9, round half even. */
void oracle_box_blur_s(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out) {
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)h; ++y)
        for (int x = 0; x < (int)w; ++x)
            for (int c = 0; c < 4; ++c) {
                int sum = 0;
                for (int dx = -1; dx <= 1; ++dx)
                    for (int dy = -1; dy <= 1; ++dy)
                        sum += in[((size_t)clampi(y + dy, 0, (int)h - 1) * w + clampi(x + dx, 0, (int)w - 1)) * 4 + c];
                out[((size_t)y * w + x) * 4 + c] = quant_raw((float)sum / 9.0f);
            }
}

/* Literal order: dx outer, dy inner, unit-scale f32 accumulation, divide by the f32 count. */
void oracle_box_blur_r(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out) {
    for (int y = 0; y < (int)h; ++y)
        for (int x = 0; x < (int)w; ++x)
            for (int c = 0; c < 4; ++c) {
                float acc = 0.0f, ct = 0.0f;
                for (int dx = -1; dx <= 1; ++dx)
                    for (int dy = -1; dy <= 1; ++dy) {
                        acc += (float)in[((size_t)clampi(y + dy, 0, (int)h - 1) * w + clampi(x + dx, 0, (int)w - 1)) * 4 + c] / 255.0f;
                        ct += 1.0f;
                    }
                out[((size_t)y * w + x) * 4 + c] = unorm8(acc / ct);
            }
}

/* ------------------------------------------------------------------------------------ */
/* Shape overlay + composite + crosshair                                                */
/* ------------------------------------------------------------------------------------ */

/* sRGB tables for the shape render texture, which has the surface's sRGB format
 * (gpu_state.rs:404-418): eotf[k] = decoded value of code k; thresh[j] (j = 1..255) = the
 * linear value at which the encoder steps from code j-1 to code j, i.e. EOTF((j-0.5)/255).
 * thresh[0] = 0.  Both evaluated in double, rounded once to f32. */
void oracle_srgb_tables(float eotf[256], float thresh[256]) {
    for (int k = 0; k < 256; ++k) {
        eotf[k] = (float)srgb_eotf((double)k / 255.0);
        thresh[k] = k == 0 ? 0.0f : (float)srgb_eotf(((double)k - 0.5) / 255.0);
    }
}

static inline unsigned srgb_code(const float thresh[256], float v) {
    unsigned lo = 0, hi = 255;
    while (lo < hi) {
        const unsigned mid = (lo + hi + 1) >> 1;
        if (v >= thresh[mid]) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

/* Surface encode of a finished frame (gpu_state.rs:94-99: the colour target is the first sRGB surface format, so
 * the value fs_main returns is sRGB-encoded on store; texture.rs:36 is the matching decode at load).  A stage
 * output is a stored linear unorm8 byte c; its code is the one whose decision interval contains c/255. */
void oracle_srgb_encode_lut(uint8_t lut[256]) {
    float eotf[256], thresh[256];
    oracle_srgb_tables(eotf, thresh);
    for (int c = 0; c < 256; ++c) lut[c] = (uint8_t)srgb_code(thresh, (float)c / 255.0f);
}

void oracle_encode_srgb(const uint8_t *in, size_t npix, uint8_t *out) {
    uint8_t lut[256];
    oracle_srgb_encode_lut(lut);
    for (size_t i = 0; i < npix; ++i) {
        out[4 * i + 0] = lut[in[4 * i + 0]], out[4 * i + 1] = lut[in[4 * i + 1]], out[4 * i + 2] = lut[in[4 * i + 2]];
        out[4 * i + 3] = in[4 * i + 3]; /* alpha is linear */
    }
}

static inline float smoothstep_f(float e0, float e1, float x) {
    float t = (x - e0) / (e1 - e0);
    t = fminf(fmaxf(t, 0.0f), 1.0f);
    return (t * t) * (3.0f - 2.0f * t);
}

/* WGSL mix(a, b, t) = a (1 - t) + b t, the second product fused into the sum */
static inline float mix_f(float a, float b, float t) { return fmaf(b, t, a * (1.0f - t)); }

/* shape_shader.wgsl:41-89 (fragment of the shape pass) -> ALPHA_BLENDING over a transparent
 * clear (gpu_state.rs:321, app_state.rs:589) -> store in the sRGB render texture -> sampled at
 * the texel centre and composited in image_shader.wgsl:176-193, then the crosshair disc.
 * Pixel (x, y) has pixel_coords (x + 0.5, y + 0.5).  A pixel that no circle, outline or
 * crosshair touches keeps its bytes.  Camera transform (vertex stage) is not modelled. */
int oracle_overlay(const uint8_t *in, uint32_t w, uint32_t h, const oracle_shape_uniform *su,
                   const oracle_circle *circles, const oracle_draw_uniform *du, uint8_t *out) {
    if (!in || !out || !su || (su->num_circles && !circles)) return ORACLE_E_INVALID;
    if (su->width != w || su->height != h || su->num_circles > 256) return ORACLE_E_INVALID;
    float eotf[256], thresh[256];
    oracle_srgb_tables(eotf, thresh);
    const float W = (float)w, H = (float)h, M = (float)(w < h ? w : h);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)h; ++y)
        for (int x = 0; x < (int)w; ++x) {
            const float px = (float)x + 0.5f, py = (float)y + 0.5f;
            float col[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            for (uint32_t i = 0; i < su->num_circles; ++i) {
                const float cx = circles[i].x * W, cy = circles[i].y * H, rp = circles[i].radius * M;
                const float dx = px - cx, dy = py - cy;
                const float d = sqrtf(fmaf(dy, dy, dx * dx));
                if (d <= rp) {
                    const float a = 1.0f - smoothstep_f(rp - 1.0f, rp, d);
                    col[0] = mix_f(col[0], 0.0f, a), col[1] = mix_f(col[1], 1.0f, a);
                    col[2] = mix_f(col[2], 1.0f, a), col[3] = mix_f(col[3], a, a);
                }
                if (i == su->selected_circle) {
                    const float od = fabsf(d - rp);
                    if (od <= 3.0f) {
                        const float a = 1.0f - smoothstep_f(0.0f, 3.0f, od);
                        col[0] = mix_f(col[0], 1.0f, a), col[1] = mix_f(col[1], 1.0f, a);
                        col[2] = mix_f(col[2], 0.0f, a), col[3] = mix_f(col[3], a, a);
                    }
                }
            }
            const uint8_t *s = in + ((size_t)y * w + x) * 4;
            uint8_t *o = out + ((size_t)y * w + x) * 4;
            /* render-target store: blend (rgb * a, a), sRGB-encode rgb, unorm8 alpha */
            const unsigned ca = unorm8(col[3]);
            int hit = 0;
            if (du && du->crosshair == 1u) {
                const float dx = px - du->circle_center_x, dy = py - du->circle_center_y;
                hit = sqrtf(fmaf(dy, dy, dx * dx)) < du->circle_radius;
            }
            if (ca == 0 && !hit) {
                memcpy(o, s, 4);
                continue;
            }
            float p[4];
            for (int c = 0; c < 4; ++c) p[c] = (float)s[c] / 255.0f;
            if (ca > 0) {
                const float sa = (float)ca / 255.0f;
                for (int c = 0; c < 3; ++c) p[c] = mix_f(p[c], eotf[srgb_code(thresh, col[c] * col[3])], sa);
                p[3] = mix_f(p[3], sa, sa);
            }
            if (hit) p[0] = 0.0f, p[1] = 1.0f, p[2] = 1.0f;
            for (int c = 0; c < 4; ++c) o[c] = unorm8(p[c]);
        }
    return ORACLE_OK;
}

/* ------------------------------------------------------------------------------------ */
/* SSIM                                                                                 */
/* ------------------------------------------------------------------------------------ */

/* ssim.rs:73-75 (RGBA branch): f32, two roundings per term, left to right, raw 0..255. */
static inline float luma_ssim(const uint8_t *p) {
    const float r = (float)p[0], g = (float)p[1], b = (float)p[2];
    return r * 0.29891f + g * 0.58661f + b * 0.11448f;
}

void oracle_ssim_moments(const uint8_t *a, const uint8_t *b, size_t npix, double m[5]) {
    double sx = 0, sy = 0, sxx = 0, syy = 0, sxy = 0;
#pragma omp parallel for schedule(static) reduction(+ : sx, sy, sxx, syy, sxy)
    for (size_t i = 0; i < npix; ++i) {
        const double x = (double)luma_ssim(a + 4 * i), y = (double)luma_ssim(b + 4 * i);
        sx += x, sy += y, sxx += x * x, syy += y * y, sxy += x * y;
    }
    m[0] = sx, m[1] = sy, m[2] = sxx, m[3] = syy, m[4] = sxy;
}

/* ssim.rs:9-14,127-183 evaluated in f64 from raw moments:
 *   mean = S/N ; var = (Sxx - Sx^2/N)/(N-1) (Bessel, :173-183) ; cov likewise (:161-170). */
double oracle_ssim_from_moments(const double m[5], double n) {
    const double K1 = 0.01, K2 = 0.03;
    const double C1 = (K1 * 255.0) * (K1 * 255.0), C2 = (K2 * 255.0) * (K2 * 255.0), C3 = C2 / 2.0;
    const double mx = m[0] / n, my = m[1] / n;
    double vx = (m[2] - m[0] * m[0] / n) / (n - 1.0);
    double vy = (m[3] - m[1] * m[1] / n) / (n - 1.0);
    const double cov = (m[4] - m[0] * m[1] / n) / (n - 1.0);
    if (vx < 0.0) vx = 0.0;
    if (vy < 0.0) vy = 0.0;
    const double sx = sqrt(vx), sy = sqrt(vy);
    const double lum = (2.0 * mx * my + C1) / (mx * mx + my * my + C1);    /* :140-145 */
    const double con = (2.0 * sx * sy + C2) / (sx * sx + sy * sy + C2);    /* :148-153 */
    const double str = (cov + C3) / (sx * sy + C3);                        /* :156-158 */
    return lum * con * str;                                                /* :136 */
}

/* Png::luma_buffer per colour type (ssim.rs:18-87), on the RGBA8 expansion of the image:
 *   mode 0  RGB / RGBA        r*0.29891 + g*0.58661 + b*0.11448 on raw bytes   (:56-58, :73-75)
 *   mode 1  Grayscale         y as f32                                            (:25)
 *   mode 2  GrayscaleAlpha    y as f32 / 255.0  -- the reference's own inconsistency (:41), kept */
static inline float luma_mode(const uint8_t *p, int mode) {
    if (mode == 1) return (float)p[0];
    if (mode == 2) return (float)p[0] / 255.0f;
    return luma_ssim(p);
}

double oracle_ssim_luma(const uint8_t *a, const uint8_t *b, size_t npix, int mode_a, int mode_b) {
    double sx = 0, sy = 0, sxx = 0, syy = 0, sxy = 0;
#pragma omp parallel for schedule(static) reduction(+ : sx, sy, sxx, syy, sxy)
    for (size_t i = 0; i < npix; ++i) {
        const double x = (double)luma_mode(a + 4 * i, mode_a), y = (double)luma_mode(b + 4 * i, mode_b);
        sx += x, sy += y, sxx += x * x, syy += y * y, sxy += x * y;
    }
    const double m[5] = {sx, sy, sxx, syy, sxy};
    return oracle_ssim_from_moments(m, (double)npix);
}

double oracle_ssim(const uint8_t *a, const uint8_t *b, size_t npix) {
    double m[5];
    oracle_ssim_moments(a, b, npix, m);
    return oracle_ssim_from_moments(m, (double)npix);
}

/* The reference's literal arithmetic: f32 everywhere, strictly sequential sums, the pass
 * structure of ssim.rs (luma + running sum -> mean; sum (l-mean)^2; covariance).
 * `a` plays `self` (the test image), `b` the reference image (ssim.rs:91-106). */
float oracle_ssim_f32_literal(const uint8_t *a, const uint8_t *b, size_t npix) {
    const float K1 = 0.01f, K2 = 0.03f;
    const float C1 = (K1 * 255.0f) * (K1 * 255.0f), C2 = (K2 * 255.0f) * (K2 * 255.0f), C3 = C2 / 2.0f;
    float *la = (float *)malloc(npix * sizeof(float)), *lb = (float *)malloc(npix * sizeof(float));
    float ma = 0.0f, mb = 0.0f;
    for (size_t i = 0; i < npix; ++i) {
        la[i] = luma_ssim(a + 4 * i);
        ma += la[i];
    }
    ma /= (float)npix;
    for (size_t i = 0; i < npix; ++i) {
        lb[i] = luma_ssim(b + 4 * i);
        mb += lb[i];
    }
    mb /= (float)npix;
    float va = 0.0f, vb = 0.0f, cov = 0.0f;
    for (size_t i = 0; i < npix; ++i) {
        const float d = la[i] - ma;
        va += d * d;
    }
    for (size_t i = 0; i < npix; ++i) {
        const float d = lb[i] - mb;
        vb += d * d;
    }
    const float denom = (float)(npix - 1);
    const float sa = sqrtf(va / denom), sb = sqrtf(vb / denom);
    for (size_t i = 0; i < npix; ++i) cov += (la[i] - ma) * (lb[i] - mb);
    cov /= denom;
    free(la);
    free(lb);
    const float lum = (2.0f * ma * mb + C1) / (ma * ma + mb * mb + C1);
    const float con = (2.0f * sa * sb + C2) / (sa * sa + sb * sb + C2);
    const float str = (cov + C3) / (sa * sb + C3);
    return lum * con * str;
}

/* ------------------------------------------------------------------------------------ */
/* Pipelines                                                                            */
/* ------------------------------------------------------------------------------------ */

static int check_uniform(const oracle_feature_uniform *f) {
    if (f->width == 0 || f->height == 0) return ORACLE_E_INVALID;
    if (f->blur == 1 && (f->blur_radius < 1 || f->blur_radius > ORACLE_MAX_RADIUS))
        return ORACLE_E_INVALID;
    return ORACLE_OK;
}

int oracle_effects(const oracle_feature_uniform *f, int cs, const uint8_t *in, uint8_t *out) {
    if (check_uniform(f) != ORACLE_OK) return ORACLE_E_INVALID;
    const uint32_t w = f->width, h = f->height;
    const size_t nb = (size_t)w * h * 4;
    uint8_t lut[256];
    oracle_front_lut(cs, f->gamma, lut);
    uint8_t *A = (uint8_t *)malloc(nb);
    /* app_state.rs:95-125: gamma (always), grayscale (+copy), invert (+copy) -> texture A */
    oracle_front_s(in, w, h, lut, f->grayscale == 1, f->invert == 1, A);
    /* image_shader.wgsl:158-174: each enabled effect overwrites `pixels` from A */
    if (f->blur == 1)
        oracle_blur_s(A, w, h, f->blur_radius, out);
    else if (f->sharpen == 1)
        oracle_sharpen_s(A, w, h, f->sharpen_factor, out);
    else if (f->edge_detect == 1)
        oracle_sobel_s(A, w, h, out);
    else
        memcpy(out, A, nb);
    /* FlipX / FlipY: feature_uniform.rs:143-175 flips the sign of the transform's x / y scale and
     * vs_main (image_shader.wgsl:43-59) applies it to the quad position: the finished frame is
     * mirrored, texture-space effects are unaffected. */
    const int flip_x = f->transform[0][0] < 0.0f, flip_y = f->transform[1][1] < 0.0f;
    if (flip_x || flip_y) {
        memcpy(A, out, nb);
        for (uint32_t y = 0; y < h; ++y)
            for (uint32_t x = 0; x < w; ++x)
                memcpy(out + ((size_t)y * w + x) * 4,
                       A + ((size_t)(flip_y ? h - 1 - y : y) * w + (flip_x ? w - 1 - x : x)) * 4, 4);
    }
    free(A);
    return ORACLE_OK;
}

int oracle_chain(const oracle_feature_uniform *f, int cs, const uint8_t *in, uint8_t *out, double *ssim) {
    if (check_uniform(f) != ORACLE_OK) return ORACLE_E_INVALID;
    const uint32_t w = f->width, h = f->height;
    const size_t nb = (size_t)w * h * 4;
    uint8_t lut[256];
    oracle_front_lut(cs, f->gamma, lut);
    uint8_t *A = (uint8_t *)malloc(nb), *B = (uint8_t *)malloc(nb);
    oracle_front_s(in, w, h, lut, f->grayscale == 1, f->invert == 1, A);
    const uint8_t *cur = A;
    if (f->blur == 1) {
        oracle_blur_s(cur, w, h, f->blur_radius, B);
        cur = B;
    }
    if (f->sharpen == 1) {
        uint8_t *dst = cur == A ? B : A;
        oracle_sharpen_s(cur, w, h, f->sharpen_factor, dst);
        cur = dst;
    }
    if (f->edge_detect == 1) {
        uint8_t *dst = cur == A ? B : A;
        oracle_sobel_s(cur, w, h, dst);
        cur = dst;
    }
    memcpy(out, cur, nb);
    if (ssim) *ssim = oracle_ssim(out, in, (size_t)w * h);
    free(A);
    free(B);
    return ORACLE_OK;
}

/* Batch of independent frames (BASELINE config 4): frames in parallel, one thread each. */
int oracle_chain_batch(const oracle_feature_uniform *f, int cs, const uint8_t *in, uint8_t *out,
                       double *ssim, int nframes) {
    if (check_uniform(f) != ORACLE_OK) return ORACLE_E_INVALID;
    const size_t nb = (size_t)f->width * f->height * 4;
    int rc = ORACLE_OK;
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < nframes; ++i) {
        int r = oracle_chain(f, cs, in + (size_t)i * nb, out + (size_t)i * nb, ssim ? ssim + i : NULL);
        if (r != ORACLE_OK) {
#pragma omp critical
            rc = r;
        }
    }
    return rc;
}
