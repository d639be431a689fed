/*
 * oracle/ggml_qref.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see ggml_qref.h for the rules and the pinning status).
 *
 * Plain-C restatement of the reference's quantized MUL_MAT / MUL_MAT_ID arithmetic.  Compile WITHOUT -march flags and
 * with -ffp-contract=off so that every float multiply/subtract rounds separately, exactly like the reference's
 * baseline-x86-64 libggml-base (the reference's CUDA build enables fast-math/FMA; its CPU quantizers do not).
 *
 * Citations: path:line relative to /root/reference/.
 */
#include "ggml_qref.h"

#include <assert.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ block formats (ggml/src/ggml-common.h) */
#define QK 32    /* QK4_0 = QK8_0 = QK8_1          ggml-common.h:166,208,216 */
#define QKK 256  /* QK_K                            ggml-common.h:89 */

#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[QK / 2]; } blk_q4_0;                           /* :167-172  18 B */
typedef struct { uint16_t d; int8_t qs[QK]; } blk_q8_0;                                /* :209-214  34 B */
typedef struct { uint16_t d; uint16_t s; int8_t qs[QK]; } blk_q8_1;                    /* :216-227  36 B */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qs[QKK / 2]; } blk_q4_K; /* :285-296 144 B */
typedef struct { uint8_t ql[QKK / 2]; uint8_t qh[QKK / 4]; int8_t scales[QKK / 16]; uint16_t d; } blk_q6_K; /* :320-326 210 B */
typedef struct { float d; int8_t qs[QKK]; int16_t bsums[QKK / 16]; } blk_q8_K;         /* :329-334 292 B */
typedef struct { uint16_t d; uint16_t scales_h; uint8_t scales_l[QKK / 64]; uint8_t qs[QKK / 2]; } blk_iq4_xs; /* :412-418 136 B */
#pragma pack(pop)

/* ggml-quants.c:2434 / ggml-cuda/common.cuh:455 */
static const int8_t iq4nl_values[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};

int qref_blck_size(int type) {
    switch (type) {
        case QREF_TYPE_F32: return 1;
        case QREF_TYPE_Q4_0: case QREF_TYPE_Q8_0: case QREF_TYPE_Q8_1: return QK;
        case QREF_TYPE_Q4_K: case QREF_TYPE_Q6_K: case QREF_TYPE_Q8_K: case QREF_TYPE_IQ4_XS: return QKK;
        default: return 0;
    }
}

size_t qref_type_size(int type) {
    switch (type) {
        case QREF_TYPE_F32:    return 4;
        case QREF_TYPE_Q4_0:   return sizeof(blk_q4_0);
        case QREF_TYPE_Q8_0:   return sizeof(blk_q8_0);
        case QREF_TYPE_Q8_1:   return sizeof(blk_q8_1);
        case QREF_TYPE_Q4_K:   return sizeof(blk_q4_K);
        case QREF_TYPE_Q6_K:   return sizeof(blk_q6_K);
        case QREF_TYPE_Q8_K:   return sizeof(blk_q8_K);
        case QREF_TYPE_IQ4_XS: return sizeof(blk_iq4_xs);
        default: return 0;
    }
}

size_t qref_row_size(int type, int64_t k) { /* ggml.c ggml_row_size */
    const int bs = qref_blck_size(type);
    assert(bs > 0 && k % bs == 0);
    return (size_t) (k / bs) * qref_type_size(type);
}

/* ------------------------------------------------------------------ fp16 (IEEE binary16, RNE) */
float qref_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    const uint32_t exp  = (h >> 10) & 0x1fu;
    uint32_t       man  = h & 0x3ffu;
    uint32_t       bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal: renormalize */
            int e = -1;
            do { man <<= 1; e++; } while (!(man & 0x400u));
            bits = sign | (uint32_t) (127 - 15 - e) << 23 | (man & 0x3ffu) << 13;
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | man << 13;
    } else {
        bits = sign | (exp + 112) << 23 | man << 13;
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

uint16_t qref_fp32_to_fp16(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t) ((x >> 16) & 0x8000u);
    const uint32_t absx = x & 0x7fffffffu;
    if (absx > 0x7f800000u) return (uint16_t) (sign | 0x7e00u);   /* NaN */
    if (absx >= 0x47800000u) {                                     /* >= 65536: inf; [65520,65536) handled below */
        return (uint16_t) (sign | 0x7c00u);
    }
    if (absx < 0x33000000u) return sign;                           /* < 2^-25 rounds to zero */
    int32_t  e = (int32_t) (absx >> 23) - 127;
    uint32_t m = (absx & 0x7fffffu) | 0x800000u;                  /* 24-bit significand */
    int shift;                                                     /* bits to drop */
    uint32_t he;
    if (e < -14) { shift = 13 + (-14 - e); he = 0; }              /* subnormal half */
    else         { shift = 13;             he = (uint32_t) (e + 15); }
    uint32_t q   = m >> shift;
    uint32_t rem = m & ((1u << shift) - 1);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1))) q++;
    /* q carries the implicit bit for normals: adding he<<10 with the carry propagates exponent overflow correctly */
    uint32_t out = (he == 0) ? q : ((he - 1) << 10) + q;
    return (uint16_t) (sign | out);
}

#define F16(x) qref_fp16_to_fp32(x)

/* ------------------------------------------------------------------ activation quantizers */

/* ggml-quants.c:220-253 quantize_row_q8_1_ref */
void qref_quantize_row_q8_1(const float * x, void * vy, int64_t k) {
    assert(k % QK == 0);
    blk_q8_1 * y = (blk_q8_1 *) vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float a = fabsf(x[i * QK + j]);
            amax = amax > a ? amax : a;   /* MAX(amax, fabsf(v)): NaN in v never replaces amax */
        }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = qref_fp32_to_fp16(d);
        int sum = 0;
        for (int j = 0; j < QK; j++) {
            const float v = x[i * QK + j] * id;
            y[i].qs[j] = (int8_t) roundf(v);
            sum += y[i].qs[j];
        }
        y[i].s = qref_fp32_to_fp16(sum * d);
    }
}

/* ggml-quants.c:194-218 quantize_row_q8_0_ref */
void qref_quantize_row_q8_0(const float * x, void * vy, int64_t k) {
    assert(k % QK == 0);
    blk_q8_0 * y = (blk_q8_0 *) vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float a = fabsf(x[i * QK + j]);
            amax = amax > a ? amax : a;
        }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = qref_fp32_to_fp16(d);
        for (int j = 0; j < QK; j++) y[i].qs[j] = (int8_t) roundf(x[i * QK + j] * id);
    }
}

/* ggml-quants.c:372-377 nearest_int */
static inline int nearest_int(float fval) {
    float val = fval + 12582912.f;
    int i;
    memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

/* ggml-quants.c:2479-2516 quantize_row_q8_K_ref */
void qref_quantize_row_q8_K(const float * x, void * vy, int64_t k) {
    assert(k % QKK == 0);
    blk_q8_K * y = (blk_q8_K *) vy;
    const int64_t nb = k / QKK;
    for (int64_t i = 0; i < nb; i++, x += QKK) {
        float max = 0, amax = 0;
        for (int j = 0; j < QKK; j++) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (!amax) {
            y[i].d = 0;
            memset(y[i].qs, 0, QKK);
            /* reference leaves bsums untouched here; callers pass zeroed or don't read them when d == 0 */
            memset(y[i].bsums, 0, sizeof(y[i].bsums));
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < QKK; j++) {
            const int v = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t) (v < 127 ? v : 127);
        }
        for (int j = 0; j < QKK / 16; j++) {
            int sum = 0;
            for (int l = 0; l < 16; l++) sum += y[i].qs[j * 16 + l];
            y[i].bsums[j] = (int16_t) sum;
        }
        y[i].d = 1 / iscale;
    }
}

/* ggml-quants.c:31-66 quantize_row_q4_0_ref */
void qref_quantize_row_q4_0(const float * x, void * vy, int64_t k) {
    assert(k % QK == 0);
    blk_q4_0 * y = (blk_q4_0 *) vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float v = x[i * QK + j];
            if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
        }
        const float d  = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = qref_fp32_to_fp16(d);
        for (int j = 0; j < QK / 2; j++) {
            const float x0 = x[i * QK + j] * id;
            const float x1 = x[i * QK + QK / 2 + j] * id;
            int a = (int8_t) (x0 + 8.5f); if (a > 15) a = 15;
            int b = (int8_t) (x1 + 8.5f); if (b > 15) b = 15;
            y[i].qs[j] = (uint8_t) ((a & 0xff) | ((b & 0xff) << 4));
        }
    }
}

/* ------------------------------------------------------------------ dequantizers */

/* ggml-quants.c:631-639 get_scale_min_k4 */
static inline void scale_min_k4(int j, const uint8_t * q, uint8_t * sc, uint8_t * mn) {
    if (j < 4) {
        *sc = q[j] & 63;
        *mn = q[j + 4] & 63;
    } else {
        *sc = (uint8_t) ((q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4));
        *mn = (uint8_t) ((q[j + 4] >> 4) | ((q[j] >> 6) << 4));
    }
}

/* the 6-bit scale of 32-element sub-block ib of an IQ4_XS super-block (ggml-quants.c:2465) */
static inline int iq4_xs_ls(const blk_iq4_xs * b, int ib) {
    return ((b->scales_l[ib / 2] >> 4 * (ib % 2)) & 0xf) | (((b->scales_h >> 2 * ib) & 3) << 4);
}

static void dq_q4_0(const blk_q4_0 * x, float * y, int64_t k) { /* ggml-quants.c:255-273 */
    for (int64_t i = 0; i < k / QK; i++) {
        const float d = F16(x[i].d);
        for (int j = 0; j < QK / 2; j++) {
            const int x0 = (x[i].qs[j] & 0x0F) - 8;
            const int x1 = (x[i].qs[j] >> 4) - 8;
            y[i * QK + j]          = x0 * d;
            y[i * QK + j + QK / 2] = x1 * d;
        }
    }
}

static void dq_q8_0(const blk_q8_0 * x, float * y, int64_t k) { /* ggml-quants.c:349-363 */
    for (int64_t i = 0; i < k / QK; i++) {
        const float d = F16(x[i].d);
        for (int j = 0; j < QK; j++) y[i * QK + j] = x[i].qs[j] * d;
    }
}

static void dq_q4_K(const blk_q4_K * x, float * y, int64_t k) { /* ggml-quants.c:1280-1302 */
    for (int64_t i = 0; i < k / QKK; i++) {
        const uint8_t * q   = x[i].qs;
        const float     d   = F16(x[i].d);
        const float     min = F16(x[i].dmin);
        for (int is = 0; is < 8; is += 2, q += 32) {
            uint8_t sc, m;
            scale_min_k4(is, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = min * m;
            scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = min * m;
            for (int l = 0; l < 32; l++) *y++ = d1 * (q[l] & 0xF) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * (q[l] >> 4) - m2;
        }
    }
}

static void dq_q6_K(const blk_q6_K * x, float * y, int64_t k) { /* ggml-quants.c:1690-1722 */
    for (int64_t i = 0; i < k / QKK; i++) {
        const float     d  = F16(x[i].d);
        const uint8_t * ql = x[i].ql;
        const uint8_t * qh = x[i].qh;
        const int8_t *  sc = x[i].scales;
        for (int half = 0; half < 2; half++, y += 128, ql += 64, qh += 32, sc += 8) {
            for (int l = 0; l < 32; l++) {
                const int    is = l / 16;
                const int8_t q1 = (int8_t) ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t) ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                y[l]      = d * sc[is + 0] * q1;
                y[l + 32] = d * sc[is + 2] * q2;
                y[l + 64] = d * sc[is + 4] * q3;
                y[l + 96] = d * sc[is + 6] * q4;
            }
        }
    }
}

static void dq_iq4_xs(const blk_iq4_xs * x, float * y, int64_t k) { /* ggml-quants.c:2454-2476 */
    for (int64_t i = 0; i < k / QKK; i++) {
        const uint8_t * qs = x[i].qs;
        const float     d  = F16(x[i].d);
        for (int ib = 0; ib < QKK / 32; ib++, y += 32, qs += 16) {
            const float dl = d * (iq4_xs_ls(&x[i], ib) - 32);
            for (int j = 0; j < 16; j++) {
                y[j]      = dl * iq4nl_values[qs[j] & 0xf];
                y[j + 16] = dl * iq4nl_values[qs[j] >> 4];
            }
        }
    }
}

void qref_dequantize_row(int type, const void * x, float * y, int64_t k) {
    assert(k % qref_blck_size(type) == 0);
    switch (type) {
        case QREF_TYPE_Q4_0:   dq_q4_0((const blk_q4_0 *) x, y, k); break;
        case QREF_TYPE_Q8_0:   dq_q8_0((const blk_q8_0 *) x, y, k); break;
        case QREF_TYPE_Q4_K:   dq_q4_K((const blk_q4_K *) x, y, k); break;
        case QREF_TYPE_Q6_K:   dq_q6_K((const blk_q6_K *) x, y, k); break;
        case QREF_TYPE_IQ4_XS: dq_iq4_xs((const blk_iq4_xs *) x, y, k); break;
        case QREF_TYPE_Q8_1: {
            const blk_q8_1 * b = (const blk_q8_1 *) x;
            for (int64_t i = 0; i < k / QK; i++)
                for (int j = 0; j < QK; j++) y[i * QK + j] = b[i].qs[j] * F16(b[i].d);
        } break;
        case QREF_TYPE_Q8_K: { /* ggml-quants.c:2518-2527 */
            const blk_q8_K * b = (const blk_q8_K *) x;
            for (int64_t i = 0; i < k / QKK; i++)
                for (int j = 0; j < QKK; j++) y[i * QKK + j] = b[i].d * b[i].qs[j];
        } break;
        default: assert(!"qref_dequantize_row: type not on the path");
    }
}

void qref_dequantize_rows(int type, const void * x, float * y, int64_t nrows, int64_t k) {
    const size_t rs = qref_row_size(type, k);
    for (int64_t r = 0; r < nrows; r++) qref_dequantize_row(type, (const char *) x + r * rs, y + r * k, k);
}

/* ------------------------------------------------------------------ ggml-cpu scalar vec_dot kernels */

/* ggml-cpu-quants.c:1912 (scalar tail) */
static float dot_q4_0_q8_0(int64_t k, const blk_q4_0 * x, const blk_q8_0 * y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < k / QK; ib++) {
        int s0 = 0, s1 = 0;
        for (int j = 0; j < QK / 2; j++) {
            s0 += ((x[ib].qs[j] & 0x0F) - 8) * y[ib].qs[j];
            s1 += ((x[ib].qs[j] >> 4) - 8) * y[ib].qs[j + QK / 2];
        }
        sumf += (s0 + s1) * F16(x[ib].d) * F16(y[ib].d);
    }
    return sumf;
}

/* ggml-cpu-quants.c:3597 (scalar tail) */
static float dot_q8_0_q8_0(int64_t k, const blk_q8_0 * x, const blk_q8_0 * y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < k / QK; ib++) {
        int sumi = 0;
        for (int j = 0; j < QK; j++) sumi += x[ib].qs[j] * y[ib].qs[j];
        sumf += sumi * (F16(x[ib].d) * F16(y[ib].d));
    }
    return sumf;
}

/* ggml-cpu-quants.c:6667 (scalar tail): 8 interleaved float partials, mins through bsums */
static float dot_q4_K_q8_K(int64_t k, const blk_q4_K * x, const blk_q8_K * y) {
    float lanes[8] = {0};
    float sumf = 0;
    for (int64_t i = 0; i < k / QKK; i++) {
        int32_t acc[8] = {0};
        int     msum   = 0;
        for (int sb = 0; sb < 8; sb++) {
            uint8_t sc, mn;
            scale_min_k4(sb, x[i].scales, &sc, &mn);
            msum += (y[i].bsums[2 * sb] + y[i].bsums[2 * sb + 1]) * mn;
            const uint8_t * q  = x[i].qs + 32 * (sb / 2);
            const int8_t *  q8 = y[i].qs + 32 * sb;
            for (int l = 0; l < 32; l++) {
                const int w = (sb & 1) ? (q[l] >> 4) : (q[l] & 0xF);
                acc[l % 8] += sc * (q8[l] * w);
            }
        }
        const float d = F16(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) lanes[l] += d * acc[l];
        sumf -= F16(x[i].dmin) * y[i].d * msum;
    }
    for (int l = 0; l < 8; l++) sumf += lanes[l];
    return sumf;
}

/* ggml-cpu-quants.c:8415 (scalar tail) */
static float dot_q6_K_q8_K(int64_t k, const blk_q6_K * x, const blk_q8_K * y) {
    float lanes[8] = {0};
    for (int64_t i = 0; i < k / QKK; i++) {
        int8_t w[QKK];
        const uint8_t * ql = x[i].ql;
        const uint8_t * qh = x[i].qh;
        int8_t *        a  = w;
        for (int half = 0; half < 2; half++, a += 128, ql += 64, qh += 32) {
            for (int l = 0; l < 32; l++) {
                a[l]      = (int8_t) ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                a[l + 32] = (int8_t) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                a[l + 64] = (int8_t) ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                a[l + 96] = (int8_t) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
            }
        }
        int32_t acc[8] = {0};
        for (int g = 0; g < QKK / 16; g++) {
            const int sc = x[i].scales[g];
            for (int l = 0; l < 16; l++) acc[l % 8] += sc * (y[i].qs[16 * g + l] * w[16 * g + l]);
        }
        const float d = F16(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) lanes[l] += d * acc[l];
    }
    float sumf = 0;
    for (int l = 0; l < 8; l++) sumf += lanes[l];
    return sumf;
}

/* ggml-cpu-quants.c:12665 (scalar tail) */
static float dot_iq4_xs_q8_K(int64_t k, const blk_iq4_xs * x, const blk_q8_K * y) {
    float sumf = 0;
    for (int64_t i = 0; i < k / QKK; i++) {
        const float     d4d8 = F16(x[i].d) * y[i].d;
        const uint8_t * qs   = x[i].qs;
        const int8_t *  q8   = y[i].qs;
        for (int ib = 0; ib < QKK / 32; ib++, qs += 16, q8 += 32) {
            const float dl = d4d8 * (iq4_xs_ls(&x[i], ib) - 32);
            int s1 = 0, s2 = 0;
            for (int j = 0; j < 16; j++) {
                s1 += q8[j] * iq4nl_values[qs[j] & 0xf];
                s2 += q8[j + 16] * iq4nl_values[qs[j] >> 4];
            }
            sumf += dl * (s1 + s2);
        }
    }
    return sumf;
}

static int vec_dot_type(int type) { /* ggml-cpu.c:223-352 */
    return (type == QREF_TYPE_Q4_0 || type == QREF_TYPE_Q8_0) ? QREF_TYPE_Q8_0 : QREF_TYPE_Q8_K;
}

static float dot_cpu(int type, int64_t k, const void * w, const void * xq) {
    switch (type) {
        case QREF_TYPE_Q4_0:   return dot_q4_0_q8_0(k, (const blk_q4_0 *) w, (const blk_q8_0 *) xq);
        case QREF_TYPE_Q8_0:   return dot_q8_0_q8_0(k, (const blk_q8_0 *) w, (const blk_q8_0 *) xq);
        case QREF_TYPE_Q4_K:   return dot_q4_K_q8_K(k, (const blk_q4_K *) w, (const blk_q8_K *) xq);
        case QREF_TYPE_Q6_K:   return dot_q6_K_q8_K(k, (const blk_q6_K *) w, (const blk_q8_K *) xq);
        case QREF_TYPE_IQ4_XS: return dot_iq4_xs_q8_K(k, (const blk_iq4_xs *) w, (const blk_q8_K *) xq);
        default: assert(!"type not on the path"); return 0;
    }
}

/* ------------------------------------------------------------------ q8_1 (CUDA-path) semantics, double accumulate */

static double dot_q8_1(int type, int64_t k, const void * w, const blk_q8_1 * y) {
    double acc = 0;
    switch (type) {
        case QREF_TYPE_Q4_0: { /* vecdotq.cuh:23-42: d4*(sumi*d8 - 8*s8) */
            const blk_q4_0 * x = (const blk_q4_0 *) w;
            for (int64_t ib = 0; ib < k / QK; ib++) {
                int sumi = 0;
                for (int j = 0; j < QK / 2; j++)
                    sumi += (x[ib].qs[j] & 0xF) * y[ib].qs[j] + (x[ib].qs[j] >> 4) * y[ib].qs[j + QK / 2];
                int ysum = 0;
                for (int j = 0; j < QK; j++) ysum += y[ib].qs[j];
                /* s = d*sum(qs) taken before its fp16 rounding (the device keeps it in f32) */
                acc += (double) F16(x[ib].d) * F16(y[ib].d) * ((double) sumi - 8.0 * ysum);
            }
        } break;
        case QREF_TYPE_Q8_0: { /* vecdotq.cuh:151-163 */
            const blk_q8_0 * x = (const blk_q8_0 *) w;
            for (int64_t ib = 0; ib < k / QK; ib++) {
                int sumi = 0;
                for (int j = 0; j < QK; j++) sumi += x[ib].qs[j] * y[ib].qs[j];
                acc += (double) F16(x[ib].d) * F16(y[ib].d) * sumi;
            }
        } break;
        case QREF_TYPE_Q4_K: { /* vecdotq.cuh:357-380, 671-715: d*sum(d8*sc*dot) - dmin*sum(m*s8) */
            const blk_q4_K * x = (const blk_q4_K *) w;
            for (int64_t i = 0; i < k / QKK; i++) {
                double sd = 0, sm = 0;
                for (int sb = 0; sb < 8; sb++) {
                    uint8_t sc, mn;
                    scale_min_k4(sb, x[i].scales, &sc, &mn);
                    const blk_q8_1 * yb = &y[i * 8 + sb];
                    const uint8_t *  q  = x[i].qs + 32 * (sb / 2);
                    int dot = 0;
                    for (int l = 0; l < 32; l++) dot += ((sb & 1) ? (q[l] >> 4) : (q[l] & 0xF)) * yb->qs[l];
                    int ysum = 0;
                    for (int l = 0; l < 32; l++) ysum += yb->qs[l];
                    sd += (double) F16(yb->d) * sc * dot;
                    sm += (double) F16(yb->d) * ysum * mn;   /* s before its fp16 rounding */
                }
                acc += (double) F16(x[i].d) * sd - (double) F16(x[i].dmin) * sm;
            }
        } break;
        case QREF_TYPE_Q6_K: { /* vecdotq.cuh:476-497, 763-787 */
            const blk_q6_K * x = (const blk_q6_K *) w;
            for (int64_t i = 0; i < k / QKK; i++) {
                /* integer weights (q - 32) and their scales, per 16 */
                const uint8_t * ql = x[i].ql;
                const uint8_t * qh = x[i].qh;
                int8_t          a[QKK];
                int8_t *        p = a;
                for (int half = 0; half < 2; half++, p += 128, ql += 64, qh += 32) {
                    for (int l = 0; l < 32; l++) {
                        p[l]      = (int8_t) ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                        p[l + 32] = (int8_t) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                        p[l + 64] = (int8_t) ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                        p[l + 96] = (int8_t) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                    }
                }
                double sd = 0;
                for (int g = 0; g < 16; g++) {
                    const blk_q8_1 * yb = &y[i * 8 + g / 2];
                    int dot = 0;
                    for (int l = 0; l < 16; l++) dot += a[16 * g + l] * yb->qs[16 * (g & 1) + l];
                    sd += (double) F16(yb->d) * x[i].scales[g] * dot;
                }
                acc += (double) F16(x[i].d) * sd;
            }
        } break;
        case QREF_TYPE_IQ4_XS: { /* vecdotq.cuh:1069-1133 */
            const blk_iq4_xs * x = (const blk_iq4_xs *) w;
            for (int64_t i = 0; i < k / QKK; i++) {
                double sd = 0;
                for (int ib = 0; ib < 8; ib++) {
                    const blk_q8_1 * yb = &y[i * 8 + ib];
                    const uint8_t *  qs = x[i].qs + 16 * ib;
                    int dot = 0;
                    for (int j = 0; j < 16; j++)
                        dot += yb->qs[j] * iq4nl_values[qs[j] & 0xf] + yb->qs[j + 16] * iq4nl_values[qs[j] >> 4];
                    sd += (double) F16(yb->d) * (iq4_xs_ls(&x[i], ib) - 32) * dot;
                }
                acc += (double) F16(x[i].d) * sd;
            }
        } break;
        default: assert(!"type not on the path");
    }
    return acc;
}

/* ------------------------------------------------------------------ MUL_MAT / MUL_MAT_ID */

static int pick_threads(int nthreads) {
#ifdef _OPENMP
    return nthreads > 0 ? nthreads : omp_get_max_threads();
#else
    (void) nthreads;
    return 1;
#endif
}

/* semantics: 0 ggml-cpu, 1 q8_1, 2 f64 */
static void mul_mat_impl(int type, const void * W, size_t rs, const float * x, float * dst, int64_t dst_stride,
                         int64_t m, int64_t n, int64_t k, int semantics, int nthreads) {
    if (m == 0 || n == 0) return;
    if (rs == 0) rs = qref_row_size(type, k);
    const int nt = pick_threads(nthreads);
    (void) nt;
    if (semantics == 2) {
#pragma omp parallel num_threads(nt)
        {
            float * wrow = (float *) malloc(sizeof(float) * (size_t) k);
#pragma omp for schedule(static)
            for (int64_t i = 0; i < m; i++) {
                qref_dequantize_row(type, (const char *) W + i * rs, wrow, k);
                for (int64_t j = 0; j < n; j++) {
                    double acc = 0;
                    for (int64_t c = 0; c < k; c++) acc += (double) wrow[c] * (double) x[j * k + c];
                    dst[j * dst_stride + i] = (float) acc;
                }
            }
            free(wrow);
        }
        return;
    }
    const int    qt  = semantics == 0 ? vec_dot_type(type) : QREF_TYPE_Q8_1;
    const size_t qrs = qref_row_size(qt, k);
    char *       xq  = (char *) calloc((size_t) n, qrs);
    for (int64_t j = 0; j < n; j++) {
        if (qt == QREF_TYPE_Q8_0) qref_quantize_row_q8_0(x + j * k, xq + j * qrs, k);
        else if (qt == QREF_TYPE_Q8_K) qref_quantize_row_q8_K(x + j * k, xq + j * qrs, k);
        else qref_quantize_row_q8_1(x + j * k, xq + j * qrs, k);
    }
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int64_t i = 0; i < m; i++) {
        const void * w = (const char *) W + i * rs;
        for (int64_t j = 0; j < n; j++) {
            dst[j * dst_stride + i] = semantics == 0 ? dot_cpu(type, k, w, xq + j * qrs)
                                                     : (float) dot_q8_1(type, k, w, (const blk_q8_1 *) (xq + j * qrs));
        }
    }
    free(xq);
}

void qref_mul_mat(int type, const void * W, size_t rs, const float * x, float * dst, int64_t m, int64_t n, int64_t k,
                  int nthreads) {
    mul_mat_impl(type, W, rs, x, dst, m, m, n, k, 0, nthreads);
}

void qref_mul_mat_q8_1(int type, const void * W, size_t rs, const float * x, float * dst, int64_t m, int64_t n,
                       int64_t k, int nthreads) {
    mul_mat_impl(type, W, rs, x, dst, m, m, n, k, 1, nthreads);
}

void qref_mul_mat_f64(int type, const void * W, size_t rs, const float * x, float * dst, int64_t m, int64_t n,
                      int64_t k, int nthreads) {
    mul_mat_impl(type, W, rs, x, dst, m, m, n, k, 2, nthreads);
}

void qref_mul_mat_id(int type, const void * as, size_t es, const float * b, const int32_t * ids, int64_t ids_stride,
                     float * dst, int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t ne11,
                     int64_t n_tokens, int semantics, int nthreads) {
    if (es == 0) es = qref_row_size(type, k) * (size_t) m;
    for (int64_t t = 0; t < n_tokens; t++) {
        for (int64_t s = 0; s < n_used; s++) {
            const int32_t e = ids[t * ids_stride + s];
            assert(e >= 0 && e < n_expert);
            (void) n_expert;
            const float * xb = b + (t * ne11 + s % ne11) * k;
            mul_mat_impl(type, (const char *) as + (size_t) e * es, 0, xb, dst + (t * n_used + s) * m, m, m, 1, k,
                         semantics, nthreads);
        }
    }
}

/* tests/test-backend-ops.cpp:177-190 */
double qref_nmse(const float * a, const float * b, size_t n) {
    double mse_a_b = 0.0, mse_a_0 = 0.0;
    for (size_t i = 0; i < n; i++) {
        const float d = a[i] - b[i];
        mse_a_b += d * d;
        mse_a_0 += a[i] * a[i];
    }
    return mse_a_b / mse_a_0;
}
