// dequant.cuh -- bit-exact dequantization of one 32-element group of a quantized row, shared by the f32 dequantizer
// (quantize.cu) and the int8 prefill-plane builder (mmq.cu).  Every float op rounds separately (__fmul_rn/__fsub_rn), in
// the reference's operand order, so results are bit-identical to dequantize_row_* (reference: ggml/src/ggml-quants.c:
// 255-273, 349-363, 1280-1302, 1690-1722, 2454-2476; get_scale_min_k4 :631-639).
#pragma once

#include "common.cuh"

namespace b200 {
#ifdef __CUDACC__

__device__ __forceinline__ float h2f(const uint8_t * p) {
    return __half2float(__ushort_as_half((unsigned short) (p[0] | (p[1] << 8))));
}

__device__ __forceinline__ void scale_min_k4(int j, const uint8_t * q, int & sc, int & mn) {
    if (j < 4) {
        sc = q[j] & 63;
        mn = q[j + 4] & 63;
    } else {
        sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

__device__ static const int8_t c_iq4nl[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};

// o[0..31] = elements 32*gi .. 32*gi+31 of the row starting at wr
template <int TYPE>
__device__ __forceinline__ void dequant_group32(const uint8_t * __restrict__ wr, int64_t gi, float * o) {
    if (TYPE == B200_TYPE_Q4_0) {
        const uint8_t * b = wr + gi * 18;
        const float     d = h2f(b);
        for (int j = 0; j < 16; j++) {
            o[j]      = __fmul_rn((float) ((b[2 + j] & 0xF) - 8), d);
            o[j + 16] = __fmul_rn((float) ((b[2 + j] >> 4) - 8), d);
        }
    } else if (TYPE == B200_TYPE_Q8_0) {
        const uint8_t * b = wr + gi * 34;
        const float     d = h2f(b);
        for (int j = 0; j < 32; j++) o[j] = __fmul_rn((float) (int8_t) b[2 + j], d);
    } else if (TYPE == B200_TYPE_Q4_K) {
        const uint8_t * b  = wr + (gi / 8) * 144;
        const int       sb = (int) (gi % 8);
        const float d = h2f(b), mn = h2f(b + 2);
        int sc, m;
        scale_min_k4(sb, b + Q4K_OFF_SCALES, sc, m);
        const float d1 = __fmul_rn(d, (float) sc), m1 = __fmul_rn(mn, (float) m);
        const uint8_t * q = b + Q4K_OFF_QS + 32 * (sb / 2);
        for (int l = 0; l < 32; l++) {
            const int v = (sb & 1) ? (q[l] >> 4) : (q[l] & 0xF);
            o[l] = __fsub_rn(__fmul_rn(d1, (float) v), m1);
        }
    } else if (TYPE == B200_TYPE_Q6_K) {
        const uint8_t * b    = wr + (gi / 8) * 210;
        const int       u    = (int) (gi % 8);       // 32-group inside the super-block
        const int       half = u / 4, j = u % 4;
        const uint8_t * ql = b + 64 * half + 32 * (j & 1);
        const uint8_t * qh = b + Q6K_OFF_QH + 32 * half;
        const int8_t *  sc = (const int8_t *) (b + Q6K_OFF_SCALES) + 8 * half + 2 * j;
        const float     d  = h2f(b + Q6K_OFF_D);
        for (int l = 0; l < 32; l++) {
            const int lo = (j & 2) ? (ql[l] >> 4) : (ql[l] & 0xF);
            const int hi = (qh[l] >> (2 * j)) & 3;
            const int q  = (int) (int8_t) (lo | (hi << 4)) - 32;
            o[l] = __fmul_rn(__fmul_rn(d, (float) sc[l / 16]), (float) (int8_t) q);
        }
    } else if (TYPE == B200_TYPE_IQ4_XS) {
        const uint8_t * b  = wr + (gi / 8) * 136;
        const int       ib = (int) (gi % 8);
        const float     d  = h2f(b);
        const int       sh = b[IQ4XS_OFF_SCALES_H] | (b[IQ4XS_OFF_SCALES_H + 1] << 8);
        const int       ls = ((b[IQ4XS_OFF_SCALES_L + ib / 2] >> (4 * (ib % 2))) & 0xf) | (((sh >> (2 * ib)) & 3) << 4);
        const float     dl = __fmul_rn(d, (float) (ls - 32));
        const uint8_t * qs = b + IQ4XS_OFF_QS + 16 * ib;
        for (int j = 0; j < 16; j++) {
            o[j]      = __fmul_rn(dl, (float) c_iq4nl[qs[j] & 0xf]);
            o[j + 16] = __fmul_rn(dl, (float) c_iq4nl[qs[j] >> 4]);
        }
    }
}

#endif  // __CUDACC__
}  // namespace b200
