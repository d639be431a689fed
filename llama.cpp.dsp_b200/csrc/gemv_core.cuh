// gemv_core.cuh -- device building blocks shared by the GEMV kernels (mmvq.cu: one launch per matmul group;
// gemvseq.cu: one persistent launch per sequence of groups): the q8_1 activation unit a lane keeps in registers, the
// per-type unit dot products (reference arithmetic: vec_dot_*_q8_1, ggml/src/ggml-cuda/vecdotq.cuh:23-42,151-163,357-380,
// 476-497,527-545,604-619,671-715,763-787,1069-1133) and the transposing butterfly reduction.
#pragma once

#include "actquant.cuh"

namespace b200 {

__device__ __forceinline__ int dp4a(uint32_t a, uint32_t b, int c) { return __dp4a((int) a, (int) b, c); }
// unsigned bytes of a  x  signed bytes of b (the high nibbles of packed 4-bit weights are used in place, as 16 * q)
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// ------------------------------------------------------------------------------------------------ q8_1 unit in registers
// One unit = 64 consecutive weights of a row as the weight format groups them:
//   Q4_K / IQ4_XS / Q4_0 / Q8_0 : two whole 32-blocks (elements 64u .. 64u+63)
//   Q6_K                        : four 16-element groups, elements 256b + 128h + 32j + l0 + [0,16), j = 0..3
struct q8_unit {
    uint32_t q[16];  // int8 x 64
    float    d[4];   // block scales (fp16-rounded): [0..1] for two-block types, [0..3] for Q6_K
    float    s[4];   // two-block types: d * sum(q) per block; Q6_K: 32 * sum(q) of the 16-group, as float
};

__device__ __forceinline__ int q8_round(float v) { return (int) roundf(v); }

// quantize 32 floats held by this lane; returns the fp16-rounded scale, packs to q[0..7]; *isum = sum of the int8s
template <bool INTERLEAVE>
__device__ __forceinline__ float quant32_inlane(const float * v, uint32_t * q, int * isum) {
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i++) amax = fmaxf(amax, fabsf(v[i]));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    int qi[32];
    int sum = 0;
#pragma unroll
    for (int i = 0; i < 32; i++) {
        qi[i] = q8_round(__fmul_rn(v[i], id));
        sum += qi[i];
    }
#pragma unroll
    for (int w = 0; w < 8; w++) {
        int a, b, c, e;
        if (INTERLEAVE) { a = qi[2 * w], b = qi[16 + 2 * w], c = qi[2 * w + 1], e = qi[16 + 2 * w + 1]; }
        else            { a = qi[4 * w], b = qi[4 * w + 1], c = qi[4 * w + 2], e = qi[4 * w + 3]; }
        q[w] = (uint32_t) (a & 0xff) | ((uint32_t) (b & 0xff) << 8) | ((uint32_t) (c & 0xff) << 16) |
               ((uint32_t) (e & 0xff) << 24);
    }
    *isum = sum;
    return __half2float(__float2half_rn(d));
}

__device__ __forceinline__ void load32(const float * p, float * v, bool vec) {
    if (vec) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float4 t = __ldg(reinterpret_cast<const float4 *>(p) + i);
            v[4 * i] = t.x, v[4 * i + 1] = t.y, v[4 * i + 2] = t.z, v[4 * i + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = __ldg(p + i);
    }
}

// Build this lane's unit from f32 activations.  `ug` = unit index in the row, k = row length.
template <int TYPE>
__device__ __forceinline__ void make_unit_f32(q8_unit & u, const float * __restrict__ x, int k, int ug, bool vec) {
    if (TYPE == B200_TYPE_Q6_K) {
        // groups j = 0..3 are the 16 elements at 256b + 128h + 32j + l0; the other half of each 32-block is held by
        // the neighbouring lane (ug ^ 1), which is always valid because k % 256 == 0
        const int base = (ug >> 2) * 256 + ((ug >> 1) & 1) * 128 + (ug & 1) * 16;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float v[16];
            const float * p = x + base + 32 * j;
            if (vec) {
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const float4 t = __ldg(reinterpret_cast<const float4 *>(p) + i);
                    v[4 * i] = t.x, v[4 * i + 1] = t.y, v[4 * i + 2] = t.z, v[4 * i + 3] = t.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = __ldg(p + i);
            }
            float amax = 0.f;
#pragma unroll
            for (int i = 0; i < 16; i++) amax = fmaxf(amax, fabsf(v[i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            const float d  = __fdiv_rn(amax, 127.0f);
            const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
            int sum = 0;
#pragma unroll
            for (int w = 0; w < 4; w++) {
                const int a = q8_round(__fmul_rn(v[4 * w], id)), b = q8_round(__fmul_rn(v[4 * w + 1], id));
                const int c = q8_round(__fmul_rn(v[4 * w + 2], id)), e = q8_round(__fmul_rn(v[4 * w + 3], id));
                sum += a + b + c + e;
                u.q[4 * j + w] = (uint32_t) (a & 0xff) | ((uint32_t) (b & 0xff) << 8) | ((uint32_t) (c & 0xff) << 16) |
                                 ((uint32_t) (e & 0xff) << 24);
            }
            u.d[j] = __half2float(__float2half_rn(d));
            u.s[j] = (float) (32 * sum);
        }
    } else {
        const int nblk = k / 32;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int b = 2 * ug + h;
            if (b < nblk) {
                float v[32];
                load32(x + (int64_t) b * 32, v, vec);
                int isum;
                u.d[h] = quant32_inlane<TYPE == B200_TYPE_IQ4_XS>(v, u.q + 8 * h, &isum);
                u.s[h] = u.d[h] * (float) isum;
            } else {  // k % 64 == 32 (Q4_0 / Q8_0 only): the unit's second block does not exist
#pragma unroll
                for (int i = 0; i < 8; i++) u.q[8 * h + i] = 0;
                u.d[h] = 0.f, u.s[h] = 0.f;
            }
        }
    }
}

// Build the unit from activations already quantized into the SoA layout (actquant.cuh) in global memory.
template <int TYPE, bool SMEM = false>
__device__ __forceinline__ void make_unit_q8(q8_unit & u, const uint8_t * __restrict__ xq, q8_layout L, int k, int ug) {
    // SMEM: xq points into shared memory (filled by the CTA's cooperative quantizer with the chunk rotation q8_swz): no
    // read-only-cache loads, rotated chunk addresses
    auto ld16 = [](const uint8_t * p) { return SMEM ? *reinterpret_cast<const uint4 *>(p) : __ldg(reinterpret_cast<const uint4 *>(p)); };
    auto ldf2 = [](const float2 * p) { return SMEM ? *p : __ldg(p); };
    auto lds2 = [](const short2 * p) { return SMEM ? *p : __ldg(p); };
    const float2 * ds  = reinterpret_cast<const float2 *>(xq + L.off_ds);
    const short2 * s16 = reinterpret_cast<const short2 *>(xq + L.off_s16);
    if (TYPE == B200_TYPE_Q6_K) {
        const int b32 = (ug >> 2) * 8 + ((ug >> 1) & 1) * 4, l0 = (ug & 1) * 16;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t o = (uint32_t) ((b32 + j) * 32 + l0);
            const uint4 t = ld16(xq + (SMEM ? q8_swz(o) : o));
            u.q[4 * j] = t.x, u.q[4 * j + 1] = t.y, u.q[4 * j + 2] = t.z, u.q[4 * j + 3] = t.w;
            u.d[j] = ldf2(&ds[b32 + j]).x;
            const short2 ss = lds2(&s16[b32 + j]);
            u.s[j] = (float) (32 * (int) (l0 ? ss.y : ss.x));
        }
    } else {
        const int nblk = k / 32;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int b = 2 * ug + h;
            if (b < nblk) {
                const uint4 t0 = ld16(xq + (SMEM ? q8_swz((uint32_t) b * 32u) : (uint32_t) b * 32u));
                const uint4 t1 = ld16(xq + (SMEM ? q8_swz((uint32_t) b * 32u + 16u) : (uint32_t) b * 32u + 16u));
                u.q[8 * h] = t0.x, u.q[8 * h + 1] = t0.y, u.q[8 * h + 2] = t0.z, u.q[8 * h + 3] = t0.w;
                u.q[8 * h + 4] = t1.x, u.q[8 * h + 5] = t1.y, u.q[8 * h + 6] = t1.z, u.q[8 * h + 7] = t1.w;
                const float2 dd = ldf2(&ds[b]);
                u.d[h] = dd.x, u.s[h] = dd.y;
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++) u.q[8 * h + i] = 0;
                u.d[h] = 0.f, u.s[h] = 0.f;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ per-type unit dots
// `row` = shared address of the row's first block, `ug` = this lane's unit, `k32` = number of 32-blocks in the row.
// Weight bytes are decoded ONCE per unit and dotted with every activation column.

__device__ __forceinline__ uint32_t iq4nl_lookup4(uint32_t h) {
    // four 4-bit codebook indices (low 16 bits of h): two 8-entry prmt lookups + a prmt select on index bit 3
    const uint32_t sel = h & 0x7777u;
    const uint32_t lo  = __byte_perm(IQ4NL_T0, IQ4NL_T1, sel);
    const uint32_t hi  = __byte_perm(IQ4NL_T2, IQ4NL_T3, sel);
    return __byte_perm(lo, hi, ((h >> 1) & 0x4444u) | 0x3210u);
}

__device__ __forceinline__ int lds_s8(uint32_t a) {
    int v;
    asm volatile("ld.shared.s8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

template <int TYPE, int NCOLS>
__device__ __forceinline__ void unit_dot(uint32_t row, int ug, int k32, const q8_unit * u, float * out) {
    if (TYPE == B200_TYPE_Q4_K) {
        const int      j   = ug & 3;
        const uint32_t blk = row + (ug >> 2) * 144;
        const uint4    hdr = lds128(blk);                       // d | dmin<<16, scales[12]
        const uint4    v0 = lds128(blk + 16 + 32 * j), v1 = lds128(blk + 32 + 32 * j);
        const float    d = half_bits_to_float(hdr.x & 0xffff), dmin = half_bits_to_float(hdr.x >> 16);
        const uint32_t sh = (j & 1) * 16;
        const uint32_t sa = (hdr.y >> sh) & 0xffff, sb = (hdr.z >> sh) & 0xffff, sc = (hdr.w >> sh) & 0xffff;
        // 6-bit scales / mins of sub-blocks 2j, 2j+1 (reference get_scale_min_k4, ggml-quants.c:631-639)
        const uint32_t sc2 = j < 2 ? (sa & 0x3f3f) : ((sc & 0x0f0f) | ((sa >> 2) & 0x3030));
        const uint32_t mn2 = j < 2 ? (sb & 0x3f3f) : (((sc >> 4) & 0x0f0f) | ((sb >> 2) & 0x3030));
        const int   sc0 = sc2 & 0xff, sc1 = sc2 >> 8;
        const float mn0 = (float) (mn2 & 0xff), mn1 = (float) (mn2 >> 8);
        const uint32_t v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        uint32_t lo[8], hi[8];  // hi keeps the nibbles in place (16 * q): the exact >> 4 is applied to the dot product
#pragma unroll
        for (int i = 0; i < 8; i++) lo[i] = v[i] & 0x0F0F0F0Fu, hi[i] = v[i] & 0xF0F0F0F0u;
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                a0 = dp4a(lo[i], u[c].q[i], a0);
                a1 = dp4a(lo[4 + i], u[c].q[4 + i], a1);
                b0 = dp4a_us(hi[i], u[c].q[8 + i], b0);
                b1 = dp4a_us(hi[4 + i], u[c].q[12 + i], b1);
            }
            const float s = u[c].d[0] * (float) (sc0 * (a0 + a1)) + u[c].d[1] * (float) (sc1 * ((b0 + b1) >> 4));
            const float t = u[c].s[0] * mn0 + u[c].s[1] * mn1;
            out[c] = d * s - dmin * t;
        }
    } else if (TYPE == B200_TYPE_IQ4_XS) {
        const int      j   = ug & 3;
        const uint32_t blk = row + (ug >> 2) * 136;
        const uint2    hdr = lds64(blk);                        // d | scales_h<<16, scales_l[4]
        const uint32_t qa  = blk + 8 + 32 * j;                  // 32 bytes: sub-blocks 2j and 2j+1
        const uint2 w0 = lds64(qa), w1 = lds64(qa + 8), w2 = lds64(qa + 16), w3 = lds64(qa + 24);
        const float    d  = half_bits_to_float(hdr.x & 0xffff);
        const uint32_t sh = hdr.x >> 16, sl = (hdr.y >> (8 * j)) & 0xff;
        const int ls0 = (int) ((sl & 0xf) | (((sh >> (4 * j)) & 3) << 4)) - 32;
        const int ls1 = (int) ((sl >> 4) | (((sh >> (4 * j + 2)) & 3) << 4)) - 32;
        const uint32_t w[8] = {w0.x, w0.y, w1.x, w1.y, w2.x, w2.y, w3.x, w3.y};
        uint32_t t[16];  // decoded int8 weights in nibble order: (j, j+16) pairs = the interleaved q8 order
#pragma unroll
        for (int i = 0; i < 8; i++) t[2 * i] = iq4nl_lookup4(w[i]), t[2 * i + 1] = iq4nl_lookup4(w[i] >> 16);
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                a0 = dp4a(t[i], u[c].q[i], a0);
                a1 = dp4a(t[4 + i], u[c].q[4 + i], a1);
                b0 = dp4a(t[8 + i], u[c].q[8 + i], b0);
                b1 = dp4a(t[12 + i], u[c].q[12 + i], b1);
            }
            out[c] = d * (u[c].d[0] * (float) (ls0 * (a0 + a1)) + u[c].d[1] * (float) (ls1 * (b0 + b1)));
        }
    } else if (TYPE == B200_TYPE_Q4_0) {
        // two 18-byte blocks = 36 bytes, 2-byte aligned: ten aligned words + clamped funnel shifts
        const uint32_t a  = row + ug * 36;
        const uint32_t a4 = a & ~3u, sh = (a & 2u) * 8u;
        uint32_t w[10];
#pragma unroll
        for (int i = 0; i < 10; i++) w[i] = lds32(a4 + 4 * i);
        uint32_t s[9];  // the 36 bytes as aligned words
#pragma unroll
        for (int i = 0; i < 9; i++) s[i] = __funnelshift_rc(w[i], w[i + 1], sh);
        const bool  two = 2 * ug + 1 < k32;
        const float d0 = half_bits_to_float(s[0] & 0xffff);
        const float d1 = two ? half_bits_to_float(s[4] >> 16) : 0.f;
        uint32_t v0[4], v1[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            v0[i] = __funnelshift_r(s[i], s[i + 1], 16);        // bytes 2..17
            v1[i] = s[5 + i];                                    // bytes 20..35
        }
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                a0 = dp4a(v0[i] & 0x0F0F0F0Fu, u[c].q[i], a0);
                a1 = dp4a_us(v0[i] & 0xF0F0F0F0u, u[c].q[4 + i], a1);   // 16 * high nibbles
                b0 = dp4a(v1[i] & 0x0F0F0F0Fu, u[c].q[8 + i], b0);
                b1 = dp4a_us(v1[i] & 0xF0F0F0F0u, u[c].q[12 + i], b1);
            }
            // sum (q - 8) * y = sum q*y - 8 * sum y
            out[c] = d0 * (u[c].d[0] * (float) (a0 + (a1 >> 4)) - 8.0f * u[c].s[0]) +
                     d1 * (u[c].d[1] * (float) (b0 + (b1 >> 4)) - 8.0f * u[c].s[1]);
        }
    } else if (TYPE == B200_TYPE_Q8_0) {
        // two 34-byte blocks = 68 bytes, 2-byte aligned
        const uint32_t a  = row + ug * 68;
        const uint32_t a4 = a & ~3u, sh = (a & 2u) * 8u;
        uint32_t w[18];
#pragma unroll
        for (int i = 0; i < 18; i++) w[i] = lds32(a4 + 4 * i);
        uint32_t s[17];
#pragma unroll
        for (int i = 0; i < 17; i++) s[i] = __funnelshift_rc(w[i], w[i + 1], sh);
        const bool  two = 2 * ug + 1 < k32;
        const float d0 = half_bits_to_float(s[0] & 0xffff);
        const float d1 = two ? half_bits_to_float(s[8] >> 16) : 0.f;
        uint32_t v0[8], v1[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            v0[i] = __funnelshift_r(s[i], s[i + 1], 16);        // bytes 2..33
            v1[i] = s[9 + i];                                    // bytes 36..67
        }
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                a0 = dp4a(v0[i], u[c].q[i], a0);
                a1 = dp4a(v0[4 + i], u[c].q[4 + i], a1);
                b0 = dp4a(v1[i], u[c].q[8 + i], b0);
                b1 = dp4a(v1[4 + i], u[c].q[12 + i], b1);
            }
            out[c] = d0 * u[c].d[0] * (float) (a0 + a1) + d1 * u[c].d[1] * (float) (b0 + b1);
        }
    } else {  // Q6_K
        const int      half = (ug >> 1) & 1, l0 = (ug & 1) * 16;
        const uint32_t blk  = row + (ug >> 2) * 210;            // 2-byte aligned
        const uint4 qa = lds128_align2(blk + 64 * half + l0);
        const uint4 qb = lds128_align2(blk + 64 * half + 32 + l0);
        const uint4 qh = lds128_align2(blk + Q6K_OFF_QH + 32 * half + l0);
        const float d  = half_bits_to_float(lds16(blk + Q6K_OFF_D));
        const uint32_t sca = blk + Q6K_OFF_SCALES + 8 * half + (l0 >> 4);
        const int sc[4] = {lds_s8(sca), lds_s8(sca + 2), lds_s8(sca + 4), lds_s8(sca + 6)};
        const uint32_t A[4] = {qa.x, qa.y, qa.z, qa.w}, B[4] = {qb.x, qb.y, qb.z, qb.w}, H[4] = {qh.x, qh.y, qh.z, qh.w};
        uint32_t v[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            v[0][i] = (A[i] & 0x0F0F0F0Fu) | ((H[i] << 4) & 0x30303030u);
            v[1][i] = (B[i] & 0x0F0F0F0Fu) | ((H[i] << 2) & 0x30303030u);
            v[2][i] = ((A[i] >> 4) & 0x0F0F0F0Fu) | (H[i] & 0x30303030u);
            v[3][i] = ((B[i] >> 4) & 0x0F0F0F0Fu) | ((H[i] >> 2) & 0x30303030u);
        }
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                int dot = dp4a(v[j][0], u[c].q[4 * j], 0);
                dot = dp4a(v[j][1], u[c].q[4 * j + 1], dot);
                dot = dp4a(v[j][2], u[c].q[4 * j + 2], dot);
                dot = dp4a(v[j][3], u[c].q[4 * j + 3], dot);
                // weights carry a +32 offset: sum (v - 32) * y = sum v*y - 32 * sum y
                s += u[c].d[j] * (float) sc[j] * ((float) dot - u[c].s[j]);
            }
            out[c] = d * s;
        }
    }
}

// Sum V per-lane values over the 32 lanes of a warp with log2(V) exchange steps + the remaining plain steps.
// On return p[0] holds the warp-wide sum of row `butterfly_row<V>(lane)`; lanes with (lane & (32/V - 1)) == 0 write it.
template <int V>
__device__ __forceinline__ void butterfly_rows(float (&p)[V], int lane) {
    static_assert(V == 2 || V == 4 || V == 8, "V");
    int mask = 16;
#pragma unroll
    for (int n = V / 2; n >= 1; n >>= 1) {
        const bool up = (lane & mask) != 0;
#pragma unroll
        for (int i = 0; i < n; i++) {
            const float send = up ? p[i] : p[i + n];
            const float keep = up ? p[i + n] : p[i];
            p[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
        }
        mask >>= 1;
    }
#pragma unroll
    for (; mask >= 1; mask >>= 1) p[0] += __shfl_xor_sync(0xffffffffu, p[0], mask);
}
template <int V>
__device__ __forceinline__ int butterfly_row(int lane) {
    return V == 8 ? (((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1))
         : V == 4 ? (((lane >> 4) & 1) * 2 + ((lane >> 3) & 1))
                  : ((lane >> 4) & 1);
}


}  // namespace b200
