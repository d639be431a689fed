// mmvq.cu -- batch 1..8 quantized GEMV for sm_100a: the decode hot path.
//
// Replaces (reference): mul_mat_vec_q<type,ncols_y> + vec_dot_*_q8_1 + quantize_q8_1 as used by
// ggml_cuda_op_mul_mat_vec_q (ggml/src/ggml-cuda/mmvq.cu:130-206,423-511; vecdotq.cuh:23-42,151-163,357-380,476-497,
// 527-545,604-619,671-715,763-787,1069-1133; quantize.cu:4-38), and for MUL_MAT_ID the host loop of
// ggml_cuda_mul_mat_id (ggml-cuda.cu:1994-2130).
//
// Design (B200-first, not a port):
//  * HBM-bound streaming.  Every warp owns a contiguous range of weight rows, i.e. one contiguous byte stream, and
//    pulls it through its PRIVATE ring of shared-memory slots with 1-D bulk TMA copies (cp.async.bulk, SASS UBLKCP)
//    completing on per-slot mbarriers.  No block-wide synchronisation in the main loop, NW*NST slots (~100 KB/SM) in
//    flight, raw ggml block bytes (18/34/144/210/136-byte blocks) re-sliced from shared memory.
//  * One launch per matmul: the f32 activation row is quantized to q8_1 in the prologue by the same routine the
//    standalone quantizer uses (actquant.cuh), into shared memory, while the first weight slots are already in flight.
//  * Programmatic dependent launch: weight prefetch is issued BEFORE griddepcontrol.wait (weights never depend on
//    the previous kernel), so back-to-back matmuls of a decode step overlap their ramp-up with the predecessor's tail.
//  * int8 dot products with dp4a, fp32 scale math, warp-shuffle reduction, one deterministic writer per output.
#include "actquant.cuh"
#include "kernels.h"

#include <cstdlib>

namespace b200 {

// blocks per ring slot ("segment"): ~4.5 KB of weights, never crossing a row
template <int TYPE> struct seg_cfg;
template <> struct seg_cfg<B200_TYPE_Q4_0>   { static constexpr int blocks = 256; };  // 4608 B, 8192 weights
template <> struct seg_cfg<B200_TYPE_Q8_0>   { static constexpr int blocks = 128; };  // 4352 B, 4096 weights
template <> struct seg_cfg<B200_TYPE_Q4_K>   { static constexpr int blocks = 32;  };  // 4608 B, 8192 weights
template <> struct seg_cfg<B200_TYPE_Q6_K>   { static constexpr int blocks = 16;  };  // 3360 B, 4096 weights
template <> struct seg_cfg<B200_TYPE_IQ4_XS> { static constexpr int blocks = 32;  };  // 4352 B, 8192 weights

template <int TYPE> __host__ __device__ constexpr int slot_bytes() {
    // payload + up to 15 bytes of leading misalignment, rounded to 16, + 16 bytes so that word-granular reads of a
    // 2-byte-aligned tail stay inside the slot
    return ((seg_cfg<TYPE>::blocks * qtraits<TYPE>::bytes + 15 + 15) & ~15) + 16;
}

__device__ __forceinline__ int dp4a(uint32_t a, uint32_t b, int c) { return __dp4a((int) a, (int) b, c); }

// q8 activations in shared memory
struct q8_smem {
    uint32_t qs, ds, s16;  // shared-space addresses of the three sections of column 0
    uint32_t pitch;        // bytes between columns
};

// ------------------------------------------------------------------------------------------------ per-type segment dots
// Each accumulates, into acc[c], this lane's share of  sum_k W[row][k] * X[c][k]  over the `nb` blocks of the segment
// that starts at shared address `seg`; `kb0` is the index of the segment's first block inside the row.

template <int NCOLS>
__device__ __forceinline__ void seg_dot_q4_K(uint32_t seg, int nb, int kb0, const q8_smem & q, int lane, float * acc) {
    for (int u = lane; u < nb * 4; u += 32) {
        const int      b = u >> 2, j = u & 3;
        const uint32_t blk = seg + b * 144;
        const uint4    hdr = lds128(blk);                       // d | dmin<<16, scales[12]
        const float    d = half_bits_to_float(hdr.x & 0xffff), dmin = half_bits_to_float(hdr.x >> 16);
        const uint32_t sh = (j & 1) * 16;
        const uint32_t sa = (hdr.y >> sh) & 0xffff, sb = (hdr.z >> sh) & 0xffff, sc = (hdr.w >> sh) & 0xffff;
        // 6-bit scales / mins of sub-blocks 2j, 2j+1 (reference get_scale_min_k4, ggml-quants.c:631-639)
        const uint32_t sc2 = j < 2 ? (sa & 0x3f3f) : ((sc & 0x0f0f) | ((sa >> 2) & 0x3030));
        const uint32_t mn2 = j < 2 ? (sb & 0x3f3f) : (((sc >> 4) & 0x0f0f) | ((sb >> 2) & 0x3030));
        const float sc0 = (float) (sc2 & 0xff), sc1 = (float) (sc2 >> 8);
        const float mn0 = (float) (mn2 & 0xff), mn1 = (float) (mn2 >> 8);
        const uint4 v0 = lds128(blk + 16 + 32 * j), v1 = lds128(blk + 32 + 32 * j);
        const uint32_t v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        const int sbg = (kb0 + b) * 8 + 2 * j;                  // q8 block of sub-block 2j
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const uint32_t qa = q.qs + c * q.pitch + sbg * 32;
            const uint4 a0 = lds128(qa), a1 = lds128(qa + 16), b0 = lds128(qa + 32), b1 = lds128(qa + 48);
            const uint4 dd = lds128(q.ds + c * q.pitch + sbg * 8);   // d(2j), d*sum(2j), d(2j+1), d*sum(2j+1)
            const uint32_t ua[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const uint32_t ub[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            int dot0 = 0, dot1 = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                dot0 = dp4a(v[i] & 0x0F0F0F0Fu, ua[i], dot0);
                dot1 = dp4a((v[i] >> 4) & 0x0F0F0F0Fu, ub[i], dot1);
            }
            const float s = __uint_as_float(dd.x) * sc0 * (float) dot0 + __uint_as_float(dd.z) * sc1 * (float) dot1;
            const float t = __uint_as_float(dd.y) * mn0 + __uint_as_float(dd.w) * mn1;
            acc[c] += d * s - dmin * t;
        }
    }
}

// codebook lookup of four 4-bit indices (the low 16 bits of h): two 8-entry prmt lookups + a prmt select on bit 3
__device__ __forceinline__ uint32_t iq4nl_lookup4(uint32_t h) {
    const uint32_t sel = h & 0x7777u;
    const uint32_t lo  = __byte_perm(IQ4NL_T0, IQ4NL_T1, sel);
    const uint32_t hi  = __byte_perm(IQ4NL_T2, IQ4NL_T3, sel);
    return __byte_perm(lo, hi, ((h >> 1) & 0x4444u) | 0x3210u);
}

template <int NCOLS>
__device__ __forceinline__ void seg_dot_iq4_xs(uint32_t seg, int nb, int kb0, const q8_smem & q, int lane,
                                               float * acc) {
    for (int u = lane; u < nb * 4; u += 32) {
        const int      b = u >> 2, j = u & 3;
        const uint32_t blk = seg + b * 136;
        const uint2    hdr = lds64(blk);                        // d | scales_h<<16, scales_l[4]
        const float    d  = half_bits_to_float(hdr.x & 0xffff);
        const uint32_t sh = hdr.x >> 16, sl = (hdr.y >> (8 * j)) & 0xff;
        const int ls0 = (int) ((sl & 0xf) | (((sh >> (4 * j)) & 3) << 4)) - 32;
        const int ls1 = (int) ((sl >> 4) | (((sh >> (4 * j + 2)) & 3) << 4)) - 32;
        const uint32_t qa = blk + 8 + 32 * j;                   // 32 bytes: sub-blocks 2j and 2j+1
        const uint2 w0 = lds64(qa), w1 = lds64(qa + 8), w2 = lds64(qa + 16), w3 = lds64(qa + 24);
        const uint32_t w[8] = {w0.x, w0.y, w1.x, w1.y, w2.x, w2.y, w3.x, w3.y};
        uint32_t t[16];  // decoded int8 weights in nibble order: (j, j+16) pairs, matching the interleaved q8 layout
#pragma unroll
        for (int i = 0; i < 8; i++) {
            t[2 * i]     = iq4nl_lookup4(w[i]);
            t[2 * i + 1] = iq4nl_lookup4(w[i] >> 16);
        }
        const int sbg = (kb0 + b) * 8 + 2 * j;
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const uint32_t ya = q.qs + c * q.pitch + sbg * 32;
            const uint4 a0 = lds128(ya), a1 = lds128(ya + 16), b0 = lds128(ya + 32), b1 = lds128(ya + 48);
            const uint4 dd = lds128(q.ds + c * q.pitch + sbg * 8);
            const uint32_t ua[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const uint32_t ub[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            int dot0 = 0, dot1 = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                dot0 = dp4a(t[i], ua[i], dot0);
                dot1 = dp4a(t[8 + i], ub[i], dot1);
            }
            acc[c] += d * (__uint_as_float(dd.x) * (float) (ls0 * dot0) + __uint_as_float(dd.z) * (float) (ls1 * dot1));
        }
    }
}

template <int NCOLS>
__device__ __forceinline__ void seg_dot_q4_0(uint32_t seg, int nb, int kb0, const q8_smem & q, int lane, float * acc) {
    for (int b = lane; b < nb; b += 32) {
        const uint32_t blk = seg + b * 18;                      // 2-byte aligned
        const uint32_t a4 = blk & ~3u, sh = (blk & 2u) * 8u;
        const uint32_t w0 = lds32(a4), w1 = lds32(a4 + 4), w2 = lds32(a4 + 8), w3 = lds32(a4 + 12), w4 = lds32(a4 + 16);
        const float    d  = half_bits_to_float((w0 >> sh) & 0xffff);
        const uint32_t v[4] = {__funnelshift_rc(w0, w1, sh + 16), __funnelshift_rc(w1, w2, sh + 16),
                               __funnelshift_rc(w2, w3, sh + 16), __funnelshift_rc(w3, w4, sh + 16)};
        const int g = kb0 + b;
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const uint32_t ya = q.qs + c * q.pitch + g * 32;
            const uint4 a0 = lds128(ya), a1 = lds128(ya + 16);
            const uint2 dd = lds64(q.ds + c * q.pitch + g * 8);
            const uint32_t ul[4] = {a0.x, a0.y, a0.z, a0.w}, uh[4] = {a1.x, a1.y, a1.z, a1.w};
            int dot = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                dot = dp4a(v[i] & 0x0F0F0F0Fu, ul[i], dot);
                dot = dp4a((v[i] >> 4) & 0x0F0F0F0Fu, uh[i], dot);
            }
            // sum (q - 8) * y = sum q*y - 8 * sum y
            acc[c] += d * (__uint_as_float(dd.x) * (float) dot - 8.0f * __uint_as_float(dd.y));
        }
    }
}

template <int NCOLS>
__device__ __forceinline__ void seg_dot_q8_0(uint32_t seg, int nb, int kb0, const q8_smem & q, int lane, float * acc) {
    for (int b = lane; b < nb; b += 32) {
        const uint32_t blk = seg + b * 34;                      // 2-byte aligned
        const float    d   = half_bits_to_float(lds16(blk));
        const uint32_t qa = blk + 2, a4 = qa & ~3u, sh = (qa & 2u) * 8u;
        uint32_t w[9];
#pragma unroll
        for (int i = 0; i < 9; i++) w[i] = lds32(a4 + 4 * i);
        uint32_t v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = __funnelshift_rc(w[i], w[i + 1], sh);
        const int g = kb0 + b;
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const uint32_t ya = q.qs + c * q.pitch + g * 32;
            const uint4 a0 = lds128(ya), a1 = lds128(ya + 16);
            const float d8 = __uint_as_float(lds32(q.ds + c * q.pitch + g * 8));
            const uint32_t u[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            int dot = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) dot = dp4a(v[i], u[i], dot);
            acc[c] += d * d8 * (float) dot;
        }
    }
}

__device__ __forceinline__ int lds_s8(uint32_t a) {
    int v;
    asm volatile("ld.shared.s8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

template <int NCOLS>
__device__ __forceinline__ void seg_dot_q6_K(uint32_t seg, int nb, int kb0, const q8_smem & q, int lane, float * acc) {
    // unit = (block, 128-half, 16-lane column group l0): 64 weights = four 16-element groups, one per 32-sub-block
    for (int u = lane; u < nb * 4; u += 32) {
        const int      b = u >> 2, half = (u >> 1) & 1, l0 = (u & 1) * 16;
        const uint32_t blk = seg + b * 210;                     // 2-byte aligned
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
        const int sbg = (kb0 + b) * 8 + 4 * half;               // q8 block of sub-block j = 0
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint4 y  = lds128(q.qs + c * q.pitch + (sbg + j) * 32 + l0);
                const float d8 = __uint_as_float(lds32(q.ds + c * q.pitch + (sbg + j) * 8));
                const int   s16 = (int) (short) lds16(q.s16 + c * q.pitch + (sbg + j) * 4 + (l0 >> 3));
                int dot = dp4a(v[j][0], y.x, 0);
                dot = dp4a(v[j][1], y.y, dot);
                dot = dp4a(v[j][2], y.z, dot);
                dot = dp4a(v[j][3], y.w, dot);
                // weights are stored with a +32 offset: sum (v - 32) * y = sum v*y - 32 * sum y
                s += d8 * (float) (sc[j] * (dot - 32 * s16));
            }
            acc[c] += d * s;
        }
    }
}

template <int TYPE, int NCOLS>
__device__ __forceinline__ void seg_dot(uint32_t seg, int nb, int kb0, const q8_smem & q, int lane, float * acc) {
    if (TYPE == B200_TYPE_Q4_K)        seg_dot_q4_K<NCOLS>(seg, nb, kb0, q, lane, acc);
    else if (TYPE == B200_TYPE_IQ4_XS) seg_dot_iq4_xs<NCOLS>(seg, nb, kb0, q, lane, acc);
    else if (TYPE == B200_TYPE_Q4_0)   seg_dot_q4_0<NCOLS>(seg, nb, kb0, q, lane, acc);
    else if (TYPE == B200_TYPE_Q8_0)   seg_dot_q8_0<NCOLS>(seg, nb, kb0, q, lane, acc);
    else                               seg_dot_q6_K<NCOLS>(seg, nb, kb0, q, lane, acc);
}

// ------------------------------------------------------------------------------------------------ the kernel
struct gemv_kargs {
    const uint8_t * W;
    int64_t         w_row_stride, w_expert_stride;
    const float *   x;
    int64_t         x_col_stride;
    const uint8_t * xq;
    float *         dst;
    int64_t         dst_col_stride;
    int             m, k, nst;
    const int32_t * ids;
    int64_t         ids_row_stride;
    int             n_used, ne11;
    int64_t         x_token_stride, x_slot_stride, dst_token_stride, dst_slot_stride;
    int             world, rank;
    int64_t         row_offset;
};
struct gemv_peers {
    float *   dst[16];
    int32_t * flag[16];
    int32_t   epoch;
};

template <int TYPE, int NCOLS, bool PEERS>
__global__ void __launch_bounds__(512, 1) k_gemv(const gemv_kargs a, const gemv_peers peers) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int BYTES = qtraits<TYPE>::bytes;
    constexpr int SEGB  = seg_cfg<TYPE>::blocks;
    constexpr int SLOT  = slot_bytes<TYPE>();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int nst  = a.nst;
    const q8_layout L = q8_soa_layout(a.k);

    // shared memory: [q8 columns][rings: nw * nst slots][mbarriers: nw * nst]
    uint8_t *      q8base = smem;
    const uint32_t ring   = smem_u32(smem) + ((NCOLS * L.bytes + 127) & ~127u) + warp * nst * SLOT;
    const uint32_t bars   = smem_u32(smem) + ((NCOLS * L.bytes + 127) & ~127u) + nw * nst * SLOT + warp * nst * 8;

    // which matrix / activation / output this CTA row (blockIdx.y) works on
    const uint8_t * W   = a.W;
    const float *   x   = a.x;
    float *         dst = a.dst;
    if (a.ids) {  // MUL_MAT_ID: pair p = (token, slot); the ids come from the previous kernel (router top-k)
        pdl_wait();
        const int64_t p = blockIdx.y, tok = p / a.n_used, slot = p % a.n_used;
        const int     e = a.ids[tok * a.ids_row_stride + slot];
        W   += (int64_t) e * a.w_expert_stride;
        x   += tok * a.x_token_stride + (slot % a.ne11) * a.x_slot_stride;
        dst += tok * a.dst_token_stride + slot * a.dst_slot_stride;
    }

    // this warp's rows and work items (row segments)
    const int64_t tw = (int64_t) gridDim.x * nw, gw = (int64_t) blockIdx.x * nw + warp;
    const int     r_begin = (int) (gw * a.m / tw), r_end = (int) ((gw + 1) * a.m / tw);
    const int     nblk = a.k / qtraits<TYPE>::blck;             // blocks per row
    const int     nseg = (nblk + SEGB - 1) / SEGB;
    const int     nitems = (r_end - r_begin) * nseg;

    if (lane == 0) {
        for (int s = 0; s < nst; s++) mbar_init(bars + 8 * s, 1);
        mbar_fence_init();
    }
    __syncwarp();

    // producer state (lane 0 only): next item to issue
    int      p_row = r_begin, p_seg = 0, p_it = 0;
    auto issue = [&]() {
        const int       blk0  = p_seg * SEGB;
        const int       nb    = min(SEGB, nblk - blk0);
        const uint8_t * src   = W + (int64_t) p_row * a.w_row_stride + (int64_t) blk0 * BYTES;
        const uint64_t  g     = (uint64_t) src;
        const uint64_t  g0    = g & ~(uint64_t) 15;
        const uint32_t  span  = (uint32_t) (((g + (uint64_t) nb * BYTES + 15) & ~(uint64_t) 15) - g0);
        const int       slot  = p_it % nst;
        mbar_expect_tx(bars + 8 * slot, span);
        tma_load_1d(ring + slot * SLOT, (const void *) g0, span, bars + 8 * slot);
        if (++p_seg == nseg) { p_seg = 0; p_row++; }
        p_it++;
    };
    if (lane == 0) {
        const int pre = min(nst, nitems);
        for (int i = 0; i < pre; i++) issue();
    }

    // everything above touched only weights; from here on we read what the previous kernel produced
    pdl_wait();
    pdl_launch_dependents();

    // activations -> shared memory, in the SoA q8_1 layout
    if (a.xq) {
        const uint4 * src = reinterpret_cast<const uint4 *>(a.xq);
        uint4 *       dq  = reinterpret_cast<uint4 *>(q8base);
        const int     n16 = NCOLS * L.bytes / 16;
        for (int i = threadIdx.x; i < n16; i += blockDim.x) dq[i] = src[i];
    } else {
        const bool vec = ((uintptr_t) x % 16 == 0) && (a.x_col_stride % 4 == 0);
#pragma unroll 1
        for (int c = 0; c < NCOLS; c++)
            quantize_row_q8_soa<TYPE == B200_TYPE_IQ4_XS>(x + c * a.x_col_stride, a.k, q8base + c * L.bytes, L, warp, nw,
                                                          lane, vec);
    }
    __syncthreads();

    q8_smem q;
    q.qs = smem_u32(q8base), q.ds = q.qs + L.off_ds, q.s16 = q.qs + L.off_s16, q.pitch = L.bytes;

    float acc[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; c++) acc[c] = 0.f;

    int row = r_begin, sgi = 0;
    for (int it = 0; it < nitems; it++) {
        const int slot = it % nst;
        mbar_wait(bars + 8 * slot, (it / nst) & 1);
        const int      blk0 = sgi * SEGB;
        const int      nb   = min(SEGB, nblk - blk0);
        const uint32_t off  = (uint32_t) ((uint64_t) (W + (int64_t) row * a.w_row_stride + (int64_t) blk0 * BYTES) & 15);
        seg_dot<TYPE, NCOLS>(ring + slot * SLOT + off, nb, blk0, q, lane, acc);
        __syncwarp();                                           // all lanes are done reading the slot
        if (lane == 0 && p_it < nitems) issue();
        if (++sgi == nseg) {
#pragma unroll
            for (int c = 0; c < NCOLS; c++) {
                const float v = warp_sum(acc[c]);
                if (lane == 0) {
                    if (PEERS) {
                        for (int r = 0; r < a.world; r++)
                            peers.dst[r][c * a.dst_col_stride + a.row_offset + row] = v;
                    } else {
                        dst[c * a.dst_col_stride + row] = v;
                    }
                }
                acc[c] = 0.f;
            }
            sgi = 0;
            row++;
        }
    }

    if (PEERS) {
        // publish: all of this CTA's peer stores are visible system-wide before its arrival is counted
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            for (int r = 0; r < a.world; r++) atomicAdd_system(peers.flag[r], 1);
        }
    }
}

// ------------------------------------------------------------------------------------------------ host side
static int env_int(const char * name, int dflt) {
    const char * s = getenv(name);
    return s && *s ? atoi(s) : dflt;
}


static int slot_bytes_of(int type) {
    switch (type) {
        case B200_TYPE_Q4_0:   return slot_bytes<B200_TYPE_Q4_0>();
        case B200_TYPE_Q8_0:   return slot_bytes<B200_TYPE_Q8_0>();
        case B200_TYPE_Q4_K:   return slot_bytes<B200_TYPE_Q4_K>();
        case B200_TYPE_Q6_K:   return slot_bytes<B200_TYPE_Q6_K>();
        case B200_TYPE_IQ4_XS: return slot_bytes<B200_TYPE_IQ4_XS>();
    }
    return 0;
}

static constexpr int SMEM_LIMIT = 227 * 1024;

size_t gemv_q8_smem_bytes(int64_t k, int ncols) { return ((size_t) ncols * q8_soa_layout(k).bytes + 127) & ~(size_t) 127; }

int gemv_max_cols(int type, int64_t k) {
    const int slot = slot_bytes_of(type);
    int best = 0;
    for (int n = 1; n <= 8; n++)
        if (gemv_q8_smem_bytes(k, n) + (size_t) 4 * 2 * (slot + 8) <= (size_t) SMEM_LIMIT) best = n;
    return best;
}

template <int TYPE, int NCOLS>
static int launch_gemv_t(const gemv_args & g, cudaStream_t stream) {
    static int sm_count = 0;
    if (!sm_count) {
        int dev = 0;
        B200_CUDA_TRY(cudaGetDevice(&dev));
        B200_CUDA_TRY(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const int    slot = slot_bytes<TYPE>();
    const size_t q8   = gemv_q8_smem_bytes(g.k, NCOLS);
    // warps per CTA and ring depth: as much in flight as shared memory allows, preferring 8 warps x 3 slots
    int nw  = env_int("B200_GEMV_WARPS", 8);
    int nst = env_int("B200_GEMV_STAGES", 3);
    while (q8 + (size_t) nw * nst * (slot + 8) > (size_t) SMEM_LIMIT) {
        if (nst > 2) nst--;
        else if (nw > 4) nw -= 4;
        else return B200_ERR_UNSUPPORTED;
    }
    const size_t smem = q8 + (size_t) nw * nst * (slot + 8);

    gemv_kargs a;
    a.W = g.W, a.w_row_stride = g.w_row_stride, a.w_expert_stride = g.w_expert_stride;
    a.x = g.x, a.x_col_stride = g.x_col_stride, a.xq = g.xq;
    a.dst = g.dst, a.dst_col_stride = g.dst_col_stride;
    a.m = (int) g.m, a.k = (int) g.k, a.nst = nst;
    a.ids = g.ids, a.ids_row_stride = g.ids_row_stride, a.n_used = g.n_used, a.ne11 = g.ne11 > 0 ? g.ne11 : 1;
    a.x_token_stride = g.x_token_stride, a.x_slot_stride = g.x_slot_stride;
    a.dst_token_stride = g.dst_token_stride, a.dst_slot_stride = g.dst_slot_stride;
    a.world = g.world, a.rank = g.rank, a.row_offset = g.row_offset;
    gemv_peers peers = {};
    const bool use_peers = g.world > 1;
    if (use_peers) {
        for (int r = 0; r < g.world && r < 16; r++) peers.dst[r] = g.peer_dst[r], peers.flag[r] = g.peer_flag[r];
        peers.epoch = g.epoch;
    }

    // grid.x: one CTA per SM, but never more warps than rows
    const int64_t pairs = g.ids ? g.npairs : 1;
    int gx = sm_count * env_int("B200_GEMV_CTAS_PER_SM", 1);
    if (g.ids && pairs > 1) gx = (int) ((gx + pairs - 1) / pairs);  // MUL_MAT_ID: the pairs share the machine
    const int64_t max_ctas = (g.m + nw - 1) / nw;
    if (gx > max_ctas) gx = (int) max_ctas;
    if (gx < 1) gx = 1;

    auto kern = use_peers ? k_gemv<TYPE, NCOLS, true> : k_gemv<TYPE, NCOLS, false>;
    B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) gx, (unsigned) pairs, 1);
    cfg.blockDim         = dim3((unsigned) nw * 32, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = env_int("B200_PDL", 1);
    cfg.attrs = attr, cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, a, peers));
    note_launch();
    return B200_OK;
}

template <int TYPE>
static int launch_gemv_n(const gemv_args & g, cudaStream_t stream) {
    switch (g.ncols) {
        case 1: return launch_gemv_t<TYPE, 1>(g, stream);
        case 2: return launch_gemv_t<TYPE, 2>(g, stream);
        case 3: return launch_gemv_t<TYPE, 3>(g, stream);
        case 4: return launch_gemv_t<TYPE, 4>(g, stream);
        case 5: return launch_gemv_t<TYPE, 5>(g, stream);
        case 6: return launch_gemv_t<TYPE, 6>(g, stream);
        case 7: return launch_gemv_t<TYPE, 7>(g, stream);
        case 8: return launch_gemv_t<TYPE, 8>(g, stream);
    }
    return B200_ERR_SHAPE;
}

int launch_gemv(const gemv_args & g, cudaStream_t stream) {
    if (g.m == 0 || g.ncols == 0) return B200_OK;
    switch (g.type) {
        case B200_TYPE_Q4_0:   return launch_gemv_n<B200_TYPE_Q4_0>(g, stream);
        case B200_TYPE_Q8_0:   return launch_gemv_n<B200_TYPE_Q8_0>(g, stream);
        case B200_TYPE_Q4_K:   return launch_gemv_n<B200_TYPE_Q4_K>(g, stream);
        case B200_TYPE_Q6_K:   return launch_gemv_n<B200_TYPE_Q6_K>(g, stream);
        case B200_TYPE_IQ4_XS: return launch_gemv_n<B200_TYPE_IQ4_XS>(g, stream);
    }
    return B200_ERR_TYPE;
}

}  // namespace b200
