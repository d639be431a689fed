// actquant.cuh -- the q8_1 activation quantizer as a warp-level device routine.
//
// Arithmetic follows quantize_row_q8_1_ref (reference: ggml/src/ggml-quants.c:220-253) operation for operation with
// IEEE round-to-nearest intrinsics (no FMA contraction, no fast-math), so the standalone kernel (quantize.cu) is
// bit-exact with the CPU reference, and the GEMV prologue (mmvq.cu), which calls the same routine, multiplies with
// exactly those int8 values and that fp16-rounded scale.
#pragma once

#include "common.cuh"

namespace b200 {

// Internal ("SoA") q8_1 layout the matmul kernels consume, per activation row of k elements:
//   int8   qs[k]            quantized values (for IQ4_XS weights: pair-interleaved inside each 32-block, see below)
//   float2 ds[k/32]         {d, d * sum(qs)}   d = fp16-rounded scale as float; the product is exact integer * d
//   short2 s16[k/32]        sums of qs over the two 16-element halves of the block (used by Q6_K)
// All three sections start 16-byte aligned; qs is padded to whole 64-byte units.
// In SHARED memory (the GEMV kernels' fused quantizer) the 16-byte chunks of qs are rotated inside each 64-byte unit
// (q8_swz): a lane of the GEMV loads the four chunks of ITS unit, lanes are 64 bytes apart, and without the rotation
// every LDS.128 of a warp falls on two groups of four banks (16-way conflict, ~0.5 us per kernel with 16 warps loading at
// once); rotated by (unit >> 1) the 32 lanes spread over all banks (4 wavefronts: the minimum for 512 bytes).
struct q8_layout {
    uint32_t off_ds, off_s16, bytes;
};
__host__ __device__ __forceinline__ q8_layout q8_soa_layout(int64_t k) {
    q8_layout L;
    const uint32_t nb = (uint32_t) (k / 32);
    L.off_ds  = (uint32_t) ((k + 63) & ~(int64_t) 63);
    L.off_s16 = L.off_ds + ((nb * 8 + 15) & ~15u);
    L.bytes   = L.off_s16 + ((nb * 4 + 15) & ~15u);
    return L;
}

// byte offset o inside qs -> where it lives in the shared-memory copy
__host__ __device__ __forceinline__ uint32_t q8_swz(uint32_t o) {
    return (o & ~0x30u) | ((((o >> 4) + (o >> 7)) & 3u) << 4);
}

#ifdef __CUDACC__

struct q8_block_result {
    uint32_t q4;     // this lane's four int8 values, packed little-endian
    float    d;      // block scale before fp16 rounding (what the reference multiplies the sum with)
    int      isum;   // sum of the 32 quantized values   (valid in all 8 lanes of the group)
    int      isum16; // sum over this lane's 16-element half
};

// Eight consecutive lanes quantize one block of 32: lane (sub = lane & 7) holds elements 4*sub .. 4*sub+3.
__device__ __forceinline__ q8_block_result quantize_block32(float4 v) {
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int) roundf(__fmul_rn(v.x, id));
    const int q1 = (int) roundf(__fmul_rn(v.y, id));
    const int q2 = (int) roundf(__fmul_rn(v.z, id));
    const int q3 = (int) roundf(__fmul_rn(v.w, id));
    q8_block_result r;
    r.q4 = (uint32_t) (q0 & 0xff) | ((uint32_t) (q1 & 0xff) << 8) | ((uint32_t) (q2 & 0xff) << 16) |
           ((uint32_t) (q3 & 0xff) << 24);
    int s = q0 + q1 + q2 + q3;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    r.isum16 = s;
    r.isum   = s + __shfl_xor_sync(0xffffffffu, s, 4);
    r.d      = d;
    return r;
}

__device__ __forceinline__ float4 load_f4(const float * p, bool vec) {
    if (vec) return *reinterpret_cast<const float4 *>(p);
    return make_float4(p[0], p[1], p[2], p[3]);
}
// same through L2 only (ld.global.cg): for data another CTA of the SAME kernel may have written (gemvseq.cu)
__device__ __forceinline__ float4 load_f4_cg(const float * p, bool vec) {
    if (vec) return __ldcg(reinterpret_cast<const float4 *>(p));
    return make_float4(__ldcg(p), __ldcg(p + 1), __ldcg(p + 2), __ldcg(p + 3));
}

// Quantize one row of k f32 (k % 32 == 0) into the SoA layout at `out` (generic pointer: shared or global).
// Called by `nwarps` warps of a CTA; warp `warp` takes blocks 4*warp + i*4*nwarps .. (+3).
// INTERLEAVE: store element j of a block at position 2*j (j < 16) or 2*(j-16)+1 (j >= 16), the order in which the
// IQ4_XS kernel's prmt codebook lookup emits weights.
// UNROLL: how many of a warp's iterations issue their global loads back to back before any arithmetic (the GEMV's
// fused prologue uses 4 so that a row costs one L2 round trip instead of one per iteration).
template <bool INTERLEAVE, int UNROLL = 1, bool CG = false, bool SWZ = false>
__device__ __forceinline__ void quantize_row_q8_soa(const float * __restrict__ x, int64_t k, uint8_t * out,
                                                    q8_layout L, int warp, int nwarps, int lane, bool vec) {
    const int nb   = (int) (k / 32);
    const int sub  = lane & 7;
    const int step = 4 * nwarps;
    float2 *  ds   = reinterpret_cast<float2 *>(out + L.off_ds);
    short2 *  s16  = reinterpret_cast<short2 *>(out + L.off_s16);
    for (int b0 = 4 * warp; b0 < nb; b0 += step * UNROLL) {
        float4 vv[UNROLL];
#pragma unroll
        for (int j = 0; j < UNROLL; j++) {
            const int b = b0 + j * step + (lane >> 3);
            vv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (b < nb) vv[j] = CG ? load_f4_cg(x + (int64_t) b * 32 + 4 * sub, vec) : load_f4(x + (int64_t) b * 32 + 4 * sub, vec);
        }
#pragma unroll
        for (int j = 0; j < UNROLL; j++) {
            if (b0 + j * step >= nb) break;                     // warp-uniform
            const int  b      = b0 + j * step + (lane >> 3);
            const bool active = b < nb;
            const q8_block_result r = quantize_block32(vv[j]);
            const int other16 = __shfl_xor_sync(0xffffffffu, r.isum16, 4);
            uint32_t w = r.q4;
            int      pos = 4 * sub;
            if (INTERLEAVE) {
                // lanes sub and sub^4 hold elements j = 4*sub+i and j+16: exchange so that each lane owns 4 adjacent
                // pairs; sub < 4 writes pairs for j, j+1; sub >= 4 writes pairs for j+2, j+3
                const uint32_t peer = __shfl_xor_sync(0xffffffffu, r.q4, 4);
                const uint32_t lo   = sub < 4 ? r.q4 : peer;   // elements j .. j+3      (j = 4*(sub&3))
                const uint32_t hi   = sub < 4 ? peer : r.q4;   // elements j+16 .. j+19
                w   = sub < 4 ? __byte_perm(lo, hi, 0x5140) : __byte_perm(lo, hi, 0x7362);
                pos = 8 * (sub & 3) + (sub < 4 ? 0 : 4);
            }
            if (active) {
                const uint32_t o = (uint32_t) b * 32u + (uint32_t) pos;
                *reinterpret_cast<uint32_t *>(out + (SWZ ? q8_swz(o) : o)) = w;
                if (sub == 0) {
                    const float dh = __half2float(__float2half_rn(r.d));
                    ds[b]  = make_float2(dh, dh * (float) r.isum);
                    s16[b] = make_short2((short) r.isum16, (short) other16);
                }
            }
        }
    }
}

#endif  // __CUDACC__
}  // namespace b200
