// mmq.cu -- batched (n > 8) quantized matmul on the 5th-generation tensor cores: tcgen05.mma kind::i8 with TMEM
// accumulators, TMA-fed shared-memory tiles, fp32 scale-accumulate epilogue.  The prefill hot path.
//
// Replaces (reference): quantize_mmq_q8_1 + mul_mat_q<type,mmq_x,8,need_check> + load_tiles_* + vec_dot_*_mma +
// mul_mat_q_stream_k_fixup as used by ggml_cuda_op_mul_mat_q (ggml/src/ggml-cuda/quantize.cu:40-127,143-169;
// mmq.cuh:231-284,563-616,1338-1432,1608-1688,2243-2301,648-727,759-835,1723-1831,2516-2762; mmq.cu:3-97).
//
// Design (B200-first, not a port):
//  * The reference applies one fp32 scale per 32 weights x one per 32 activations after EVERY K=32 mma.sync, on CUDA
//    cores.  On B200 an M128 x N256 x K32 int8 MMA takes ~128 clocks while scaling its 32768 outputs takes ~1500 FMA-pipe
//    clocks, so that scheme is capped near 10 % of the tensor peak.  Here scales are made SEPARABLE over
//    chunks of MMQ_KC = 256 elements of k: weights are re-expressed once per tensor as an int8 "prefill plane" with one
//    fp32 scale per (row, chunk) -- w ~= s[i][c] * P[i][k] -- and activations are quantized with one scale per (column,
//    chunk) -- x ~= t[j][c] * Q[j][k], exactly the granularity of ggml-cpu's q8_K, which is what ggml-cpu multiplies the
//    K-quants and IQ4_XS with (ggml-common.h:329-334, ggml-cpu.c:282-297).  Eight K=32 MMAs accumulate a chunk in TMEM as
//    int32; the epilogue then does acc[i][j] += s[i][c] * t[j][c] * D[i][j] in fp32 registers while the tensor core works
//    on the next chunk in the other TMEM buffer.
//  * Precision where a single int8 per chunk is not enough comes from RESIDUAL planes laid side by side along k, so the
//    GEMM stays ONE pass over a longer "virtual k":
//      - Q8_0 weights keep two planes (value + residual of the first rounding): their per-32 fp16 scales have more dynamic
//        range than one int8 per 256 can hold.
//      - activations multiplied with the 32-block types (Q4_0, Q8_0) get two planes as well: ggml-cpu quantizes those
//        activations per 32 (q8_0), so one scale per 256 is 8x coarser than the reference wherever a chunk holds an outlier
//        channel; value + residual is ~15 bits per element and holds the harness bound on such data
//        (profiles/r02/mmq_nmse_schemes_cpu_study.log).
//    The kernel walks a short list of segments (weight-plane chunk offset, activation-plane chunk offset):
//      K-quants / IQ4_XS: W0.X0      Q4_0: W0.X0 + W0.X1      Q8_0: W0.X0 + W0.X1 + W1.X0   (residual x residual dropped)
//    Cost: +1 byte of HBM per weight for the plane (+2 for Q8_0; built lazily, cached by the caller) and a rounding error
//    of 4e-5 .. 1.7e-4 NMSE against ggml-cpu (bar 5e-4; tests/test_gpu_parity.py measures it).
//  * Warp roles (one CTA per SM, 576 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA
//    issuer, warps 2..17 = epilogue (each owns 32 TMEM lanes x a quarter of the tile's columns).
#include "actquant.cuh"
#include "dequant.cuh"
#include "kernels.h"

#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

// ------------------------------------------------------------------------------------------------ layouts
#ifndef B200_MMQ_KC
#define B200_MMQ_KC 256
#endif
// scale granularity along k (one TMEM accumulation chunk).  256 = ggml-cpu's q8_K block, the default: parity with
// ggml-cpu holds on any data.  512 (make EXTRA=-DB200_MMQ_KC=512) halves the TMEM read-outs (+9 % pp512 in round 1) but is
// 2x coarser than the reference: NMSE vs ggml-cpu 7.8e-4 with outlier channels, above the 5e-4 bar
// (profiles/r02/mmq_nmse_schemes_cpu_study.log) -- not for production use.
constexpr int MMQ_KC = B200_MMQ_KC;
constexpr int MMQ_BM = 128;   // rows of W per CTA tile (= TMEM lanes)
constexpr int MMQ_BK = 128;   // int8 elements per shared-memory stage row (= one 128-byte swizzle atom)
constexpr int MMQ_KBC = MMQ_KC / MMQ_BK;   // k-blocks (shared-memory stages) per chunk
constexpr int MMQ_G32 = MMQ_KC / 32;       // 32-element groups per chunk
static_assert(MMQ_KC == 256 || MMQ_KC == 512, "chunk");
// epilogue flavour the chunk size allows: the magic-number conversion is exact for |D| < 2^22 = 256 * 127 * 127 + ...;
// a 512-chunk can exceed that, so it converts with I2FP (half as many conversions per weight as at 256 anyway)
constexpr int MMQ_EPI_DEFAULT = MMQ_KC > 256 ? 0 : 2;
constexpr int MMQ_PEPI_DEFAULT = MMQ_KC > 256 ? 0 : 3;       // persistent kernel: conversion bias kept in TMEM (epi_accumulate)

int mmq_chunk() { return MMQ_KC; }
static inline int64_t mmq_kpad(int64_t k) { return (k + MMQ_KC - 1) / MMQ_KC * MMQ_KC; }
static inline size_t  align256(size_t v) { return (v + 255) & ~(size_t) 255; }

// Q8_0 keeps two weight planes (value + residual); everything else one
int mmq_nplanes(int type) { return type == B200_TYPE_Q8_0 ? 2 : 1; }
// activations against the 32-block types are quantized to two planes (value + residual), see the file header
int mmq_xplanes(int type) { return type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 ? 2 : 1; }
// weight plane(s) of one tensor = [P: m x (nwp * kpad) int8, row-major, plane p in columns [p * kpad, (p+1) * kpad)]
//                                 [s: (nwp * nchunks) x mpad f32, chunk-major so that a tile's 128 row scales of one chunk
//                                  are one contiguous bulk copy]   (mpad = m rounded up to the tile height)
static inline int64_t mmq_mpad(int64_t m) { return (m + MMQ_BM - 1) / MMQ_BM * MMQ_BM; }
size_t mmq_plane_bytes(int type, int64_t m, int64_t k) {
    const int64_t kp = mmq_kpad(k), nwp = mmq_nplanes(type);
    return align256((size_t) m * nwp * kp) + align256((size_t) mmq_mpad(m) * nwp * (kp / MMQ_KC) * 4);
}
// activation scratch = [Q: n x (nxp * kpad) int8][t: (nxp * nchunks) x npad f32]   (npad = n rounded up to 4)
static inline int64_t mmq_npad(int64_t n) { return (n + 3) & ~(int64_t) 3; }
size_t mmq_act_bytes(int64_t n, int64_t k, int nxp) {
    const int64_t kp = mmq_kpad(k);
    return align256((size_t) n * nxp * kp) + align256((size_t) nxp * (kp / MMQ_KC) * mmq_npad(n) * 4);
}

// ------------------------------------------------------------------------------------------------ plane builder
// One thread per (row, 32-element group): dequantize the group bit-exactly, find the chunk's amax over the 8 threads
// of the chunk, re-quantize to int8 with s = amax / 127.  P has row pitch ldp = nwp * kp; plane p occupies columns
// [p * kp, (p+1) * kp) and scale rows [p * nch, (p+1) * nch).
// RESID: build plane 1 = the residual w - s0 * P0 left by plane 0: two int8 planes give ~15 bits of weight precision,
// used for Q8_0 whose 8-bit blocks would otherwise lose their per-32 scales' dynamic range.
template <int TYPE, bool RESID>
__global__ void __launch_bounds__(256) k_plane_build(const uint8_t * __restrict__ w, int64_t w_stride,
                                                     int8_t * __restrict__ P, float * __restrict__ S, int64_t m, int64_t k,
                                                     int64_t kp, int64_t ldp, int64_t mpad) {
    const int64_t gpr = kp / 32, nch = kp / MMQ_KC;
    const int64_t g   = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const bool    in  = g < gpr * m;
    const int64_t row = in ? g / gpr : 0;
    const int64_t gi  = in ? g % gpr : 0;
    float v[32];
    if (in && gi * 32 < k) {
        dequant_group32<TYPE>(w + row * w_stride, gi, v);
    } else {
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = 0.f;
    }
    if (RESID && in) {
        const float    s0 = S[(gi / MMQ_G32) * mpad + row];
        const int8_t * p0 = P + row * ldp + gi * 32;
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = __fsub_rn(v[i], __fmul_rn(s0, (float) p0[i]));
    }
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i++) amax = fmaxf(amax, fabsf(v[i]));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
    if (MMQ_G32 == 16) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 8));
    if (!in) return;
    const float s  = amax / 127.0f;
    const float is = s > 0.f ? 1.0f / s : 0.f;
    uint32_t out[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int a = __float2int_rn(v[4 * i] * is), b = __float2int_rn(v[4 * i + 1] * is);
        const int c = __float2int_rn(v[4 * i + 2] * is), d = __float2int_rn(v[4 * i + 3] * is);
        out[i] = (uint32_t) (a & 0xff) | ((uint32_t) (b & 0xff) << 8) | ((uint32_t) (c & 0xff) << 16) |
                 ((uint32_t) (d & 0xff) << 24);
    }
    uint4 * dst = reinterpret_cast<uint4 *>(P + row * ldp + (RESID ? kp : 0) + gi * 32);
    dst[0] = make_uint4(out[0], out[1], out[2], out[3]);
    dst[1] = make_uint4(out[4], out[5], out[6], out[7]);
    if ((gi % MMQ_G32) == 0) S[((RESID ? nch : 0) + gi / MMQ_G32) * mpad + row] = s;
}

// ------------------------------------------------------------------------------------------------ activation quantizer
// One warp per (column j, chunk c): KC/32 floats per lane, t = amax / 127, Q = rint(x / t).  NXP = 2 adds the residual
// plane: r = x - t * Q (rounded per operation, like the plane builder), t1 = amax(r) / 127, Q1 = rint(r / t1), stored in
// columns [kp, 2 kp) of the row and scale rows [nch, 2 nch).
// Also zeroes the results of a split-K GEMM (mmq_zero: up to 4 dst matrices of n rows, m % 4 == 0, 16-byte aligned rows),
// whose two K halves are then added with red.global.add -- two addends onto zero give the same bits in either order.
struct mmq_zero {
    int     cnt;
    float * p[4];
    int64_t stride[4];   // elements between rows
    int64_t m[4];
};
// MUL_MAT_ID on the tensor cores: output row j of the quantizer is the activation row of the (token, slot) pair that the
// device-side plan sorted to position j (sorted[j] < 0: padding between two experts' ranges -> a row of zeros)
struct mmq_gather {
    const int32_t * sorted;      // nullptr: no gather, row j of x
    int             n_used, ne11;
    int64_t         x_token_stride, x_slot_stride;   // elements
};
template <int NXP>
__global__ void __launch_bounds__(256) k_quantize_x8(const float * __restrict__ x, int64_t x_stride, int8_t * __restrict__ Q,
                                                     float * __restrict__ Tt, int64_t n, int64_t k, int64_t kp,
                                                     int64_t npad, const mmq_zero z, bool vec, const mmq_gather gt) {
    pdl_launch_dependents();   // the GEMM behind us sets up its barriers / TMEM and prefetches its tensor maps meanwhile
    pdl_wait();                // x comes from the previous kernel, and dst may still be read by it
    if (z.cnt) {
        const int64_t tid = (int64_t) blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t) gridDim.x * blockDim.x;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i < z.cnt) {
                const int64_t m4 = z.m[i] / 4;
                for (int64_t e = tid; e < n * m4; e += nthr)
                    reinterpret_cast<float4 *>(z.p[i] + (e / m4) * z.stride[i])[e % m4] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    const int     lane = threadIdx.x & 31;
    const int64_t wid  = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nch  = kp / MMQ_KC;
    if (wid >= n * nch) return;
    const int64_t j = wid / nch, c = wid % nch;
    constexpr int PL = MMQ_KC / 32;                             // elements per lane: 8 or 16
    const int64_t e0 = c * MMQ_KC + lane * PL;
    float v[PL];
    const float * xr = x + j * x_stride + e0;
    bool          pad_row = false;
    if (gt.sorted) {
        const int p = gt.sorted[j];
        pad_row = p < 0;
        if (!pad_row) xr = x + (int64_t) (p / gt.n_used) * gt.x_token_stride + (int64_t) ((p % gt.n_used) % gt.ne11) * gt.x_slot_stride + e0;
    }
    if (pad_row) {
#pragma unroll
        for (int i = 0; i < PL; i++) v[i] = 0.f;
    } else if (vec && e0 + PL <= k) {                           // 16-byte loads (x and its row stride 16-byte aligned)
#pragma unroll
        for (int i = 0; i < PL / 4; i++) {
            const float4 f = __ldg(reinterpret_cast<const float4 *>(xr) + i);
            v[4 * i] = f.x, v[4 * i + 1] = f.y, v[4 * i + 2] = f.z, v[4 * i + 3] = f.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < PL; i++) v[i] = e0 + i < k ? xr[i] : 0.f;
    }
#pragma unroll
    for (int pl = 0; pl < NXP; pl++) {
        float amax = 0.f;
#pragma unroll
        for (int i = 0; i < PL; i++) amax = fmaxf(amax, fabsf(v[i]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float t  = amax / 127.0f;
        const float it = t > 0.f ? 1.0f / t : 0.f;
        uint32_t o2[PL / 4];
#pragma unroll
        for (int h = 0; h < PL / 4; h++) {
            int q[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                q[e] = __float2int_rn(v[4 * h + e] * it);
                if (pl + 1 < NXP) v[4 * h + e] = __fsub_rn(v[4 * h + e], __fmul_rn(t, (float) q[e]));   // residual
            }
            o2[h] = (uint32_t) (q[0] & 0xff) | ((uint32_t) (q[1] & 0xff) << 8) | ((uint32_t) (q[2] & 0xff) << 16) |
                    ((uint32_t) (q[3] & 0xff) << 24);
        }
        int8_t * qd = Q + j * (NXP * kp) + pl * kp + e0;
        if (PL == 8) *reinterpret_cast<uint2 *>(qd) = make_uint2(o2[0], o2[1]);
        else *reinterpret_cast<uint4 *>(qd) = make_uint4(o2[0], o2[1], o2[PL / 4 - 2], o2[PL / 4 - 1]);
        if (lane == 0) Tt[(pl * nch + c) * npad + j] = t;
    }
}

// ------------------------------------------------------------------------------------------------ MUL_MAT_ID plan (tensor cores)
// Replaces the host side of ggml_cuda_mul_mat_id for batches (ggml-cuda.cu:2008-2011 copies ids to the host and SYNCHRONIZES,
// :2059-2118 gathers rows per expert with k_copy_src1_to_contiguous, runs one GEMM per expert, scatters back): here one
// CTA sorts the (token, slot) pairs by expert on the device -- every expert's range starts at a multiple of 32 positions so
// that tiles and scale vectors stay 16-byte aligned -- and writes the GEMM's work items; ids never leave the device.
//   plan = [hdr: n_items, positions used, 0, 0][sorted: maxpos ints, -1 = padding][items: max_items x {row tile in the stacked
//           expert matrix, first position, columns, UMMA N}]
constexpr int MMIDQ_PAD = 32, MMIDQ_MAX_EXPERTS = 1024;
static inline int64_t mmidq_maxpos(int64_t npairs, int64_t n_expert) { return npairs + MMIDQ_PAD * (n_expert < npairs ? n_expert : npairs); }
static inline int64_t mmidq_max_items(int64_t npairs, int64_t n_expert, int64_t mt) {
    return mt * ((n_expert < npairs ? n_expert : npairs) + npairs / 256);
}
__global__ void __launch_bounds__(1024) k_mmid_sort(const int32_t * __restrict__ ids, int64_t ids_row_stride, int npairs, int n_used,
                                                    int n_expert, int mt, int maxpos, int32_t * __restrict__ hdr,
                                                    int32_t * __restrict__ sorted, int4 * __restrict__ items) {
    __shared__ int cnt[MMIDQ_MAX_EXPERTS], off[MMIDQ_MAX_EXPERTS], cur[MMIDQ_MAX_EXPERTS], ibase[MMIDQ_MAX_EXPERTS];
    pdl_launch_dependents();
    pdl_wait();                                                 // ids come from the router kernels before us
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) cnt[e] = 0, cur[e] = 0;
    for (int i = threadIdx.x; i < maxpos; i += blockDim.x) sorted[i] = -1;
    __syncthreads();
    for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
        const int e = ids[(int64_t) (p / n_used) * ids_row_stride + p % n_used];
        if (e >= 0 && e < n_expert) atomicAdd(&cnt[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int pos = 0, it = 0;
        for (int e = 0; e < n_expert; e++) {
            off[e] = pos, ibase[e] = it;
            pos += (cnt[e] + MMIDQ_PAD - 1) / MMIDQ_PAD * MMIDQ_PAD;
            it  += (cnt[e] + 255) / 256 * mt;
        }
        hdr[0] = it, hdr[1] = pos, hdr[2] = hdr[3] = 0;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
        const int e = ids[(int64_t) (p / n_used) * ids_row_stride + p % n_used];
        if (e < 0 || e >= n_expert) continue;                   // (an invalid id leaves its dst row untouched)
        sorted[off[e] + atomicAdd(&cur[e], 1)] = p;             // order inside an expert's range does not matter: every
    }                                                           // column of the GEMM is computed on its own
    // items: per expert, 256-position tiles outermost, the expert's m-tiles innermost (they share the activation tile)
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) {
        int it = ibase[e];
        for (int n0 = 0; n0 < cnt[e]; n0 += 256) {
            const int ncols = min(256, cnt[e] - n0), bn = (ncols + 31) / 32 * 32;
            for (int t = 0; t < mt; t++) items[it++] = make_int4(e * mt + t, off[e] + n0, ncols, bn);
        }
    }
}

// ------------------------------------------------------------------------------------------------ PTX helpers (sm_100a)
__device__ __forceinline__ void tma_load_2d(uint32_t dst_smem, const CUtensorMap * map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            dst_smem),
        "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap * map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_result) {  // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result), "n"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {  // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 x int8 -> int32, M x N x 32 per instruction; issued by ONE thread
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier when all previously issued MMAs of this thread have completed (implies fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns: thread t of the warp gets lane (taddr.lane + t), columns taddr.col .. +31
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ldw(uint32_t taddr, uint32_t (&r)[8]) { tmem_ld8(taddr, r); }
__device__ __forceinline__ void tmem_ldw(uint32_t taddr, uint32_t (&r)[32]) { tmem_ld32(taddr, r); }
__device__ __forceinline__ void tmem_ldw(uint32_t taddr, uint32_t (&r)[16]) { tmem_ld16(taddr, r); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// eight columns of this warp's 32 lanes <- one 32-bit value (EPI 3: the accumulator's conversion bias, see below)
// Measured (profiles/r02/ncu_mmq_p_v3_tmem_bias_*): the store's source is a vector of DISTINCT registers, and under the
// 96-register cap ptxas re-creates the constant in them before every store (8 IMAD.MOV per 8 columns), so the instruction
// count per element stays where it was and the gain is small (+1 % on the 8B step, +3.5 % on 28672 x 4096 x 512).  Two
// alternatives were built and dropped: holding the fill registers for the whole kernel (ptxas merges equal values and
// copies them anyway; spills), and refilling from shared memory with tcgen05.cp from the MMA thread (correct, but 32
// copies of 4 KB per chunk run in order with the MMAs at ~100 B/clk: 79 -> 111 us).
__device__ __forceinline__ void tmem_st8_fill(uint32_t taddr, uint32_t v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(v) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// int32 accumulator -> fp32 without the conversion unit.  I2FP issues to the XU pipe (4 lanes/clk per SM sub-partition,
// 8 clocks per warp instruction): ncu showed it 64 % busy with the tensor pipe at 25 % -- the conversion, not the MMA,
// paced the kernel.  |D| <= 256 * 127 * 127 < 2^22, so  bits(1.5 * 2^23) + D  is the float 12582912 + D exactly and one
// FADD removes the bias exactly: same bits as I2FP, on the INT and FMA pipes at full rate.
constexpr uint32_t MAGIC_BITS = 0x4B400000u;
constexpr float    MAGIC_F    = 12582912.0f;
__device__ __forceinline__ float i32_to_f32_magic(uint32_t v) { return __fsub_rn(__uint_as_float(v + MAGIC_BITS), MAGIC_F); }
// packed fp32 pairs (sm_100 FADD2 / FMUL2 / FFMA2): one issue slot per two elements
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float & lo, float & hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
// acc[0..N) += float(r[0..N)) * (sic * t[0..N)); EPI 0 = I2FP, 1 = magic-number conversion, 2 = magic + packed fp32,
// 3 = packed fp32 on accumulators that ALREADY carry the bias: the persistent kernel's epilogue warps refill every
// accumulator column they have read with MAGIC_BITS (tcgen05.st, one instruction per 8 columns) and the tensor core always
// accumulates onto it, so r = MAGIC_BITS + D arrives as the float 12582912 + D and the two integer adds per pair are gone
// (2.5 -> 1.75 issue slots per accumulator element; the epilogue warps are issue-bound, profiles/r02/ncu_mmq_p_v2_*).
// All four produce the same bits (every operation is a separately rounded fp32 op on the same operands).
template <int EPI, int N>
__device__ __forceinline__ void epi_accumulate(const uint32_t (&r)[N], const float4 * tp, float sic, float * acc) {
    if (EPI == 3) {
        const uint64_t s2 = pack2(sic, sic), nm = pack2(-MAGIC_F, -MAGIC_F);
        uint64_t * acc2 = reinterpret_cast<uint64_t *>(acc);
#pragma unroll
        for (int i4 = 0; i4 < N / 4; i4++) {
            const float4 t = tp[i4];
            const uint64_t st0 = mul2(s2, pack2(t.x, t.y)), st1 = mul2(s2, pack2(t.z, t.w));
            const uint64_t f0 = add2(pack2(__uint_as_float(r[4 * i4 + 0]), __uint_as_float(r[4 * i4 + 1])), nm);
            const uint64_t f1 = add2(pack2(__uint_as_float(r[4 * i4 + 2]), __uint_as_float(r[4 * i4 + 3])), nm);
            acc2[2 * i4 + 0] = fma2(f0, st0, acc2[2 * i4 + 0]);
            acc2[2 * i4 + 1] = fma2(f1, st1, acc2[2 * i4 + 1]);
        }
    } else if (EPI == 2) {
        const uint64_t s2 = pack2(sic, sic), nm = pack2(-MAGIC_F, -MAGIC_F);
        uint64_t * acc2 = reinterpret_cast<uint64_t *>(acc);
#pragma unroll
        for (int i4 = 0; i4 < N / 4; i4++) {
            const float4 t = tp[i4];
            const uint64_t st0 = mul2(s2, pack2(t.x, t.y)), st1 = mul2(s2, pack2(t.z, t.w));
            const uint64_t f0 = add2(pack2(__uint_as_float(r[4 * i4 + 0] + MAGIC_BITS), __uint_as_float(r[4 * i4 + 1] + MAGIC_BITS)), nm);
            const uint64_t f1 = add2(pack2(__uint_as_float(r[4 * i4 + 2] + MAGIC_BITS), __uint_as_float(r[4 * i4 + 3] + MAGIC_BITS)), nm);
            acc2[2 * i4 + 0] = fma2(f0, st0, acc2[2 * i4 + 0]);
            acc2[2 * i4 + 1] = fma2(f1, st1, acc2[2 * i4 + 1]);
        }
    } else {
#pragma unroll
        for (int i4 = 0; i4 < N / 4; i4++) {
            const float4 t = tp[i4];
            const float  tt[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float f = EPI == 1 ? i32_to_f32_magic(r[4 * i4 + e]) : (float) (int) r[4 * i4 + e];
                acc[4 * i4 + e] = fmaf(f, sic * tt[e], acc[4 * i4 + e]);
            }
        }
    }
}

// shared-memory matrix descriptor: K-major operand, 128-byte swizzle, rows of 128 bytes, 8-row groups 1024 bytes apart
// (cute::UMMA::SmemDescriptor: start >> 4 at [0,14), LBO >> 4 at [16,30), SBO >> 4 at [32,46), version 1 at [46,48),
// layout SWIZZLE_128B = 2 at [61,64))
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
    return (uint64_t) ((saddr & 0x3FFFFu) >> 4) | ((uint64_t) 1 << 16) | ((uint64_t) (1024 >> 4) << 32) |
           ((uint64_t) 1 << 46) | ((uint64_t) 2 << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D = S32 (2 at [4,6)), A = B = signed int8 (1 at [7,10) and
// [10,13)), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t idesc_i8(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (M >> 4) << 24);
}

// One accumulator chunk of an epilogue warp: CW columns of this warp's 32 TMEM lanes starting at `tbase`, software
// pipelined over LW-column loads -- the TMEM load of group g+1 is in flight while group g is scaled and accumulated
// (tcgen05.wait::ld waits for ALL outstanding loads, so load g+1 is issued right after the wait that delivered g).
// `release()` hands the accumulator buffer back to the tensor core as soon as its last column is in registers.
template <int EPI, int CW, typename Release>
__device__ __forceinline__ void epi_chunk(uint32_t tbase, const float4 * tp, float sic, float * acc, Release release) {
    constexpr int LW = CW > 32 ? 16 : 32;                        // columns per load: two loads' registers must fit
    constexpr int NG = CW / LW;
    uint32_t r[2][LW];
    tmem_ldw(tbase, r[0]);
#pragma unroll
    for (int g = 0; g < NG; g++) {
        tmem_ld_wait();
        if (g + 1 < NG) {
            tmem_ldw(tbase + (g + 1) * LW, r[(g + 1) & 1]);
        } else {
            tc_fence_before();
            __syncwarp();
            release();
        }
        epi_accumulate<EPI, LW>(r[g & 1], tp + g * (LW / 4), sic, acc + g * LW);
    }
}

// The same for the persistent kernel, where 64 accumulators per thread leave room for two 8-column load buffers only
// (96 registers with 18 warps).  NG8 = 8-column groups of this warp's slice (compile time: no per-group predicates, TMEM
// addresses are immediates).
template <int EPI, int NG8, typename Release>
__device__ __forceinline__ void epi_chunk8(uint32_t tbase, const float4 * tp, float sic, float * acc, Release release) {
    uint32_t r[2][8];
    tmem_ldw(tbase, r[0]);
#pragma unroll
    for (int g = 0; g < NG8; g++) {
        tmem_ld_wait();
        if (EPI == 3) tmem_st8_fill(tbase + g * 8, MAGIC_BITS);  // these columns are in registers: re-arm them for the next chunk
        if (g + 1 < NG8) {
            tmem_ldw(tbase + (g + 1) * 8, r[(g + 1) & 1]);
        } else {                                                 // accumulator fully read: hand it back to the tensor core
            if (EPI == 3) tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            release();
        }
        epi_accumulate<EPI, 8>(r[g & 1], tp + g * 2, sic, acc + g * 8);
    }
}

// ------------------------------------------------------------------------------------------------ the GEMM kernel
struct mmq_kargs {
    const float * S;      // weight scales  [nch][mpad]
    const float * Tt;     // activation scales [nch][npad] (+ slack: tiles may overhang)
    float *       dst;    // [n][dst_stride], m contiguous
    int64_t       dst_stride;
    int           m, n, nch, npad, mpad;   // nch = chunks per PLANE (kpad / MMQ_KC)
    // virtual k: nseg segments of nch chunks each; segment s multiplies weight-plane chunks seg_a[s] + i with
    // activation-plane chunks seg_b[s] + i  (W0.X0 | W0.X1 | W1.X0, see the file header)
    int           nseg, seg_a[3], seg_b[3];
    int           ksplit;       // 1, or 2: grid.z halves of the virtual chunk range, results added with red.global.add onto
                                // a zeroed dst (few-tile shapes: m = 4096 at n = 512 fills 64 of 148 SMs otherwise)
    int           dbg;          // experiments (B200_MMQ_DEBUG): 1 = persistent kernel skips the second B box (wrong results,
                                // 2/3 of the L2->SM traffic), 2 = every item reads the A tile of m-tile 0
    long long *   trace;        // -DB200_MMQ_TRACE builds only: 16 clock counters per CTA (tools/mmq_trace.py), else unused
    // several weight matrices against the same activations in one launch (q/k/v, gate/up): the grid's x dimension runs
    // over the m-tiles of all of them; matrix i >= 1 starts at tile tile0[i-1] and is described by the x-fields
    int           nmat;
    int           tile0[3];
    const float * Sx[3];
    float *       dstx[3];
    int64_t       dst_stridex[3];
    int           mx[3], mpadx[3];
};
struct mmq_maps3 {
    CUtensorMap m[3];
};
// virtual chunk -> (weight-plane chunk, activation-plane chunk)
__device__ __forceinline__ void mmq_vchunk(const mmq_kargs & a, int vc, int & ac, int & bc) {
    int seg = 0, i = vc;
    if (i >= a.nch) seg = 1, i -= a.nch;
    if (i >= a.nch) seg = 2, i -= a.nch;
    ac = (seg == 0 ? a.seg_a[0] : seg == 1 ? a.seg_a[1] : a.seg_a[2]) + i;
    bc = (seg == 0 ? a.seg_b[0] : seg == 1 ? a.seg_b[1] : a.seg_b[2]) + i;
}

// Role-level wait accounting (trace builds only): TW(slot, wait) adds the clocks spent in `wait` to counter `slot`.
#ifdef B200_MMQ_TRACE
#define TW(slot, stmt)                                   \
    do {                                                 \
        const long long t0_ = clock64();                 \
        stmt;                                            \
        tw[slot] += clock64() - t0_;                     \
    } while (0)
#else
#define TW(slot, stmt) stmt
#endif

template <int BN> struct mmq_shape {
    static constexpr int NS      = BN == 256 ? 4 : 6;                 // shared-memory stages
    static constexpr int A_BYTES = MMQ_BM * MMQ_BK;                   // 16 KB
    static constexpr int B_BYTES = BN * MMQ_BK;
    static constexpr int STAGE   = A_BYTES + B_BYTES;
    static constexpr int NSC     = 4;                                 // ring of per-chunk scale vectors
    static constexpr int SC_FLTS = MMQ_BM + BN;                       // [128 row scales][BN column scales]
    static constexpr int SMEM    = NS * STAGE + NSC * SC_FLTS * 4 + 1024;   // + alignment slack (128-byte swizzle: 1024)
    // Epilogue warps.  The epilogue is what paces this kernel: the accumulator is read out of TMEM once per 256-chunk at
    // ~64 B/clk per SM, and a warp cannot overlap its own tcgen05.ld with its own scale-accumulate math.  16 warps of 32
    // columns each (BN = 128) let one warp's math run under another's load; at BN = 256 that would need 64 accumulator
    // registers per thread in a 576-thread CTA (96-register cap: it spills), so BN = 256 keeps 8 warps x 128 columns.
    static constexpr int EW      = 16;
    static constexpr int THREADS = (2 + EW) * 32;
    static constexpr int NB      = 512 / BN;                          // TMEM accumulator buffers (all 512 columns in use)
};

template <int BN, int EPI>
__global__ void __launch_bounds__(mmq_shape<BN>::THREADS, 1) k_mmq(const __grid_constant__ CUtensorMap mapA,
                                                const __grid_constant__ CUtensorMap mapB, const mmq_kargs a,
                                                const __grid_constant__ mmq_maps3 mapAx) {
    using SH = mmq_shape<BN>;
    constexpr int NS = SH::NS;
    extern __shared__ uint8_t mmq_smem_raw[];
    constexpr int NB = SH::NB;
    __shared__ __align__(8) uint64_t bars[2 * NS + 2 * NB + 2 * SH::NSC];
    __shared__ uint32_t tmem_slot;

    const int      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tiles = (smem_u32(mmq_smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * NS, bar_tfull = bar_empty + 8 * NS,
                   bar_tempty = bar_tfull + 8 * NB, bar_sfull = bar_tempty + 8 * NB, bar_sempty = bar_sfull + 8 * SH::NSC;
    const uint32_t scales = tiles + NS * SH::STAGE;             // shared address of the scale ring
    // which weight matrix this CTA works on (compile-time indices only: the parameter arrays stay in param space)
    const CUtensorMap * pA = &mapA;
    const float *       S  = a.S;
    float *             dstp = a.dst;
    int64_t             dstride = a.dst_stride;
    int                 mrows = a.m, mpad = a.mpad, tile = (int) blockIdx.x;
    if (a.nmat > 1) {
        int first = 0;
#pragma unroll
        for (int i = 0; i < 3; i++)
            if (i + 1 < a.nmat && (int) blockIdx.x >= a.tile0[i])
                pA = &mapAx.m[i], S = a.Sx[i], dstp = a.dstx[i], dstride = a.dst_stridex[i], mrows = a.mx[i], mpad = a.mpadx[i],
                first = a.tile0[i];
        tile -= first;
    }
    const int m0 = tile * MMQ_BM, n0 = blockIdx.y * BN;
    // this CTA's range [cb, cb + nch) of VIRTUAL chunks: everything below counts chunks and k-blocks locally
    const int nvc = a.nseg * a.nch, cper = (nvc + a.ksplit - 1) / a.ksplit, cb = (int) blockIdx.z * cper;
    const int nch = min(cper, nvc - cb), nkb = MMQ_KBC * nch;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; s++) mbar_init(bar_full + 8 * s, 1), mbar_init(bar_empty + 8 * s, 1);
        for (int b = 0; b < NB; b++) mbar_init(bar_tfull + 8 * b, 1), mbar_init(bar_tempty + 8 * b, SH::EW);
        for (int b = 0; b < SH::NSC; b++) mbar_init(bar_sfull + 8 * b, 1), mbar_init(bar_sempty + 8 * b, SH::EW);
        mbar_fence_init();
        prefetch_tmap(pA);
        prefetch_tmap(&mapB);
    }
    if (warp == 1) tmem_alloc<NB * BN>(smem_u32(&tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    pdl_launch_dependents();
    pdl_wait();                                                 // activations come from the quantizer kernel before us

#ifdef B200_MMQ_TRACE
    long long tw[4] = {0, 0, 0, 0};
    const long long t_start = clock64();
#endif
    if (warp == 0) {
        // ------------------------------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int ac = 0, bc = 0;
            for (int kb = 0; kb < nkb; kb++) {
                const int h = kb % MMQ_KBC;
                if (h == 0) {                                    // this chunk's scale vectors, ahead of its tiles
                    const int c = kb / MMQ_KBC, b = c % SH::NSC;
                    mmq_vchunk(a, cb + c, ac, bc);
                    TW(0, mbar_wait(bar_sempty + 8 * b, ((c / SH::NSC) & 1) ^ 1));
                    const uint32_t sd = scales + b * SH::SC_FLTS * 4;
                    mbar_expect_tx(bar_sfull + 8 * b, SH::SC_FLTS * 4);
                    tma_load_1d(sd, S + (int64_t) ac * mpad + m0, MMQ_BM * 4, bar_sfull + 8 * b);
                    tma_load_1d(sd + MMQ_BM * 4, a.Tt + (int64_t) bc * a.npad + n0, BN * 4, bar_sfull + 8 * b);
                }
                const int s = kb % NS, ph = (kb / NS) & 1;
                TW(1, mbar_wait(bar_empty + 8 * s, ph ^ 1));
                const uint32_t sa = tiles + s * SH::STAGE, sb = sa + SH::A_BYTES;
                mbar_expect_tx(bar_full + 8 * s, SH::STAGE);
                tma_load_2d(sa, pA, (MMQ_KBC * ac + h) * MMQ_BK, m0, bar_full + 8 * s);
                tma_load_2d(sb, &mapB, (MMQ_KBC * bc + h) * MMQ_BK, n0, bar_full + 8 * s);
            }
#ifdef B200_MMQ_TRACE
            if (a.trace) {
                long long * o = a.trace + ((size_t) blockIdx.y * gridDim.x + blockIdx.x) * 16;
                o[0] = clock64() - t_start, o[1] = tw[0], o[2] = tw[1];
            }
#endif
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = idesc_i8(MMQ_BM, BN);
            for (int c = 0; c < nch; c++) {
                const int buf = c % NB;
                TW(0, mbar_wait(bar_tempty + 8 * buf, ((c / NB) & 1) ^ 1));   // epilogue has drained this accumulator
                tc_fence_after();
                for (int h = 0; h < MMQ_KBC; h++) {
                    const int kb = MMQ_KBC * c + h, s = kb % NS, ph = (kb / NS) & 1;
                    TW(1, mbar_wait(bar_full + 8 * s, ph));
                    tc_fence_after();
                    const uint32_t sa = tiles + s * SH::STAGE, sb = sa + SH::A_BYTES;
                    const uint64_t da = smem_desc_sw128(sa), db = smem_desc_sw128(sb);
#pragma unroll
                    for (int kk = 0; kk < MMQ_BK / 32; kk++)   // +32 bytes along k inside the swizzle atom = +2 (>> 4)
                        mma_i8(tmem + buf * BN, da + 2 * kk, db + 2 * kk, idesc, (uint32_t) ((h | kk) != 0));
                    mma_commit(bar_empty + 8 * s);               // stage reusable once these MMAs have read it
                }
                mma_commit(bar_tfull + 8 * buf);                 // chunk accumulated
            }
#ifdef B200_MMQ_TRACE
            if (a.trace) {
                long long * o = a.trace + ((size_t) blockIdx.y * gridDim.x + blockIdx.x) * 16;
                o[4] = clock64() - t_start, o[5] = tw[0], o[6] = tw[1];
            }
#endif
        }
    } else {
        // ------------------------------------------------------------------------------------------ epilogue warps
        constexpr int CW = BN / (SH::EW / 4);                    // columns per warp
        const int q = warp & 3, half = (warp - 2) >> 2;          // TMEM lane quarter (hardware: warp id % 4), column half
        const int row = m0 + q * 32 + lane, col0 = half * CW;
        __align__(8) float acc[CW];
#pragma unroll
        for (int i = 0; i < CW; i++) acc[i] = 0.f;
        const float * sring = reinterpret_cast<const float *>(mmq_smem_raw + (scales - smem_u32(mmq_smem_raw)));
        for (int c = 0; c < nch; c++) {
            const int buf = c % NB, sb = c % SH::NSC;
            TW(0, mbar_wait(bar_sfull + 8 * sb, (c / SH::NSC) & 1));
            const float *  sv  = sring + sb * SH::SC_FLTS;
            const float    sic = sv[q * 32 + lane];
            const float4 * tp  = reinterpret_cast<const float4 *>(sv + MMQ_BM + col0);   // warp-uniform: broadcast reads
            TW(1, mbar_wait(bar_tfull + 8 * buf, (c / NB) & 1));
            tc_fence_after();
            epi_chunk<EPI, CW>(tmem + ((uint32_t) (q * 32) << 16) + buf * BN + col0, tp, sic, acc,
                               [&] { if (lane == 0) mbar_arrive(bar_tempty + 8 * buf); });
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_sempty + 8 * sb);     // scale vectors consumed
        }
#ifdef B200_MMQ_TRACE
        const long long t_loop = clock64();
#endif
        if (row < mrows) {
#pragma unroll
            for (int i = 0; i < CW; i++) {
                const int j = n0 + col0 + i;
                if (j < a.n) {                                   // lanes = consecutive rows: coalesced
                    float * d = dstp + (int64_t) j * dstride + row;
                    if (a.ksplit > 1) atomicAdd(d, acc[i]);      // RED.ADD.F32: fire and forget
                    else *d = acc[i];
                }
            }
        }
#ifdef B200_MMQ_TRACE
        if (a.trace && warp == 2 && lane == 0) {
            long long * o = a.trace + ((size_t) blockIdx.y * gridDim.x + blockIdx.x) * 16;
            o[8] = t_loop - t_start, o[9] = tw[0], o[10] = tw[1], o[11] = clock64() - t_loop;
            unsigned long long g;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
            o[12] = (long long) g;
        }
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<NB * BN>(tmem);
}

// ------------------------------------------------------------------------------------------------ persistent GEMM kernel
// One CTA per SM walks a host-made list of work items (item i -> CTA i % gridDim.x).  Items are 128 x 256 tiles in
// n-fastest order (the CTAs running at one time share a few weight tiles through L2); when the last round would give
// fewer than half of the CTAs a tile, those tiles are cut into 128 x 128 halves so that the round costs half as much
// (gate+up of an 8B layer at n = 512: 448 tiles on 148 SMs = 3 rounds + 4 tiles -> 3 rounds + 8 half tiles).
// Against one CTA per tile this removes, per round, the barrier/TMEM set-up, the pipeline fill and most of the 128 KB
// result store (~6.5 us of a ~25 us round, profiles/r01/mmq_trace_v1.log): the producer and the MMA warp run ahead into
// the next item while the epilogue warps still store the previous one.
// B tiles are always loaded as 128-row boxes (two per k-block for a 256-wide item), so one tensor map serves both widths;
// the UMMA N comes from a per-item instruction descriptor; TMEM holds two 256-column accumulator buffers.
struct mmq_pargs {
    int ntn;       // 256-wide n-tiles per m-tile
    int mtiles;    // m-tiles of all matrices of the launch
    int raster;    // 0: consecutive items walk n first, 1: m first (B200_MMQ_RASTER)
    int n_main;    // 128 x 256 items
    int n_items;   // n_main + 128 x 128 tail items
    // MUL_MAT_ID (k_mmq_p<EPI, true>): the items come from the device-side plan (k_mmid_sort), A is the stacked plane of all
    // experts, B the activation rows gathered in sorted order, and the epilogue scatters column j to the dst row of pair
    // sorted[n0 + j]
    const int4 *    items;
    const int32_t * hdr;         // hdr[0] = number of items
    const int32_t * sorted;
    int             n_used, mt_expert, m_expert;
    int64_t         dst_token_stride, dst_slot_stride;
    // split-mode row (world > 1): every result element is stored at the same offset of every rank's symmetric dst allocation
    // (peer_delta[r] bytes from this rank's, NVLink-mapped; [rank] = 0) straight from the epilogue, and the launch counts one
    // ARRIVAL on every rank's sync block when all its CTAs are done (the protocol of the row-split GEMV, mmvq.cu)
    int             world, rank;
    long long       peer_delta[16];
    int32_t *       peer_flag[16];
    int32_t *       local_count;
};
struct mmq_pshape {
    static constexpr int NS      = 4;
    static constexpr int A_BYTES = MMQ_BM * MMQ_BK;                   // 16 KB
    static constexpr int B_BYTES = 256 * MMQ_BK;                      // 32 KB (two 128-row boxes)
    static constexpr int STAGE   = A_BYTES + B_BYTES;
    static constexpr int NSC     = 8;
    static constexpr int SC_FLTS = MMQ_BM + 256;
    static constexpr int SMEM    = NS * STAGE + NSC * SC_FLTS * 4 + 1024;
    static constexpr int EW      = 16;
    static constexpr int THREADS = (2 + EW) * 32;
    static constexpr int NB      = 2;
};
struct mmq_item {
    int mt, n0, bn;    // global m-tile (over all matrices of the launch), first column, width
};
template <bool MMID>
__device__ __forceinline__ mmq_item mmq_item_get(int it, const mmq_pargs & p, int & ncols);
__device__ __forceinline__ mmq_item mmq_item_of(int it, const mmq_pargs & p) {
    int t, half = 0, bn = 256;
    if (it < p.n_main) t = it;
    else t = p.n_main + ((it - p.n_main) >> 1), half = (it - p.n_main) & 1, bn = 128;
    mmq_item r;
    int      nt;
    if (p.raster) nt = t / p.mtiles, r.mt = t - nt * p.mtiles;
    else r.mt = t / p.ntn, nt = t - r.mt * p.ntn;
    r.n0 = nt * 256 + half * 128;
    r.bn = bn;
    return r;
}

template <bool MMID>
__device__ __forceinline__ mmq_item mmq_item_get(int it, const mmq_pargs & p, int & ncols) {
    if (MMID) {
        const int4 v = __ldg(p.items + it);
        mmq_item r;
        r.mt = v.x, r.n0 = v.y, r.bn = v.w, ncols = v.z;
        return r;
    }
    const mmq_item r = mmq_item_of(it, p);
    ncols = r.bn;
    return r;
}

template <int EPI, bool MMID>
__global__ void __launch_bounds__(mmq_pshape::THREADS, 1) k_mmq_p(const __grid_constant__ CUtensorMap mapA,
                                                const __grid_constant__ CUtensorMap mapB, const mmq_kargs a,
                                                const __grid_constant__ mmq_maps3 mapAx, const mmq_pargs p) {
    using SH = mmq_pshape;
    constexpr int NS = SH::NS, NB = SH::NB, NSC = SH::NSC;
    extern __shared__ uint8_t mmq_smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * NS + 2 * NB + 2 * NSC];
    __shared__ uint32_t tmem_slot;

    const int      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tiles = (smem_u32(mmq_smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * NS, bar_tfull = bar_empty + 8 * NS,
                   bar_tempty = bar_tfull + 8 * NB, bar_sfull = bar_tempty + 8 * NB, bar_sempty = bar_sfull + 8 * NSC;
    const uint32_t scales = tiles + NS * SH::STAGE;
    const int      nch = a.nseg * a.nch;                        // virtual chunks per item

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; s++) mbar_init(bar_full + 8 * s, 1), mbar_init(bar_empty + 8 * s, 1);
        for (int b = 0; b < NB; b++) mbar_init(bar_tfull + 8 * b, 1), mbar_init(bar_tempty + 8 * b, SH::EW);
        for (int b = 0; b < NSC; b++) mbar_init(bar_sfull + 8 * b, 1), mbar_init(bar_sempty + 8 * b, SH::EW);
        mbar_fence_init();
        prefetch_tmap(&mapA);
        prefetch_tmap(&mapB);
    }
    if (warp == 1) tmem_alloc<512>(smem_u32(&tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    pdl_launch_dependents();
    pdl_wait();                                                 // activations come from the quantizer kernel before us
    const int n_items = MMID ? __ldg(p.hdr) : p.n_items;        // (MUL_MAT_ID: written by the plan kernel two launches ago)

#ifdef B200_MMQ_TRACE
    long long tw[4] = {0, 0, 0, 0};
    const long long t_start = clock64();
#endif
    if (warp == 0) {
        // ------------------------------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            uint32_t st = 0, st_ph = 0, sc = 0, sc_ph = 0;      // stage / scale-slot cursors with their phase bits
            for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
                int            ncols_;
                const mmq_item I = mmq_item_get<MMID>(it, p, ncols_);
                const CUtensorMap * pA = &mapA;
                const float *       S  = a.S;
                int                 mpad = a.mpad, tile = I.mt;
                if (a.nmat > 1) {
                    int first = 0;
#pragma unroll
                    for (int i = 0; i < 3; i++)
                        if (i + 1 < a.nmat && I.mt >= a.tile0[i]) pA = &mapAx.m[i], S = a.Sx[i], mpad = a.mpadx[i], first = a.tile0[i];
                    tile -= first;
                }
                const int      m0 = (a.dbg & 2) ? 0 : tile * MMQ_BM;
                const bool     two = I.bn > 128 && !(a.dbg & 1);
                const uint32_t stage_tx = SH::A_BYTES + (two ? 256u : 128u) * MMQ_BK, scale_tx = (MMQ_BM + (uint32_t) I.bn) * 4;
                for (int c = 0; c < nch; c++) {
                    int ac, bc;
                    mmq_vchunk(a, c, ac, bc);
                    TW(0, mbar_wait(bar_sempty + 8 * sc, sc_ph ^ 1));  // this chunk's scale vectors, ahead of its tiles
                    const uint32_t sd = scales + sc * SH::SC_FLTS * 4;
                    mbar_expect_tx(bar_sfull + 8 * sc, scale_tx);
                    tma_load_1d(sd, S + (int64_t) ac * mpad + m0, MMQ_BM * 4, bar_sfull + 8 * sc);
                    tma_load_1d(sd + MMQ_BM * 4, a.Tt + (int64_t) bc * a.npad + I.n0, I.bn * 4, bar_sfull + 8 * sc);
                    if (++sc == NSC) sc = 0, sc_ph ^= 1;
#pragma unroll
                    for (int h = 0; h < MMQ_KBC; h++) {
                        TW(1, mbar_wait(bar_empty + 8 * st, st_ph ^ 1));
                        const uint32_t sa = tiles + st * SH::STAGE, sb = sa + SH::A_BYTES;
                        mbar_expect_tx(bar_full + 8 * st, stage_tx);
                        tma_load_2d(sa, pA, (MMQ_KBC * ac + h) * MMQ_BK, m0, bar_full + 8 * st);
                        tma_load_2d(sb, &mapB, (MMQ_KBC * bc + h) * MMQ_BK, I.n0, bar_full + 8 * st);
                        if (two) tma_load_2d(sb + 128 * MMQ_BK, &mapB, (MMQ_KBC * bc + h) * MMQ_BK, I.n0 + 128, bar_full + 8 * st);
                        if (++st == NS) st = 0, st_ph ^= 1;
                    }
                }
            }
#ifdef B200_MMQ_TRACE
            if (a.trace) {
                long long * o = a.trace + (size_t) blockIdx.x * 16;
                o[0] = clock64() - t_start, o[1] = tw[0], o[2] = tw[1];
            }
#endif
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            uint32_t st = 0, st_ph = 0, buf = 0, buf_ph = 0;
            for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
                const uint32_t idesc = MMID ? idesc_i8(MMQ_BM, __ldg(p.items + it).w)
                                            : it < p.n_main ? idesc_i8(MMQ_BM, 256) : idesc_i8(MMQ_BM, 128);
                for (int c = 0; c < nch; c++) {
                    // epilogue has drained this accumulator (EPI 3: and refilled it with the bias -- one more phase, the
                    // first one being the initial fill)
                    TW(0, mbar_wait(bar_tempty + 8 * buf, EPI == 3 ? buf_ph : buf_ph ^ 1));
                    tc_fence_after();
#pragma unroll
                    for (int h = 0; h < MMQ_KBC; h++) {
                        TW(1, mbar_wait(bar_full + 8 * st, st_ph));
                        tc_fence_after();
                        const uint32_t sa = tiles + st * SH::STAGE, sb = sa + SH::A_BYTES;
                        const uint64_t da = smem_desc_sw128(sa), db = smem_desc_sw128(sb);
#pragma unroll
                        for (int kk = 0; kk < MMQ_BK / 32; kk++)
                            mma_i8(tmem + buf * 256, da + 2 * kk, db + 2 * kk, idesc, EPI == 3 ? 1u : (uint32_t) ((h | kk) != 0));
                        mma_commit(bar_empty + 8 * st);
                        if (++st == NS) st = 0, st_ph ^= 1;
                    }
                    mma_commit(bar_tfull + 8 * buf);
                    if (++buf == NB) buf = 0, buf_ph ^= 1;
                }
            }
#ifdef B200_MMQ_TRACE
            if (a.trace) {
                long long * o = a.trace + (size_t) blockIdx.x * 16;
                o[4] = clock64() - t_start, o[5] = tw[0], o[6] = tw[1];
            }
#endif
        }
    } else {
        // ------------------------------------------------------------------------------------------ epilogue warps
        const int q = warp & 3, part = (warp - 2) >> 2;          // TMEM lane quarter (warp id % 4), column part 0..3
        const float * sring = reinterpret_cast<const float *>(mmq_smem_raw + (scales - smem_u32(mmq_smem_raw)));
        uint32_t sc = 0, sc_ph = 0, buf = 0, buf_ph = 0;
        if (EPI == 3) {
            // every accumulator column starts as the conversion bias; from here on a column is either freshly refilled by
            // the warp that read it, or belongs to the chunk in flight (columns an item does not use stay untouched)
#pragma unroll
            for (int b = 0; b < NB; b++) {
#pragma unroll
                for (int g = 0; g < 8; g++) tmem_st8_fill(tmem + ((uint32_t) (q * 32) << 16) + b * 256 + part * 64 + g * 8, MAGIC_BITS);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0)
                for (int b = 0; b < NB; b++) mbar_arrive(bar_tempty + 8 * b);
        }
        for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
            const int ncw = (MMID ? __ldg(p.items + it).w : it < p.n_main ? 256 : 128) >> 2;   // columns per warp (8 .. 64)
            const int col0 = part * ncw;
            __align__(8) float acc[64];
#pragma unroll
            for (int i = 0; i < 64; i++) acc[i] = 0.f;
            for (int c = 0; c < nch; c++) {
                TW(0, mbar_wait(bar_sfull + 8 * sc, sc_ph));
                const float *  sv  = sring + sc * SH::SC_FLTS;
                const float    sic = sv[q * 32 + lane];
                const float4 * tp  = reinterpret_cast<const float4 *>(sv + MMQ_BM + col0);
                TW(1, mbar_wait(bar_tfull + 8 * buf, buf_ph));
                tc_fence_after();
                const uint32_t tbase = tmem + ((uint32_t) (q * 32) << 16) + buf * 256 + col0;
                // Software pipeline over 8-column loads: the load of group g + 1 is in flight while group g is scaled and
                // accumulated (tcgen05.wait::ld waits for ALL outstanding loads, so the next load is issued right after the
                // wait that delivered the current one).
                auto release = [&] { if (lane == 0) mbar_arrive(bar_tempty + 8 * buf); };
                if (ncw == 64) epi_chunk8<EPI, 8>(tbase, tp, sic, acc, release);
                else if (!MMID || ncw == 32) epi_chunk8<EPI, 4>(tbase, tp, sic, acc, release);
                else if (ncw == 8)  epi_chunk8<EPI, 1>(tbase, tp, sic, acc, release);
                else if (ncw == 16) epi_chunk8<EPI, 2>(tbase, tp, sic, acc, release);
                else if (ncw == 24) epi_chunk8<EPI, 3>(tbase, tp, sic, acc, release);
                else if (ncw == 40) epi_chunk8<EPI, 5>(tbase, tp, sic, acc, release);
                else if (ncw == 48) epi_chunk8<EPI, 6>(tbase, tp, sic, acc, release);
                else                epi_chunk8<EPI, 7>(tbase, tp, sic, acc, release);
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_sempty + 8 * sc);
                if (++sc == NSC) sc = 0, sc_ph ^= 1;
                if (++buf == NB) buf = 0, buf_ph ^= 1;
            }
#ifdef B200_MMQ_TRACE
            const long long t_st = clock64();
#endif
            if (MMID) {
                // scatter: column j of the item belongs to pair sorted[n0 + j]; lanes look the dst rows up together
                int            ncols;
                const mmq_item I = mmq_item_get<true>(it, p, ncols);
                const int      row = (I.mt % p.mt_expert) * MMQ_BM + q * 32 + lane;
                int64_t        off[2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int j = col0 + 32 * h + lane;
                    off[h] = -1;
                    if (32 * h + lane < ncw && j < ncols) {
                        const int pr = __ldg(p.sorted + I.n0 + j);
                        if (pr >= 0) off[h] = (int64_t) (pr / p.n_used) * p.dst_token_stride + (int64_t) (pr % p.n_used) * p.dst_slot_stride;
                    }
                }
#pragma unroll
                for (int i = 0; i < 64; i++) {
                    if (i < ncw) {                               // warp-uniform
                        const int64_t o = __shfl_sync(0xffffffffu, off[i >> 5], i & 31);
                        if (o >= 0 && row < p.m_expert) a.dst[o + row] = acc[i];
                    }
                }
                continue;
            }
            // where the item's results go (worked out only now: it would cost ten live registers across the chunk loop)
            const mmq_item I = mmq_item_of(it, p);
            float *  dstp = a.dst;
            int64_t  dstride = a.dst_stride;
            int      mrows = a.m, tile = I.mt;
            if (a.nmat > 1) {
                int first = 0;
#pragma unroll
                for (int i = 0; i < 3; i++)
                    if (i + 1 < a.nmat && I.mt >= a.tile0[i]) dstp = a.dstx[i], dstride = a.dst_stridex[i], mrows = a.mx[i], first = a.tile0[i];
                tile -= first;
            }
            const int row = tile * MMQ_BM + q * 32 + lane;
            if (row < mrows) {
                const int jn = min(ncw, a.n - (I.n0 + col0));    // columns of this warp inside the matrix
                float *   d  = dstp + (int64_t) (I.n0 + col0) * dstride + row;
                if (p.world > 1) {
                    for (int r = 0; r < p.world; r++) {          // 128-byte warp stores, local or over NVLink
                        float * dr = reinterpret_cast<float *>(reinterpret_cast<char *>(d) + p.peer_delta[r]);
#pragma unroll
                        for (int i = 0; i < 64; i++) {
                            if (i < jn) *dr = acc[i];
                            dr += dstride;
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 64; i++) {               // lanes = consecutive rows: coalesced
                        if (i < jn) *d = acc[i];
                        d += dstride;
                    }
                }
            }
#ifdef B200_MMQ_TRACE
            tw[2] += clock64() - t_st;
#endif
        }
#ifdef B200_MMQ_TRACE
        if (a.trace && warp == 2 && lane == 0) {
            long long * o = a.trace + (size_t) blockIdx.x * 16;
            o[8] = clock64() - t_start - tw[2], o[9] = tw[0], o[10] = tw[1], o[11] = tw[2];
            unsigned long long g;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
            o[12] = (long long) g;
        }
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem);
    if (!MMID && p.world > 1 && threadIdx.x == 0) {
        // the barrier ordered every warp's result stores before this thread: one system-scope fence publishes them, then
        // the CTA counts itself on every rank; the last CTA of the grid tops the launch up to B200_ARRIVAL_UNIT
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        int add = 1;
        if (atomicAdd(p.local_count, 1) == (int) gridDim.x - 1) {
            atomicExch(p.local_count, 0);
            add += B200_ARRIVAL_UNIT - (int) gridDim.x;
        }
        for (int r = 0; r < p.world; r++)
            asm volatile("red.relaxed.sys.global.add.s32 [%0], %1;" ::"l"(p.peer_flag[r]), "r"(add) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
static encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = nullptr;
    if (!fn) {
        void *                           p = nullptr;
        cudaDriverEntryPointQueryResult  qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = (encode_tiled_fn) p;
    }
    return fn;
}

static int env_flag(const char * name, int dflt) { return env_knob(name, dflt); }

// row-major int8 matrix [rows][kp] -> tiles of box_rows x 128 bytes, 128-byte swizzle, zero fill outside
// A tensor map depends only on (base, rows, row length, box): planes and the activation workspace keep their addresses from
// call to call, so the encoded maps are kept in a small per-thread table (cuTensorMapEncodeTiled costs 1-2 us, 2-5 of them
// per batched matmul: on the path of an eager caller such as the plug-in, irrelevant under graph replay).
struct map_cache_entry {
    const void * base = nullptr;
    int64_t      rows = 0, kp = 0;
    int          box_rows = 0;
    CUtensorMap  map;
};
static int make_map(CUtensorMap * map, const void * base, int64_t rows, int64_t kp, int box_rows) {
    constexpr int SLOTS = 64;
    static thread_local map_cache_entry cache[SLOTS];
    const size_t h = ((uintptr_t) base >> 8) * 0x9E3779B97F4A7C15ull;
    map_cache_entry & e = cache[((h >> 40) ^ (size_t) rows ^ (size_t) box_rows) % SLOTS];
    if (e.base == base && e.rows == rows && e.kp == kp && e.box_rows == box_rows) {
        *map = e.map;
        return B200_OK;
    }
    encode_tiled_fn enc = get_encode_tiled();
    if (!enc) return B200_ERR_CUDA;
    const cuuint64_t dims[2]    = {(cuuint64_t) kp, (cuuint64_t) rows};
    const cuuint64_t strides[1] = {(cuuint64_t) kp};
    const cuuint32_t box[2]     = {(cuuint32_t) MMQ_BK, (cuuint32_t) box_rows};
    const cuuint32_t estr[2]    = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_last_cuda_error, sizeof(g_last_cuda_error), "cuTensorMapEncodeTiled failed: %d", (int) r);
        return B200_ERR_CUDA;
    }
    e.base = base, e.rows = rows, e.kp = kp, e.box_rows = box_rows, e.map = *map;
    return B200_OK;
}

int launch_plane_build(int type, const void * W, int64_t w_stride, void * plane, int64_t m, int64_t k,
                       cudaStream_t stream) {
    if (m == 0 || k == 0) return B200_OK;
    const int64_t kp = mmq_kpad(k), nwp = mmq_nplanes(type), ldp = nwp * kp;
    int8_t *      P  = (int8_t *) plane;
    float *       S  = (float *) ((uint8_t *) plane + align256((size_t) m * ldp));
    const int64_t total = (kp / 32) * m, mpad = mmq_mpad(m);
    const unsigned grid = (unsigned) ((total + 255) / 256);
    const uint8_t * wp = (const uint8_t *) W;
    switch (type) {
        case B200_TYPE_Q4_0:   k_plane_build<B200_TYPE_Q4_0, false><<<grid, 256, 0, stream>>>(wp, w_stride, P, S, m, k, kp, ldp, mpad); break;
        case B200_TYPE_Q8_0:   k_plane_build<B200_TYPE_Q8_0, false><<<grid, 256, 0, stream>>>(wp, w_stride, P, S, m, k, kp, ldp, mpad); break;
        case B200_TYPE_Q4_K:   k_plane_build<B200_TYPE_Q4_K, false><<<grid, 256, 0, stream>>>(wp, w_stride, P, S, m, k, kp, ldp, mpad); break;
        case B200_TYPE_Q6_K:   k_plane_build<B200_TYPE_Q6_K, false><<<grid, 256, 0, stream>>>(wp, w_stride, P, S, m, k, kp, ldp, mpad); break;
        case B200_TYPE_IQ4_XS: k_plane_build<B200_TYPE_IQ4_XS, false><<<grid, 256, 0, stream>>>(wp, w_stride, P, S, m, k, kp, ldp, mpad); break;
        default: return B200_ERR_TYPE;
    }
    note_launch();
    if (nwp == 2) {  // residual plane beside the first one (stream order: plane 0 is complete)
        k_plane_build<B200_TYPE_Q8_0, true><<<grid, 256, 0, stream>>>(wp, w_stride, P, S, m, k, kp, ldp, mpad);
        note_launch();
    }
    return check_cuda(cudaGetLastError(), "k_plane_build");
}

// workspace of the GEMM proper: quantized activations + their scales, sized for two activation planes whatever the type
// (the C ABI's size query has no type argument), + slack for overhanging tiles
size_t mmq_workspace_bytes(int64_t n, int64_t k) { return mmq_act_bytes(n, k, 2) + 2048; }

// epilogue flavour: B200_MMQ_EPI = 0 (I2FP), 1 (magic-number conversion), 2 (magic + packed fp32; default at chunk 256)
template <int BN, int EPI>
static int launch_mmq_bn_epi(const CUtensorMap & mapA, const CUtensorMap & mapB, const mmq_kargs & a, const mmq_maps3 & mapAx,
                             int mtiles, cudaStream_t stream) {
    auto kern = k_mmq<BN, EPI>;
    B200_SMEM_ATTR_ONCE(kern, mmq_shape<BN>::SMEM);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) mtiles, (unsigned) ((a.n + BN - 1) / BN), (unsigned) (a.ksplit > 1 ? a.ksplit : 1));
    cfg.blockDim         = dim3(mmq_shape<BN>::THREADS, 1, 1);
    cfg.dynamicSmemBytes = mmq_shape<BN>::SMEM;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    mmq_kargs a2 = a;
    a2.trace = nullptr;
    if (a2.ksplit < 1) a2.ksplit = 1;
#ifdef B200_MMQ_TRACE
    if (const char * tp = getenv("B200_MMQ_TRACE_PTR")) a2.trace = (long long *) strtoull(tp, nullptr, 0);
#endif
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, mapA, mapB, a2, mapAx));
    note_launch();
    return B200_OK;
}
template <int BN>
static int launch_mmq_bn(const CUtensorMap & mapA, const CUtensorMap & mapB, const mmq_kargs & a, const mmq_maps3 & mapAx,
                         int mtiles, cudaStream_t stream) {
    const int epi = MMQ_KC > 256 ? 0 : env_flag("B200_MMQ_EPI", MMQ_EPI_DEFAULT);
    switch (epi) {
        case 0: return launch_mmq_bn_epi<BN, 0>(mapA, mapB, a, mapAx, mtiles, stream);
        case 1: return launch_mmq_bn_epi<BN, 1>(mapA, mapB, a, mapAx, mtiles, stream);
        default: return launch_mmq_bn_epi<BN, 2>(mapA, mapB, a, mapAx, mtiles, stream);
    }
}

// persistent launch: items = 128 x 256 tiles (+ 128 x 128 halves for a short last round), one CTA per SM
static int launch_mmq_persistent(const CUtensorMap & mapA, const CUtensorMap & mapB128, const mmq_kargs & a,
                                 const mmq_maps3 & mapAx, int mtiles, int sm_count, cudaStream_t stream,
                                 const mmq_rowsplit * rs = nullptr) {
    // B200_MMQ_TMEM_BIAS=0: the epilogue that adds the conversion bias itself (EPI 2) instead of finding it in TMEM (EPI 3)
    const bool bias  = MMQ_PEPI_DEFAULT == 3 && env_flag("B200_MMQ_TMEM_BIAS", 1);
    auto kern3 = k_mmq_p<MMQ_PEPI_DEFAULT, false>;
    auto kern2 = k_mmq_p<MMQ_EPI_DEFAULT, false>;
    B200_SMEM_ATTR_ONCE(kern3, mmq_pshape::SMEM);
    B200_SMEM_ATTR_ONCE(kern2, mmq_pshape::SMEM);
    auto kern = bias ? kern3 : kern2;
    mmq_pargs p = {};
    p.world = 1;
    if (rs && rs->world > 1) {
        p.world = rs->world, p.rank = rs->rank, p.local_count = rs->local_count;
        for (int r = 0; r < rs->world; r++) p.peer_delta[r] = rs->peer_delta[r], p.peer_flag[r] = rs->peer_flag[r];
    }
    p.ntn = (a.n + 255) / 256;
    p.mtiles = mtiles;
    p.raster = env_flag("B200_MMQ_RASTER", 0);
    const int tiles = mtiles * p.ntn;
    const int G     = tiles < sm_count ? tiles : sm_count;
    const int rem   = tiles % G;
    const bool halves = tiles > G && rem > 0 && rem * 2 <= G && env_flag("B200_MMQ_TAIL_HALVES", 1);
    p.n_main  = halves ? tiles - rem : tiles;
    p.n_items = p.n_main + (halves ? 2 * rem : 0);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) G, 1, 1);
    cfg.blockDim         = dim3(mmq_pshape::THREADS, 1, 1);
    cfg.dynamicSmemBytes = mmq_pshape::SMEM;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    mmq_kargs a2 = a;
    a2.trace = nullptr, a2.ksplit = 1, a2.dbg = env_flag("B200_MMQ_DEBUG", 0);
#ifdef B200_MMQ_TRACE
    if (const char * tp = getenv("B200_MMQ_TRACE_PTR")) a2.trace = (long long *) strtoull(tp, nullptr, 0);
#endif
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, mapA, mapB128, a2, mapAx, p));
    note_launch();
    return B200_OK;
}

static int launch_quantize_x8(int nxp, const float * x, int64_t x_stride, int8_t * Q, float * Tt, int64_t n, int64_t k,
                              int64_t kp, int64_t npad, const mmq_zero & z, cudaStream_t stream, const mmq_gather & gt = {},
                              bool vec_ok = true) {
    const int64_t warps = n * (kp / MMQ_KC);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim  = dim3((unsigned) ((warps + 7) / 8), 1, 1);
    cfg.blockDim = dim3(256, 1, 1);
    cfg.stream   = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    const bool vec = vec_ok && (uintptr_t) x % 16 == 0 && x_stride % 4 == 0;
    if (nxp == 2) B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_quantize_x8<2>, x, x_stride, Q, Tt, n, k, kp, npad, z, vec, gt));
    else          B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_quantize_x8<1>, x, x_stride, Q, Tt, n, k, kp, npad, z, vec, gt));
    note_launch();
    return B200_OK;
}

// split-K decision: two halves of the (virtual) chunk range when one launch would leave more than half of the SMs without
// a tile (B200_MMQ_KSPLIT=1/2 forces)
static int mmq_ksplit(int64_t tiles, int64_t nvc, int sm_count, bool allowed) {
    const int e = env_flag("B200_MMQ_KSPLIT", 0);
    if (!allowed || nvc < 4) return 1;
    if (e) return e == 2 ? 2 : 1;
    return tiles * 2 <= sm_count ? 2 : 1;
}

// Prepared planes of ONE precision class (same number of weight and activation planes: K-quants + IQ4_XS | Q4_0 | Q8_0)
// and one k against the same activations: x is quantized ONCE and the m-tiles of all matrices share one grid (up to 4
// matrices per launch), so that small matrices (k/v projections: 8 m-tiles) do not leave the machine empty and every
// launch's fixed costs are paid once per group.
int launch_mmq_multi(int type, int nmat, const void * const * planes, const float * x, int64_t x_stride, float * const * dst,
                     const int64_t * dst_stride, const int64_t * m, int64_t n, int64_t k, void * workspace,
                     cudaStream_t stream, const mmq_rowsplit * rs) {
    if (nmat == 0 || n == 0) return B200_OK;
    if (rs && (nmat > 4 || MMQ_KC > 256)) return B200_ERR_UNSUPPORTED;   // row split: ONE persistent launch = one arrival
    const int sm_count = sm_count_current_device();
    if (!sm_count) return B200_ERR_CUDA;
    const int     nwp = mmq_nplanes(type), nxp = mmq_xplanes(type);
    const int64_t kp = mmq_kpad(k), nch = kp / MMQ_KC, npad = mmq_npad(n);
    int8_t * Q  = (int8_t *) workspace;
    float *  Tt = (float *) ((uint8_t *) workspace + align256((size_t) n * nxp * kp));
    // segments of the virtual k (chunk offsets into the weight planes / the activation planes)
    int nseg = 1, seg_a[3] = {0, 0, 0}, seg_b[3] = {0, 0, 0};
    if (nxp == 2) seg_a[nseg] = 0, seg_b[nseg] = (int) nch, nseg++;          // W0 . X1
    if (nwp == 2) seg_a[nseg] = (int) nch, seg_b[nseg] = 0, nseg++;          // W1 . X0
    int64_t total_mt = 0;
    bool    can_split = nmat <= 4;
    for (int i = 0; i < nmat; i++) {
        total_mt += (m[i] + MMQ_BM - 1) / MMQ_BM;
        can_split = can_split && m[i] % 4 == 0 && dst_stride[i] % 4 == 0 && (uintptr_t) dst[i] % 16 == 0;
    }
    // tile width: 256 columns (the tensor core takes ~120 clocks per instruction whether N is 128 or 256, see
    // profiles/r01/mmq_trace_v1.log) unless n is small; shapes with few tiles split the virtual k in two instead
    const int bn     = rs ? 256 : env_flag("B200_MMQ_BN", 0) > 0 ? env_flag("B200_MMQ_BN", 0) : (n <= 128 ? 128 : 256);
    const int ksplit = rs ? 1 : mmq_ksplit(total_mt * ((n + bn - 1) / bn), nseg * nch, sm_count, can_split);
    {
        mmq_zero z = {};
        if (ksplit > 1)
            for (int i = 0; i < nmat; i++) z.p[z.cnt] = dst[i], z.stride[z.cnt] = dst_stride[i], z.m[z.cnt] = m[i], z.cnt++;
        const int st = launch_quantize_x8(nxp, x, x_stride, Q, Tt, n, k, kp, npad, z, stream);
        if (st != B200_OK) return st;
    }
    const bool persist = rs || (ksplit == 1 && bn == 256 && env_flag("B200_MMQ_PERSIST", 1));
    CUtensorMap mapB;
    int st = make_map(&mapB, Q, n, nxp * kp, persist ? 128 : bn);
    if (st != B200_OK) return st;
    for (int i0 = 0; i0 < nmat; i0 += 4) {
        const int cnt = nmat - i0 < 4 ? nmat - i0 : 4;
        CUtensorMap mapA;
        mmq_maps3   mx = {};
        mmq_kargs   a  = {};
        a.Tt = Tt, a.n = (int) n, a.nch = (int) nch, a.npad = (int) npad, a.nmat = cnt, a.ksplit = ksplit;
        a.nseg = nseg;
        for (int s = 0; s < 3; s++) a.seg_a[s] = seg_a[s], a.seg_b[s] = seg_b[s];
        int tiles = 0;
        for (int j = 0; j < cnt; j++) {
            const int64_t   mj = m[i0 + j];
            const uint8_t * pb = (const uint8_t *) planes[i0 + j];
            const float *   Sj = (const float *) (pb + align256((size_t) mj * nwp * kp));
            st = make_map(j == 0 ? &mapA : &mx.m[j - 1], pb, mj, nwp * kp, MMQ_BM);
            if (st != B200_OK) return st;
            if (j == 0) {
                a.S = Sj, a.dst = dst[i0], a.dst_stride = dst_stride[i0], a.m = (int) mj, a.mpad = (int) mmq_mpad(mj);
            } else {
                a.tile0[j - 1] = tiles, a.Sx[j - 1] = Sj, a.dstx[j - 1] = dst[i0 + j], a.dst_stridex[j - 1] = dst_stride[i0 + j];
                a.mx[j - 1] = (int) mj, a.mpadx[j - 1] = (int) mmq_mpad(mj);
            }
            tiles += (int) ((mj + MMQ_BM - 1) / MMQ_BM);
        }
        if (tiles == 0) continue;
        st = persist ? launch_mmq_persistent(mapA, mapB, a, mx, tiles, sm_count, stream, rs)
           : bn == 256 ? launch_mmq_bn<256>(mapA, mapB, a, mx, tiles, stream)
                       : launch_mmq_bn<128>(mapA, mapB, a, mx, tiles, stream);
        if (st != B200_OK) return st;
    }
    return B200_OK;
}

int launch_mmq(int type, const void * plane, const float * x, int64_t x_stride, float * dst, int64_t dst_stride, int64_t m,
               int64_t n, int64_t k, void * workspace, cudaStream_t stream) {
    return launch_mmq_multi(type, 1, &plane, x, x_stride, &dst, &dst_stride, &m, n, k, workspace, stream, nullptr);
}

// ------------------------------------------------------------------------------------------------ MUL_MAT_ID on tensor cores
// workspace = [plan: hdr | sorted | items][Q: maxpos x (nxp * kpad) int8][t: (nxp * nchunks) x npad f32]
static inline size_t mmidq_plan_bytes(int64_t npairs, int64_t n_expert, int64_t mt) {
    return align256(16 + (size_t) mmidq_maxpos(npairs, n_expert) * 4) + align256((size_t) mmidq_max_items(npairs, n_expert, mt) * 16);
}
size_t mmid_mmq_workspace_bytes(int64_t m, int64_t k, int64_t n_expert, int64_t npairs) {
    if (m <= 0 || m % MMQ_BM != 0 || n_expert <= 0 || n_expert > MMIDQ_MAX_EXPERTS || npairs <= 0) return 0;
    const int64_t maxpos = mmidq_maxpos(npairs, n_expert), kp = mmq_kpad(k);
    return mmidq_plan_bytes(npairs, n_expert, m / MMQ_BM) + align256((size_t) maxpos * 2 * kp) +
           align256((size_t) 2 * (kp / MMQ_KC) * (((maxpos + 3) & ~(int64_t) 3) + 256) * 4) + 2048;
}
int launch_mmid_mmq(int type, const void * plane, const float * b, int64_t b_token_stride, int64_t b_slot_stride, int64_t ne11,
                    const int32_t * ids, int64_t ids_row_stride, float * dst, int64_t dst_token_stride, int64_t dst_slot_stride,
                    int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens, void * workspace,
                    cudaStream_t stream) {
    const int sm_count = sm_count_current_device();
    if (!sm_count) return B200_ERR_CUDA;
    const int64_t npairs = n_tokens * n_used, mt = m / MMQ_BM;
    if (m % MMQ_BM != 0 || n_expert > MMIDQ_MAX_EXPERTS || npairs > INT32_MAX / 2 || n_expert * m > INT32_MAX) return B200_ERR_UNSUPPORTED;
    const int     nwp = mmq_nplanes(type), nxp = mmq_xplanes(type);
    const int64_t kp = mmq_kpad(k), nch = kp / MMQ_KC, maxpos = mmidq_maxpos(npairs, n_expert),
                  npad = ((maxpos + 3) & ~(int64_t) 3) + 256;   // pitch of the scale rows: a multiple of 16 bytes (bulk copies)
    const int64_t max_items = mmidq_max_items(npairs, n_expert, mt), M = n_expert * m;
    uint8_t * w      = (uint8_t *) workspace;
    int32_t * hdr    = (int32_t *) w;
    int32_t * sorted = hdr + 4;
    int4 *    items  = (int4 *) (w + align256(16 + (size_t) maxpos * 4));
    int8_t *  Q      = (int8_t *) (w + mmidq_plan_bytes(npairs, n_expert, mt));
    float *   Tt     = (float *) ((uint8_t *) Q + align256((size_t) maxpos * nxp * kp));
    {   // 1. the plan: pairs sorted by expert, work items
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(1, 1, 1), cfg.blockDim = dim3(1024, 1, 1), cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr, cfg.numAttrs = 1;
        B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_mmid_sort, ids, ids_row_stride, (int) npairs, (int) n_used, (int) n_expert, (int) mt,
                                         (int) maxpos, hdr, sorted, items));
        note_launch();
    }
    {   // 2. activation rows gathered in sorted order while they are quantized
        mmq_gather gt = {};
        gt.sorted = sorted, gt.n_used = (int) n_used, gt.ne11 = (int) ne11;
        gt.x_token_stride = b_token_stride, gt.x_slot_stride = b_slot_stride;
        const bool vec_ok = b_token_stride % 4 == 0 && b_slot_stride % 4 == 0;
        const int st = launch_quantize_x8(nxp, b, 0, Q, Tt, maxpos, k, kp, npad, mmq_zero{}, stream, gt, vec_ok);
        if (st != B200_OK) return st;
    }
    // 3. the GEMM over the stacked expert planes
    CUtensorMap mapA, mapB;
    int st = make_map(&mapB, Q, maxpos, nxp * kp, 128);
    if (st != B200_OK) return st;
    st = make_map(&mapA, plane, M, nwp * kp, MMQ_BM);
    if (st != B200_OK) return st;
    mmq_kargs a = {};
    a.S = (const float *) ((const uint8_t *) plane + align256((size_t) M * nwp * kp));
    a.Tt = Tt, a.dst = dst, a.dst_stride = m, a.m = (int) M, a.n = (int) maxpos, a.nch = (int) nch, a.npad = (int) npad;
    a.mpad = (int) mmq_mpad(M), a.nmat = 1, a.ksplit = 1, a.nseg = 1;
    if (nxp == 2) a.seg_a[a.nseg] = 0, a.seg_b[a.nseg] = (int) nch, a.nseg++;
    if (nwp == 2) a.seg_a[a.nseg] = (int) nch, a.seg_b[a.nseg] = 0, a.nseg++;
    mmq_pargs p = {};
    p.items = items, p.hdr = hdr, p.sorted = sorted, p.n_used = (int) n_used, p.mt_expert = (int) mt, p.m_expert = (int) m;
    p.dst_token_stride = dst_token_stride, p.dst_slot_stride = dst_slot_stride;
    const bool bias  = MMQ_PEPI_DEFAULT == 3 && env_flag("B200_MMQ_TMEM_BIAS", 1);
    auto kern3 = k_mmq_p<MMQ_PEPI_DEFAULT, true>;
    auto kern2 = k_mmq_p<MMQ_EPI_DEFAULT, true>;
    B200_SMEM_ATTR_ONCE(kern3, mmq_pshape::SMEM);
    B200_SMEM_ATTR_ONCE(kern2, mmq_pshape::SMEM);
    auto kern = bias ? kern3 : kern2;
    const int64_t grid = max_items < sm_count ? max_items : sm_count;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) grid, 1, 1), cfg.blockDim = dim3(mmq_pshape::THREADS, 1, 1);
    cfg.dynamicSmemBytes = mmq_pshape::SMEM, cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    static const mmq_maps3 none = {};
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, mapA, mapB, a, none, p));
    note_launch();
    return B200_OK;
}

}  // namespace b200
