// quantize.cu -- standalone activation quantizer (ggml block_q8_1 output, bit-exact with the CPU reference), the SoA
// quantizer that feeds the batched matmul kernels, and the bit-exact dequantizers.
//
// Replaces (reference): ggml/src/ggml-cuda/quantize.cu:4-38,129-141 (quantize_q8_1) and
// ggml/src/ggml-cuda/convert.cu:6-25,70-96,189-218,254-278,440-455 (dequantize_block*).
// Oracles: ggml/src/ggml-quants.c:220-253 (quantize_row_q8_1_ref), :255-273, 349-363, 1280-1302, 1690-1722,
// 2454-2476 (dequantize_row_*).
#include "actquant.cuh"
#include "kernels.h"

namespace b200 {

// ---------------------------------------------------------------------------------------------- q8_1, ggml layout
// grid (ceil(k/32 / 16), nrows), block 128: each group of 8 lanes quantizes one 36-byte block.
__global__ void __launch_bounds__(128) k_quantize_q8_1_ggml(const float * __restrict__ x, int64_t x_stride,
                                                            uint8_t * __restrict__ y, int k, bool vec) {
    const int     nb  = k / 32;
    const int64_t row = blockIdx.y;
    const int     b   = blockIdx.x * 16 + (threadIdx.x >> 3);
    const int     sub = threadIdx.x & 7;
    const bool    active = b < nb;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active) v = load_f4(x + row * x_stride + (int64_t) b * 32 + 4 * sub, vec);
    const q8_block_result r = quantize_block32(v);
    if (!active) return;
    uint8_t * blk = y + (row * nb + b) * 36;
    // 36-byte blocks are 4-byte aligned: {half d, half s} then qs
    *reinterpret_cast<uint32_t *>(blk + 4 + 4 * sub) = r.q4;
    if (sub == 0) {
        const __half hd = __float2half_rn(r.d);
        const __half hs = __float2half_rn(__fmul_rn((float) r.isum, r.d));
        *reinterpret_cast<uint32_t *>(blk) = (uint32_t) __half_as_ushort(hd) | ((uint32_t) __half_as_ushort(hs) << 16);
    }
}

int launch_quantize_q8_1_ggml(const float * x, int64_t x_stride, void * y, int64_t k, int64_t nrows,
                              cudaStream_t stream) {
    if (k == 0 || nrows == 0) return B200_OK;
    const bool vec = ((uintptr_t) x % 16 == 0) && (x_stride % 4 == 0);
    int64_t done = 0;
    while (done < nrows) {  // grid.y limit
        const int64_t rows = nrows - done < 65535 ? nrows - done : 65535;
        dim3 grid((unsigned) ((k / 32 + 15) / 16), (unsigned) rows);
        k_quantize_q8_1_ggml<<<grid, 128, 0, stream>>>(x + done * x_stride, x_stride,
                                                       (uint8_t *) y + done * (k / 32) * 36, (int) k, vec);
        note_launch();
        done += rows;
    }
    return check_cuda(cudaGetLastError(), "k_quantize_q8_1_ggml");
}

// ---------------------------------------------------------------------------------------------- q8_1, SoA layout
// grid (ceil(k/32 / 32), nrows), 8 warps: every 8-lane group quantizes one block; output rows are L.bytes apart.
// Launched with programmatic stream serialization so the GEMV that consumes it can prefetch weights meanwhile.
template <bool INTERLEAVE>
__global__ void __launch_bounds__(256) k_quantize_q8_soa(const float * __restrict__ x, int64_t x_stride,
                                                         uint8_t * __restrict__ y, int64_t y_pitch, int64_t k, bool vec) {
    pdl_launch_dependents();   // first: the consuming GEMV may become resident and prefetch its weights right away
    pdl_wait();
    const q8_layout L = q8_soa_layout(k);
    quantize_row_q8_soa<INTERLEAVE>(x + blockIdx.y * x_stride, k, y + (int64_t) blockIdx.y * y_pitch, L,
                                    blockIdx.x * 8 + (threadIdx.x >> 5), gridDim.x * 8, threadIdx.x & 31, vec);
}

int launch_quantize_q8_soa_strided(const float * x, int64_t x_stride, void * y, int64_t y_pitch, int64_t k,
                                   int64_t nrows, bool interleave, cudaStream_t stream) {
    if (k == 0 || nrows == 0) return B200_OK;
    const bool vec = ((uintptr_t) x % 16 == 0) && (x_stride % 4 == 0);
    int64_t done = 0;
    while (done < nrows) {
        const int64_t rows = nrows - done < 65535 ? nrows - done : 65535;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim  = dim3((unsigned) ((k / 32 + 31) / 32), (unsigned) rows, 1);
        cfg.blockDim = dim3(256, 1, 1);
        cfg.stream   = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr, cfg.numAttrs = 1;
        const float * xp = x + done * x_stride;
        uint8_t *     yp = (uint8_t *) y + done * y_pitch;
        if (interleave) B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_quantize_q8_soa<true>, xp, x_stride, yp, y_pitch, k, vec));
        else            B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_quantize_q8_soa<false>, xp, x_stride, yp, y_pitch, k, vec));
        note_launch();
        done += rows;
    }
    return B200_OK;
}

int launch_quantize_q8_soa(const float * x, int64_t x_stride, void * y, int64_t k, int64_t nrows, bool interleave,
                           cudaStream_t stream) {
    return launch_quantize_q8_soa_strided(x, x_stride, y, (int64_t) q8_soa_layout(k).bytes, k, nrows, interleave, stream);
}

// ---------------------------------------------------------------------------------------------- dequantizers
// One thread per (row, 32-element group); every float op rounds separately (__fmul_rn/__fsub_rn), in the reference's
// operand order, so results are bit-identical to dequantize_row_*.
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

__device__ const int8_t c_iq4nl[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};

template <int TYPE>
__global__ void __launch_bounds__(256) k_dequantize(const uint8_t * __restrict__ w, int64_t w_stride,
                                                    float * __restrict__ y, int64_t k, int64_t nrows) {
    const int64_t groups_per_row = k / 32;
    const int64_t g = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= groups_per_row * nrows) return;
    const int64_t row = g / groups_per_row;
    const int64_t gi  = g % groups_per_row;          // 32-element group within the row
    const uint8_t * wr = w + row * w_stride;
    float *         o  = y + row * k + gi * 32;
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

int launch_dequantize_f32(int type, const void * w, int64_t w_stride, float * y, int64_t k, int64_t nrows,
                          cudaStream_t stream) {
    if (k == 0 || nrows == 0) return B200_OK;
    const int64_t  total = (k / 32) * nrows;
    const unsigned grid  = (unsigned) ((total + 255) / 256);
    const uint8_t * wp = (const uint8_t *) w;
    switch (type) {
        case B200_TYPE_Q4_0:   k_dequantize<B200_TYPE_Q4_0><<<grid, 256, 0, stream>>>(wp, w_stride, y, k, nrows); break;
        case B200_TYPE_Q8_0:   k_dequantize<B200_TYPE_Q8_0><<<grid, 256, 0, stream>>>(wp, w_stride, y, k, nrows); break;
        case B200_TYPE_Q4_K:   k_dequantize<B200_TYPE_Q4_K><<<grid, 256, 0, stream>>>(wp, w_stride, y, k, nrows); break;
        case B200_TYPE_Q6_K:   k_dequantize<B200_TYPE_Q6_K><<<grid, 256, 0, stream>>>(wp, w_stride, y, k, nrows); break;
        case B200_TYPE_IQ4_XS: k_dequantize<B200_TYPE_IQ4_XS><<<grid, 256, 0, stream>>>(wp, w_stride, y, k, nrows); break;
        default: return B200_ERR_TYPE;
    }
    note_launch();
    return check_cuda(cudaGetLastError(), "k_dequantize");
}

}  // namespace b200
