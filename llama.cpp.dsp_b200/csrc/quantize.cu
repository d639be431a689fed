// quantize.cu -- standalone activation quantizer (ggml block_q8_1 output, bit-exact with the CPU reference), the SoA
// quantizer that feeds the batched matmul kernels, and the bit-exact dequantizers.
//
// Replaces (reference): ggml/src/ggml-cuda/quantize.cu:4-38,129-141 (quantize_q8_1) and
// ggml/src/ggml-cuda/convert.cu:6-25,70-96,189-218,254-278,440-455 (dequantize_block*).
// Oracles: ggml/src/ggml-quants.c:220-253 (quantize_row_q8_1_ref), :255-273, 349-363, 1280-1302, 1690-1722,
// 2454-2476 (dequantize_row_*).
#include "actquant.cuh"
#include "dequant.cuh"
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
// row-split waits that gave up (per device; read and cleared by rowsplit_timeouts_fetch)
__device__ int g_rowsplit_timeouts = 0;

template <bool INTERLEAVE>
__global__ void __launch_bounds__(256) k_quantize_q8_soa(const float * __restrict__ x, int64_t x_stride,
                                                         uint8_t * __restrict__ y, int64_t y_pitch, int64_t k, bool vec,
                                                         int32_t * wait_sync, int wait_target) {
    pdl_launch_dependents();   // first: the consuming GEMV may become resident and prefetch its weights right away
    pdl_wait();
    if (wait_sync) {
        // row-split: x is only whole once `wait_target` ranks' slices have landed (their GEMVs bump wait_sync[0] over
        // NVLink).  The producers are kernels on OTHER GPUs, so this cannot starve them; the spin is bounded anyway
        // (~2 s) and counts a time-out in g_rowsplit_timeouts (b200_rowsplit_timeout_count) instead of hanging the
        // stream.  (The mark used to go into wait_sync[1], which is the block's local CTA counter: one time-out broke
        // every later use of the block.)  The kernel then quantizes whatever x holds: the caller must check the count.
        if (threadIdx.x == 0) {
            const long long t0 = clock64();
            int v;
            for (;;) {
                asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(wait_sync) : "memory");
                if (v >= wait_target) break;
                if (clock64() - t0 > (1ll << 32)) {
                    atomicAdd(&g_rowsplit_timeouts, 1);
                    break;
                }
            }
        }
        __syncthreads();
    }
    const q8_layout L = q8_soa_layout(k);
    quantize_row_q8_soa<INTERLEAVE>(x + blockIdx.y * x_stride, k, y + (int64_t) blockIdx.y * y_pitch, L,
                                    blockIdx.x * 8 + (threadIdx.x >> 5), gridDim.x * 8, threadIdx.x & 31, vec);
}

int launch_quantize_q8_soa_strided(const float * x, int64_t x_stride, void * y, int64_t y_pitch, int64_t k,
                                   int64_t nrows, bool interleave, cudaStream_t stream, int32_t * wait_sync,
                                   int wait_target) {
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
        if (interleave) B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_quantize_q8_soa<true>, xp, x_stride, yp, y_pitch, k, vec, wait_sync, wait_target));
        else            B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_quantize_q8_soa<false>, xp, x_stride, yp, y_pitch, k, vec, wait_sync, wait_target));
        note_launch();
        done += rows;
    }
    return B200_OK;
}

int launch_quantize_q8_soa(const float * x, int64_t x_stride, void * y, int64_t k, int64_t nrows, bool interleave,
                           cudaStream_t stream, int32_t * wait_sync, int wait_target) {
    return launch_quantize_q8_soa_strided(x, x_stride, y, (int64_t) q8_soa_layout(k).bytes, k, nrows, interleave, stream,
                                          wait_sync, wait_target);
}

// ---------------------------------------------------------------------------------------------- dequantizers
// One thread per (row, 32-element group); every float op rounds separately (__fmul_rn/__fsub_rn), in the reference's
// operand order, so results are bit-identical to dequantize_row_*.
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
    dequant_group32<TYPE>(wr, gi, o);
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

int rowsplit_timeouts_fetch(int * count, cudaStream_t stream) {
    int zero = 0;
    B200_CUDA_TRY(cudaMemcpyFromSymbolAsync(count, g_rowsplit_timeouts, sizeof(int), 0, cudaMemcpyDeviceToHost, stream));
    B200_CUDA_TRY(cudaStreamSynchronize(stream));
    if (*count) B200_CUDA_TRY(cudaMemcpyToSymbolAsync(g_rowsplit_timeouts, &zero, sizeof(int), 0, cudaMemcpyHostToDevice, stream));
    return B200_OK;
}

}  // namespace b200
