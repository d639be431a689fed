// tmem_read_bench.cu -- micro-benchmark: how fast can epilogue warps read TMEM (tcgen05.ld 32x32b) on sm_100a?
// Decides the prefill design: the tensor-core path reads its int32 accumulators once per scale chunk, so the chunk size
// at which the read-out hides behind the MMAs follows from this number.  Build: see tools/ubench/Makefile.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

template <int X> __device__ __forceinline__ uint32_t ld_cols(uint32_t taddr);
template <> __device__ __forceinline__ uint32_t ld_cols<16>(uint32_t taddr) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) x ^= r[i];
    return x;
}
template <> __device__ __forceinline__ uint32_t ld_cols<32>(uint32_t taddr) {
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                   "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 32; i++) x ^= r[i];
    return x;
}
// two x32 loads in flight before one wait
template <> __device__ __forceinline__ uint32_t ld_cols<64>(uint32_t taddr) {
    uint32_t r[64];
#pragma unroll
    for (int h = 0; h < 2; h++)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                     "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(r[32 * h + 0]), "=r"(r[32 * h + 1]), "=r"(r[32 * h + 2]), "=r"(r[32 * h + 3]), "=r"(r[32 * h + 4]),
                       "=r"(r[32 * h + 5]), "=r"(r[32 * h + 6]), "=r"(r[32 * h + 7]), "=r"(r[32 * h + 8]), "=r"(r[32 * h + 9]),
                       "=r"(r[32 * h + 10]), "=r"(r[32 * h + 11]), "=r"(r[32 * h + 12]), "=r"(r[32 * h + 13]),
                       "=r"(r[32 * h + 14]), "=r"(r[32 * h + 15]), "=r"(r[32 * h + 16]), "=r"(r[32 * h + 17]),
                       "=r"(r[32 * h + 18]), "=r"(r[32 * h + 19]), "=r"(r[32 * h + 20]), "=r"(r[32 * h + 21]),
                       "=r"(r[32 * h + 22]), "=r"(r[32 * h + 23]), "=r"(r[32 * h + 24]), "=r"(r[32 * h + 25]),
                       "=r"(r[32 * h + 26]), "=r"(r[32 * h + 27]), "=r"(r[32 * h + 28]), "=r"(r[32 * h + 29]),
                       "=r"(r[32 * h + 30]), "=r"(r[32 * h + 31])
                     : "r"(taddr + 32 * h) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 64; i++) x ^= r[i];
    return x;
}

// EW epilogue warps (warp id % 4 = TMEM lane quarter) sweep the 512 columns `iters` times, X columns per wait
template <int X>
__global__ void __launch_bounds__(18 * 32, 1) k_tmem_read(int ew, int iters, long long * clocks, uint32_t * sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    if (warp >= 2 && warp < 2 + ew) {
        const int q = warp & 3, part = (warp - 2) >> 2, nparts = ew / 4 > 0 ? ew / 4 : 1;
        const int cols = 512 / nparts, c0 = part * cols;            // this warp's column range of its lane quarter
        uint32_t x = 0;
        const long long t0 = clock64();
        for (int it = 0; it < iters; it++)
            for (int c = 0; c < cols; c += X) x ^= ld_cols<X>(tmem + ((uint32_t) (q * 32) << 16) + c0 + c);
        const long long t1 = clock64();
        if (lane == 0) clocks[blockIdx.x * 16 + (warp - 2)] = t1 - t0;
        if (x == 0x12345678u) sink[0] = x;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <int X> static void run(int ew, int iters, long long * d_clk, uint32_t * d_sink) {
    cudaMemset(d_clk, 0, 148 * 16 * 8);
    k_tmem_read<X><<<148, 18 * 32>>>(ew, iters, d_clk, d_sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("X=%d ew=%d: %s\n", X, ew, cudaGetErrorString(e)); return; }
    static long long h[148 * 16];
    cudaMemcpy(h, d_clk, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0; double sum = 0; int cnt = 0;
    for (int i = 0; i < 148 * 16; i++) if (h[i]) { mx = h[i] > mx ? h[i] : mx; sum += h[i]; cnt++; }
    // per SM: every warp quarter-group reads 32 lanes x its columns x 4 B per sweep; all EW warps together read
    // min(ew,4) quarters x 512 columns -> bytes per sweep = min(ew,4)/4 * 128 * 512 * 4
    const double bytes = (double) iters * (ew < 4 ? ew : 4) * 32 * 512 * 4;
    printf("cols/wait %3d  epilogue warps %2d : %8.0f clk mean %8lld max for %5.1f KB x %d sweeps -> %6.1f B/clk/SM (mean)  %6.1f (max-paced)\n",
           X, ew, sum / cnt, mx, bytes / iters / 1024, iters, bytes / (sum / cnt), bytes / mx);
}

int main() {
    long long * d_clk; uint32_t * d_sink;
    cudaMalloc(&d_clk, 148 * 16 * 8); cudaMalloc(&d_sink, 4);
    for (int ew : {1, 4, 8, 16}) { run<16>(ew, 64, d_clk, d_sink); run<32>(ew, 64, d_clk, d_sink); run<64>(ew, 64, d_clk, d_sink); }
    return 0;
}
