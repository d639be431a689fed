// noalloc.cu -- b200_mul_mat_noalloc: the batch 1..8 MUL_MAT (activation quantizer + k_gemv) for callers that cannot
// allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  It is b200_mul_mat's GEMV branch with
// one difference: the q8_1 activation scratch that b200_mul_mat takes from the caller's workspace lives in a file-scope
// __device__ array of this translation unit, so the call allocates nothing and takes no workspace argument.
//
//  * The array is B200_NOALLOC_SCRATCH_BYTES (384 KiB) and every GPU that loads the library holds it until exit.  That
//    is enough for the largest problem the GEMV path computes at all: 8 columns of k = 32768
//    (8 x q8_soa_layout(32768).bytes = 360448).  Larger problems are rejected with B200_ERR_TOO_LARGE.
//  * In host code the array's NAME is a host-side stand-in, not a device address.  The device address is taken once per
//    device, with that device current, by cudaGetSymbolAddress (return code checked) and that pointer is what the
//    kernels receive.
//  * Every call on a device shares the array and it keeps the last call's contents.  Nothing in it has to start at
//    zero: the quantizer kernel overwrites every byte the GEMV of the same call reads.  Calls that OVERLAP on one GPU
//    (two streams, two threads) race on it -- serialize them, or use b200_mul_mat with a workspace per stream.
//  * Calls no cudaMalloc / cudaFree / allocating library routine: only cudaGetDevice, cudaGetSymbolAddress (first call per
//    device) and kernel launches.
#include "actquant.cuh"
#include "kernels.h"

#include <cstdio>

namespace b200 {

static __device__ __align__(256) uint8_t g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

// device address of g_noalloc_scratch on the CURRENT device; nullptr + status on failure
static int noalloc_scratch_ptr(uint8_t ** out) {
    static std::atomic<uint8_t *> ptrs[MAX_DEVICES];
    const int dev = current_device();
    if (dev < 0 || dev >= MAX_DEVICES) return B200_ERR_CUDA;
    uint8_t * p = ptrs[dev].load(std::memory_order_acquire);
    if (!p) {
        void * sym = nullptr;
        B200_CUDA_TRY(cudaGetSymbolAddress(&sym, g_noalloc_scratch));   // this device's copy: the device is current
        if (!sym) return B200_ERR_CUDA;
        p = (uint8_t *) sym;
        ptrs[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_noalloc_scratch_size(void) { return B200_NOALLOC_SCRATCH_BYTES; }

int b200_mul_mat_noalloc(int type, const void * W, int64_t w_row_stride, const float * x, int64_t x_row_stride, float * dst,
                         int64_t dst_row_stride, int64_t m, int64_t n, int64_t k, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (!b200_type_supported(type)) return B200_ERR_TYPE;
    const int64_t blck = b200_block_size(type);
    if (m < 0 || n < 0 || k <= 0 || k % blck != 0 || x_row_stride < k || dst_row_stride < m) return B200_ERR_SHAPE;
    if (m > INT32_MAX || k > INT32_MAX) return B200_ERR_SHAPE;
    const int64_t rs = (int64_t) b200_row_size(type, k);
    if (w_row_stride == 0) w_row_stride = rs;
    if (w_row_stride < rs) return B200_ERR_SHAPE;
    const int al = type == B200_TYPE_Q4_K ? 16 : type == B200_TYPE_IQ4_XS ? 8 : 2;
    if ((uintptr_t) W % al != 0 || w_row_stride % al != 0) return B200_ERR_ALIGNMENT;
    if ((uintptr_t) x % 4 != 0 || (uintptr_t) dst % 4 != 0) return B200_ERR_ALIGNMENT;
    if (m == 0 || n == 0) return B200_OK;
    // the limits of this entry point: the GEMV path (n <= 8, one CTA spans a row) and the built-in scratch
    const int       maxc = gemv_max_cols(type, k);
    const q8_layout L    = q8_soa_layout(k);
    if (n > 8 || maxc < 1 || (size_t) n * L.bytes > (size_t) B200_NOALLOC_SCRATCH_BYTES) {
        snprintf(g_last_cuda_error, sizeof(g_last_cuda_error),
                 "b200_mul_mat_noalloc: n = %lld, k = %lld needs %zu bytes of q8 scratch on the batch <= 8 path; the built-in "
                 "scratch holds %d bytes, n <= 8, k <= 32768 (use b200_mul_mat with a caller-owned workspace)",
                 (long long) n, (long long) k, (size_t) n * L.bytes, (int) B200_NOALLOC_SCRATCH_BYTES);
        return B200_ERR_TOO_LARGE;
    }
    uint8_t * scratch = nullptr;
    int st = noalloc_scratch_ptr(&scratch);
    if (st != B200_OK) return st;
    st = launch_quantize_q8_soa(x, x_row_stride, scratch, k, n, type == B200_TYPE_IQ4_XS, stream);
    if (st != B200_OK) return st;
    gemv_args g = {};
    g.type = type, g.W = (const uint8_t *) W, g.w_row_stride = w_row_stride;
    g.m = m, g.k = k, g.dst_col_stride = dst_row_stride, g.x_col_stride = x_row_stride;
    g.world = 1;
    for (int64_t c0 = 0; c0 < n; c0 += maxc) {
        g.ncols = (int) (n - c0 < maxc ? n - c0 : maxc);
        g.xq    = scratch + (size_t) c0 * L.bytes;
        g.dst   = dst + c0 * dst_row_stride;
        st = launch_gemv(g, stream);
        if (st != B200_OK) return st;
    }
    return B200_OK;
}

}  // extern "C"
