// api.cu -- the extern "C" boundary declared in include/b200_qmm.h: argument validation, kernel selection, workspace
// carving.  Host-only logic; every device kernel lives in quantize.cu / mmvq.cu / mmq.cu.
#include "actquant.cuh"
#include "kernels.h"

#include <cuda.h>

#include <atomic>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

thread_local char g_last_cuda_error[256] = "";
static std::atomic<int64_t> g_launches{0};

void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int check_cuda(cudaError_t e, const char * what) {
    if (e == cudaSuccess) return B200_OK;
    snprintf(g_last_cuda_error, sizeof(g_last_cuda_error), "%s: %s (%s)", what, cudaGetErrorName(e),
             cudaGetErrorString(e));
    return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? B200_ERR_NO_DEVICE : B200_ERR_CUDA;
}

int current_device() {
    int dev = -1;
    return cudaGetDevice(&dev) == cudaSuccess ? dev : -1;
}
int sm_count_current_device() {
    static std::atomic<int> cache[MAX_DEVICES];
    const int dev = current_device();
    if (dev < 0 || dev >= MAX_DEVICES) return 0;
    int n = cache[dev].load(std::memory_order_relaxed);
    if (!n) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

// Environment knobs are read once per process (a getenv per knob per launch is ~1 us of host time on the eager path);
// b200_reload_env() forgets the cached values (tests and sweep tools flip knobs inside one process).
struct knob_entry {
    const char * name;
    int          value;
};
static knob_entry           g_knobs[64];
static std::atomic<int>     g_nknobs{0};
static std::mutex           g_knob_mtx;
int env_knob(const char * name, int dflt) {
    const int n = g_nknobs.load(std::memory_order_acquire);
    for (int i = 0; i < n; i++)
        if (g_knobs[i].name == name || strcmp(g_knobs[i].name, name) == 0) return g_knobs[i].value;
    const char * e = getenv(name);
    const int    v = e && *e ? atoi(e) : dflt;
    std::lock_guard<std::mutex> lock(g_knob_mtx);
    const int m = g_nknobs.load(std::memory_order_relaxed);
    if (m < 64) {
        g_knobs[m].name = name, g_knobs[m].value = v;
        g_nknobs.store(m + 1, std::memory_order_release);
    }
    return v;
}
static void env_forget() {
    std::lock_guard<std::mutex> lock(g_knob_mtx);
    g_nknobs.store(0, std::memory_order_release);
}

static int blck_of(int type) {
    switch (type) {
        case B200_TYPE_Q4_0: case B200_TYPE_Q8_0: case B200_TYPE_Q8_1: return 32;
        case B200_TYPE_Q4_K: case B200_TYPE_Q6_K: case B200_TYPE_IQ4_XS: return 256;
    }
    return 0;
}
static size_t tsize_of(int type) {
    switch (type) {
        case B200_TYPE_Q4_0: return 18;
        case B200_TYPE_Q8_0: return 34;
        case B200_TYPE_Q8_1: return 36;
        case B200_TYPE_Q4_K: return 144;
        case B200_TYPE_Q6_K: return 210;
        case B200_TYPE_IQ4_XS: return 136;
    }
    return 0;
}
static int align_of(int type) { return type == B200_TYPE_Q4_K ? 16 : type == B200_TYPE_IQ4_XS ? 8 : 2; }
static bool weight_type(int type) { return type != B200_TYPE_Q8_1 && blck_of(type) != 0; }

static int check_weights(int type, const void * W, int64_t & stride, int64_t k) {
    if (!weight_type(type)) return B200_ERR_TYPE;
    if (k < 0 || k % blck_of(type) != 0) return B200_ERR_SHAPE;
    const int64_t rs = (int64_t) (k / blck_of(type)) * (int64_t) tsize_of(type);
    if (stride == 0) stride = rs;
    if (stride < rs) return B200_ERR_SHAPE;
    const int al = align_of(type);
    if ((uintptr_t) W % al != 0 || stride % al != 0) return B200_ERR_ALIGNMENT;
    return B200_OK;
}

}  // namespace b200

using namespace b200;

struct rowsplit_signal_args {
    int32_t * flag[16];
    int       world;
};
// the arrival of a rank that owns no rows of this matrix (or of the world == 1 degenerate case)
static __global__ void k_rowsplit_signal(const rowsplit_signal_args a) {
    if (threadIdx.x == 0) {
        __threadfence_system();
        for (int r = 0; r < a.world; r++) atomicAdd_system(a.flag[r], B200_ARRIVAL_UNIT);
    }
}

// results -> pinned host memory with coalesced 16-byte stores over PCIe: a kernel boundary (~2 us under programmatic
// dependent launch) instead of a copy node (~10 us each in a stream of 129 per token)
static __global__ void __launch_bounds__(256) k_copy_out(const float * __restrict__ src, float * __restrict__ dst, size_t count,
                                                         bool vec) {
    pdl_launch_dependents();
    pdl_wait();
    const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t) gridDim.x * blockDim.x;
    if (vec) {
        for (size_t i = tid; i < count / 4; i += nthr) reinterpret_cast<float4 *>(dst)[i] = reinterpret_cast<const float4 *>(src)[i];
        for (size_t i = count / 4 * 4 + tid; i < count; i += nthr) dst[i] = src[i];
    } else {
        for (size_t i = tid; i < count; i += nthr) dst[i] = src[i];
    }
}
// device results -> host: the copy kernel when dst_host is pinned (UVA-mapped), else cudaMemcpyAsync
static int results_to_host(float * dst_host, const float * src_dev, size_t count, cudaStream_t stream) {
    if (!count) return B200_OK;
    cudaPointerAttributes at;
    float * alias = nullptr;
    if (env_knob("B200_HOST_ZEROCOPY", 1) != 0) {
        if (cudaPointerGetAttributes(&at, dst_host) == cudaSuccess) {
            if (at.type == cudaMemoryTypeHost && at.devicePointer) alias = (float *) at.devicePointer;
        } else {
            cudaGetLastError();
        }
    }
    if (!alias) {
        B200_CUDA_TRY(cudaMemcpyAsync(dst_host, src_dev, count * 4, cudaMemcpyDeviceToHost, stream));
        return B200_OK;
    }
    const bool vec = (uintptr_t) alias % 16 == 0 && (uintptr_t) src_dev % 16 == 0;
    size_t blocks = (count / 4 + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 296) blocks = 296;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim  = dim3((unsigned) blocks, 1, 1);
    cfg.blockDim = dim3(256, 1, 1);
    cfg.stream   = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_copy_out, src_dev, alias, count, vec));
    note_launch();
    return B200_OK;
}

// small stream-ordered copies between device memory and PINNED host memory as a kernel (16-byte accesses over PCIe, chained
// with programmatic dependent launch): a kernel boundary costs ~2 us in a stream of launches, a copy node ~8-10 us
static __global__ void __launch_bounds__(256) k_copy_bytes(const uint8_t * __restrict__ src, uint8_t * __restrict__ dst, size_t bytes,
                                                           int width) {
    pdl_launch_dependents();
    pdl_wait();
    const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t) gridDim.x * blockDim.x;
    if (width == 16) {
        for (size_t i = tid; i < bytes / 16; i += nthr) reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(src)[i];
        for (size_t i = bytes / 16 * 16 + tid; i < bytes; i += nthr) dst[i] = src[i];
    } else if (width == 4) {
        for (size_t i = tid; i < bytes / 4; i += nthr) reinterpret_cast<uint32_t *>(dst)[i] = reinterpret_cast<const uint32_t *>(src)[i];
        for (size_t i = bytes / 4 * 4 + tid; i < bytes; i += nthr) dst[i] = src[i];
    } else {
        for (size_t i = tid; i < bytes; i += nthr) dst[i] = src[i];
    }
}
// the device-side alias of a host pointer when it is pinned and mapped (UVA), else null
static void * mapped_alias(const void * host) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host) == cudaSuccess) {
        if (at.type == cudaMemoryTypeHost && at.devicePointer) return at.devicePointer;
        if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) return at.devicePointer;
    } else {
        cudaGetLastError();
    }
    return nullptr;
}

extern "C" {

int b200_qmm_abi_version(void) { return 1; }

static int copy_with_kernel(void * d, const void * s, size_t bytes, cudaStream_t stream);

int b200_copy_async(void * dst, const void * src, size_t bytes, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (!bytes) return B200_OK;
    if (!dst || !src) return B200_ERR_SHAPE;
    void * d = nullptr, * s = nullptr;
    if (bytes <= (size_t) env_knob("B200_COPY_KERNEL_MAX", 1 << 20) && env_knob("B200_HOST_ZEROCOPY", 1) != 0)
        d = mapped_alias(dst), s = d ? mapped_alias(src) : nullptr;
    if (!d || !s) {   // pageable host memory, or a large copy: the copy engine
        B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, stream));
        return B200_OK;
    }
    return copy_with_kernel(d, s, bytes, stream);
}

// the caller KNOWS that both pointers are usable from the device as they are (device memory, or host memory from
// cudaHostAlloc under UVA): no driver query per call
int b200_copy_mapped_async(void * dst, const void * src, size_t bytes, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (!bytes) return B200_OK;
    if (!dst || !src) return B200_ERR_SHAPE;
    if (bytes > (size_t) env_knob("B200_COPY_KERNEL_MAX", 1 << 20) || env_knob("B200_HOST_ZEROCOPY", 1) == 0) {
        B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, stream));
        return B200_OK;
    }
    return copy_with_kernel(dst, src, bytes, stream);
}

static int copy_with_kernel(void * d, const void * s, size_t bytes, cudaStream_t stream) {
    const uintptr_t both = (uintptr_t) d | (uintptr_t) s;
    const int width = both % 16 == 0 ? 16 : both % 4 == 0 ? 4 : 1;
    size_t blocks = (bytes / 16 + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 296) blocks = 296;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim  = dim3((unsigned) blocks, 1, 1);
    cfg.blockDim = dim3(256, 1, 1);
    cfg.stream   = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = env_knob("B200_PDL", 1);
    cfg.attrs = attr, cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_copy_bytes, (const uint8_t *) s, (uint8_t *) d, bytes, width));
    note_launch();
    return B200_OK;
}

void b200_reload_env(void) { env_forget(); }
void b200_debug_gemv_trace(void * base, int64_t launches) { gemv_trace_set(base, launches); }

const char * b200_status_string(int s) {
    switch (s) {
        case B200_OK: return "ok";
        case B200_ERR_TYPE: return "type not on the path (Q4_0, Q8_0, Q4_K, Q6_K, IQ4_XS)";
        case B200_ERR_SHAPE: return "bad shape or stride";
        case B200_ERR_ALIGNMENT: return "pointer or stride alignment";
        case B200_ERR_WORKSPACE: return "workspace missing or too small";
        case B200_ERR_NO_DEVICE: return "no sm_100 CUDA device";
        case B200_ERR_CUDA: return "CUDA runtime error";
        case B200_ERR_UNSUPPORTED: return "unsupported configuration";
        case B200_ERR_TOO_LARGE: return "problem exceeds the built-in scratch of the no-alloc entry point";
    }
    return "unknown status";
}

const char * b200_last_cuda_error(void) { return g_last_cuda_error; }

int b200_device_ok(int device) {
    int major = 0, n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess) return 0;
    return major == 10;
}

int b200_sm_count(int device) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
    return n;
}

int64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int    b200_type_supported(int type) { return weight_type(type) ? 1 : 0; }
int    b200_block_size(int type) { return blck_of(type); }
size_t b200_type_size(int type) { return tsize_of(type); }
size_t b200_row_size(int type, int64_t k) {
    const int b = blck_of(type);
    return b && k >= 0 && k % b == 0 ? (size_t) (k / b) * tsize_of(type) : 0;
}
size_t b200_weight_alloc_size(int type, int64_t k, int64_t nrows) {
    const size_t n = b200_row_size(type, k) * (size_t) nrows;
    return ((n + 15) & ~(size_t) 15) + 16;
}

int b200_quantize_q8_1(const float * x, int64_t x_row_stride, void * y, int64_t k, int64_t nrows,
                       b200_stream_t stream) {
    if (k < 0 || nrows < 0 || k % 32 != 0 || x_row_stride < k) return B200_ERR_SHAPE;
    if ((uintptr_t) x % 4 != 0 || (uintptr_t) y % 4 != 0) return B200_ERR_ALIGNMENT;
    return launch_quantize_q8_1_ggml(x, x_row_stride, y, k, nrows, (cudaStream_t) stream);
}

int b200_dequantize_f32(int type, const void * w, int64_t w_row_stride, float * y, int64_t k, int64_t nrows,
                        b200_stream_t stream) {
    if (!weight_type(type)) return B200_ERR_TYPE;
    if (k < 0 || nrows < 0 || k % blck_of(type) != 0) return B200_ERR_SHAPE;
    if (w_row_stride == 0) w_row_stride = (int64_t) b200_row_size(type, k);
    return launch_dequantize_f32(type, w, w_row_stride, y, k, nrows, (cudaStream_t) stream);
}

// ---------------------------------------------------------------------------------------------------------- MUL_MAT
// Activation quantization policy (measured on B200 with in-kernel timestamps, profiles/r01/gemv_trace_*.log):
//  * separate: one small kernel writes the SoA q8_1 row(s) to the workspace; it triggers its dependents before waiting,
//              so under programmatic dependent launch the GEMV's CTAs are resident and have their first weight stages
//              in shared memory while the previous kernel drains.  Cost: one more kernel boundary (~1.6 us).
//  * fused   : every CTA quantizes the f32 row into its shared memory cooperatively.  One launch per matmul, but 150-300
//              CTAs re-reading and re-quantizing the same row cost 2-5 us between the dependency release and the first
//              dot product (k = 4096: 6.8 us vs 6.0 us per matmul; k = 14336: 18 us vs 12 us).
// Round 2 re-measured it after the quantizer got 8 loads per lane in flight (one L2 round trip for k <= 16384), conflict-free
// unit loads and grouped launches: on whole decode tokens the fused form now wins for k <= 8192 (8B token 660 -> 687 tok/s
// with k = 4096 fused; 70B 122.0 -> 122.7 with k = 8192) and still loses for the long rows (k = 14336: 676; k = 28672:
// 117.6), where 148 CTAs re-quantizing 57-115 KB each cost more than the launch they save.  Default: fused for
// single-column matmuls with k <= 8192 (B200_FUSED_QUANT_MAX_K, 0 = never) -- except when x is pinned HOST memory read in
// place (the _host entries): 148 CTAs reading it over PCIe is far worse than one quantizer kernel doing it once.
static thread_local bool tl_x_in_host_memory = false;
static int fused_quant_max_k() { return env_knob("B200_FUSED_QUANT_MAX_K", 8192); }
static bool use_fused_quant(int64_t n, int64_t k) { return n == 1 && k <= fused_quant_max_k() && !tl_x_in_host_memory; }
// in-lane activation quantizer inside the GEMV (no quantizer kernel): B200_INLANE_QUANT=1
static bool use_inlane_quant() { return env_knob("B200_INLANE_QUANT", 0) != 0; }

static int mmq_min_batch() { return env_knob("B200_MMQ_MIN_N", 9); }
static size_t ws_align(size_t v) { return (v + 255) & ~(size_t) 255; }
// batches above MMVQ_MAX_BATCH_SIZE go to the tensor cores; so do rows too long for one GEMV CTA (k > 32768)
static bool use_mmq(int type, int64_t n, int64_t k) { return n >= mmq_min_batch() || gemv_max_cols(type, k) < 1; }

size_t b200_mul_mat_workspace_size(int type, int64_t m, int64_t n, int64_t k) {
    if (!weight_type(type) || k <= 0 || n <= 0 || m <= 0) return 0;
    if (use_mmq(type, n, k)) return ws_align(mmq_plane_bytes(type, m, k)) + mmq_workspace_bytes(n, k);
    // (the fused quantizer needs no scratch, but which form runs may differ between this query and the call -- a knob
    //  re-read, a host-resident x -- so the size never depends on it)
    return (size_t) n * q8_soa_layout(k).bytes;
}

int    b200_mmq_min_batch(void) { return mmq_min_batch(); }
int    b200_mmq_chunk(void) { return mmq_chunk(); }
size_t b200_mmq_plane_size(int type, int64_t m, int64_t k) {
    if (!weight_type(type) || m < 0 || k < 0 || k % blck_of(type) != 0) return 0;
    return mmq_plane_bytes(type, m, k);
}
int b200_mmq_build_plane(int type, const void * W, int64_t w_row_stride, void * plane, size_t plane_bytes, int64_t m,
                         int64_t k, b200_stream_t stream) {
    int st = check_weights(type, W, w_row_stride, k);
    if (st != B200_OK) return st;
    if (m < 0) return B200_ERR_SHAPE;
    if (!plane || plane_bytes < mmq_plane_bytes(type, m, k)) return B200_ERR_WORKSPACE;
    if ((uintptr_t) plane % 256 != 0) return B200_ERR_ALIGNMENT;
    return launch_plane_build(type, W, w_row_stride, plane, m, k, (cudaStream_t) stream);
}
size_t b200_mul_mat_mmq_workspace_size(int64_t m, int64_t n, int64_t k) {
    (void) m;
    return n > 0 && k > 0 ? mmq_workspace_bytes(n, k) : 0;
}
// 1 when a caller that HOLDS a prepared plane should use it at this batch: always from b200_mmq_min_batch() on, and from
// 5 columns on for the types that need one weight plane and one activation plane (Q4_K, Q6_K, IQ4_XS): the GEMV takes two
// passes above 4 columns (27-43 us for 4096 x 14336), one pass over the plane takes 24 us (tools/small_batch_mmq.py).
// Building a plane for such a batch does not pay; B200_MMQ_SMALL_N=0 switches the rule off.
int b200_mmq_use_cached_plane(int type, int64_t n) {
    if (!weight_type(type) || n <= 0) return 0;
    if (n >= mmq_min_batch()) return 1;
    const int small = env_knob("B200_MMQ_SMALL_N", 5);
    if (small <= 0 || n < small) return 0;
    return mmq_nplanes(type) == 1 && mmq_xplanes(type) == 1;
}

int b200_mul_mat_mmq(int type, const void * plane, const float * x, int64_t x_row_stride, float * dst, int64_t dst_row_stride,
                     int64_t m, int64_t n, int64_t k, void * workspace, size_t workspace_bytes, b200_stream_t stream) {
    if (!weight_type(type)) return B200_ERR_TYPE;
    if (m < 0 || n < 0 || k <= 0 || x_row_stride < k || dst_row_stride < m) return B200_ERR_SHAPE;
    if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return B200_ERR_SHAPE;
    if (m == 0 || n == 0) return B200_OK;
    if (!plane || (uintptr_t) plane % 256 != 0) return B200_ERR_ALIGNMENT;
    if (!workspace || workspace_bytes < mmq_workspace_bytes(n, k)) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
    return launch_mmq(type, plane, x, x_row_stride, dst, dst_row_stride, m, n, k, workspace, (cudaStream_t) stream);
}

int b200_mul_mat_mmq_multi(int nmat, const int * types, const void * const * planes, const float * x, int64_t x_row_stride,
                           float * const * dst, const int64_t * dst_row_stride, const int64_t * m, int64_t n, int64_t k,
                           void * workspace, size_t workspace_bytes, b200_stream_t stream) {
    if (nmat < 0 || (nmat > 0 && (!types || !planes || !dst || !dst_row_stride || !m))) return B200_ERR_SHAPE;
    if (nmat == 0 || n == 0) return B200_OK;
    if (n < 0 || k <= 0 || x_row_stride < k || n > INT32_MAX || k > INT32_MAX) return B200_ERR_SHAPE;
    if (!workspace || workspace_bytes < mmq_workspace_bytes(n, k)) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
    // matrices of one precision class (same number of weight / activation planes) share launches and one quantizer pass
    const void * pl[64];
    float *      ds[64];
    int64_t      st_[64], ms[64];
    int cnt = 0, cls_type = 0;
    auto cls = [](int t) { return mmq_nplanes(t) * 4 + mmq_xplanes(t); };
    auto flush = [&]() -> int {
        if (!cnt) return B200_OK;
        const int st = launch_mmq_multi(cls_type, cnt, pl, x, x_row_stride, ds, st_, ms, n, k, workspace, (cudaStream_t) stream);
        cnt = 0;
        return st;
    };
    for (int i = 0; i < nmat; i++) {
        if (!weight_type(types[i])) return B200_ERR_TYPE;
        if (m[i] < 0 || m[i] > INT32_MAX || dst_row_stride[i] < m[i]) return B200_ERR_SHAPE;
        if (m[i] == 0) continue;
        if (!planes[i] || (uintptr_t) planes[i] % 256 != 0 || !dst[i]) return B200_ERR_ALIGNMENT;
        if (cnt && (cls(types[i]) != cls(cls_type) || cnt == 64)) {
            const int st = flush();
            if (st != B200_OK) return st;
        }
        cls_type = types[i];
        pl[cnt] = planes[i], ds[cnt] = dst[i], st_[cnt] = dst_row_stride[i], ms[cnt] = m[i];
        cnt++;
    }
    return flush();
}

int b200_mul_mat(int type, const void * W, int64_t w_row_stride, const float * x, int64_t x_row_stride, float * dst,
                 int64_t dst_row_stride, int64_t m, int64_t n, int64_t k, void * workspace, size_t workspace_bytes,
                 b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    int st = check_weights(type, W, w_row_stride, k);
    if (st != B200_OK) return st;
    if (m < 0 || n < 0 || x_row_stride < k || dst_row_stride < m) return B200_ERR_SHAPE;
    if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return B200_ERR_SHAPE;
    if ((uintptr_t) x % 4 != 0 || (uintptr_t) dst % 4 != 0) return B200_ERR_ALIGNMENT;
    if (m == 0 || n == 0) return B200_OK;
    if (k == 0) {
        B200_CUDA_TRY(cudaMemset2DAsync(dst, dst_row_stride * 4, 0, m * 4, n, stream));
        return B200_OK;
    }
    if (use_mmq(type, n, k)) {  // tensor-core path with a transient plane carved from the workspace
        const size_t pb = ws_align(mmq_plane_bytes(type, m, k));
        if (!workspace || workspace_bytes < pb + mmq_workspace_bytes(n, k)) return B200_ERR_WORKSPACE;
        if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
        st = launch_plane_build(type, W, w_row_stride, workspace, m, k, stream);
        if (st != B200_OK) return st;
        return launch_mmq(type, workspace, x, x_row_stride, dst, dst_row_stride, m, n, k, (uint8_t *) workspace + pb, stream);
    }
    const int maxc = gemv_max_cols(type, k);

    gemv_args g = {};
    g.type = type, g.W = (const uint8_t *) W, g.w_row_stride = w_row_stride;
    g.m = m, g.k = k, g.dst_col_stride = dst_row_stride, g.x_col_stride = x_row_stride;
    g.world = 1;

    if (use_fused_quant(n, k)) {
        g.x = x, g.dst = dst, g.ncols = 1;
        return launch_gemv(g, stream);
    }
    if (use_inlane_quant()) {
        for (int64_t c0 = 0; c0 < n; c0 += maxc) {
            g.ncols = (int) (n - c0 < maxc ? n - c0 : maxc);
            g.x = x + c0 * x_row_stride, g.xq = nullptr, g.inlane = true;
            g.dst = dst + c0 * dst_row_stride;
            st = launch_gemv(g, stream);
            if (st != B200_OK) return st;
        }
        return B200_OK;
    }
    const q8_layout L = q8_soa_layout(k);
    if (!workspace || workspace_bytes < (size_t) n * L.bytes) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 16 != 0) return B200_ERR_ALIGNMENT;
    st = launch_quantize_q8_soa(x, x_row_stride, workspace, k, n, type == B200_TYPE_IQ4_XS, stream);
    if (st != B200_OK) return st;
    for (int64_t c0 = 0; c0 < n; c0 += maxc) {
        g.ncols = (int) (n - c0 < maxc ? n - c0 : maxc);
        g.xq    = (const uint8_t *) workspace + (size_t) c0 * L.bytes;
        g.dst   = dst + c0 * dst_row_stride;
        st = launch_gemv(g, stream);
        if (st != B200_OK) return st;
    }
    return B200_OK;
}

// Several MUL_MATs that read the SAME activations (q/k/v; gate/up): x is quantized once, and consecutive matrices of one
// type go out as ONE GEMV launch over their concatenated rows.
size_t b200_mul_mat_multi_workspace_size(int nmat, const int * types, const int64_t * m, int64_t n, int64_t k) {
    size_t need = 0;
    for (int i = 0; i < nmat; i++) {
        const size_t w = b200_mul_mat_workspace_size(types[i], m[i], n, k);
        if (w > need) need = w;
    }
    // both activation layouts may be live at once (IQ4_XS wants its own)
    if (k > 0 && n > 0 && n < mmq_min_batch()) {
        const size_t both = 2 * ws_align((size_t) n * q8_soa_layout(k).bytes);
        if (both > need) need = both;   // (never less than what a per-matrix fall-back to b200_mul_mat needs)
    }
    return need;
}

int b200_mul_mat_multi(int nmat, const int * types, const void * const * W, const float * x, int64_t x_row_stride,
                       float * const * dst, const int64_t * dst_row_stride, const int64_t * m, int64_t n, int64_t k,
                       void * workspace, size_t workspace_bytes, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (nmat < 0 || (nmat > 0 && (!types || !W || !dst || !dst_row_stride || !m))) return B200_ERR_SHAPE;
    if (nmat == 0 || n == 0) return B200_OK;
    bool gemv_ok = n > 0 && k > 0 && x_row_stride >= k && (uintptr_t) x % 4 == 0;
    for (int i = 0; i < nmat && gemv_ok; i++) gemv_ok = weight_type(types[i]) && !use_mmq(types[i], n, k);
    if (!gemv_ok) {  // tensor-core sizes, odd arguments: one full b200_mul_mat per matrix (which validates them)
        for (int i = 0; i < nmat; i++) {
            int st = b200_mul_mat(types[i], W[i], 0, x, x_row_stride, dst[i], dst_row_stride[i], m[i], n, k, workspace,
                                  workspace_bytes, stream_);
            if (st != B200_OK) return st;
        }
        return B200_OK;
    }
    for (int i = 0; i < nmat; i++) {
        int64_t stride = 0;
        int st = check_weights(types[i], W[i], stride, k);
        if (st != B200_OK) return st;
        if (m[i] < 0 || m[i] > INT32_MAX || dst_row_stride[i] < m[i] || (uintptr_t) dst[i] % 4 != 0) return B200_ERR_SHAPE;
    }
    const q8_layout L    = q8_soa_layout(k);
    const size_t    half = ws_align((size_t) n * L.bytes);
    if (!workspace || workspace_bytes < 2 * half) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 16 != 0) return B200_ERR_ALIGNMENT;
    uint8_t * xq[2]   = {(uint8_t *) workspace, (uint8_t *) workspace + half};   // [0] plain, [1] IQ4_XS-interleaved
    bool      have[2] = {false, false};
    int i = 0;
    while (i < nmat) {
        const int type = types[i], il = type == B200_TYPE_IQ4_XS ? 1 : 0;
        const bool inlane = use_inlane_quant(), fusedq = use_fused_quant(n, k);
        if (!inlane && !fusedq && !have[il]) {
            int st = launch_quantize_q8_soa(x, x_row_stride, xq[il], k, n, il == 1, stream);
            if (st != B200_OK) return st;
            have[il] = true;
        }
        // the run of matrices that can share a launch: same type, rows a multiple of 8, total rows fit an int
        int     j     = i + 1;
        int64_t total = m[i];
        if (m[i] % 8 == 0 && m[i] > 0)
            while (j < nmat && j - i < 4 && types[j] == type && m[j] % 8 == 0 && m[j] > 0 && total + m[j] <= INT32_MAX)
                total += m[j++];
        const int maxc = gemv_max_cols(type, k);
        for (int64_t c0 = 0; c0 < n; c0 += maxc) {
            gemv_args g = {};
            g.type = type, g.k = k, g.world = 1;
            g.w_row_stride = (k / blck_of(type)) * (int64_t) tsize_of(type);
            g.ncols = (int) (n - c0 < maxc ? n - c0 : maxc);
            g.xq    = xq[il] + (size_t) c0 * L.bytes;
            if (inlane) g.xq = nullptr, g.x = x + c0 * x_row_stride, g.x_col_stride = x_row_stride, g.inlane = true;
            if (fusedq) g.xq = nullptr, g.x = x, g.x_col_stride = x_row_stride;   // every CTA quantizes x into its shared memory
            g.W = (const uint8_t *) W[i], g.dst = dst[i] + c0 * dst_row_stride[i], g.dst_col_stride = dst_row_stride[i];
            g.m = total, g.nmat = j - i;
            int64_t at = m[i];
            for (int q = i + 1; q < j; q++) {
                g.mstart[q - i - 1] = at, g.Wm[q - i - 1] = (const uint8_t *) W[q];
                g.dstm[q - i - 1] = dst[q] + c0 * dst_row_stride[q], g.dcsm[q - i - 1] = dst_row_stride[q];
                at += m[q];
            }
            if (g.m > 0) {
                int st = launch_gemv(g, stream);
                if (st != B200_OK) return st;
            }
        }
        i = j;
    }
    return B200_OK;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t) 255; }

// Host activations that live in pinned (page-locked, UVA-mapped) memory are read by the quantizer kernel straight over
// PCIe instead of being staged by a copy: one stream operation and one boundary less per call, the same bytes crossing
// the bus inside the call.  Pageable memory (or B200_HOST_ZEROCOPY=0) takes the staged copy.  n <= 8 only: the
// tensor-core path reads x more than once.
static const float * host_x_device_alias(const float * x_host, int64_t n) {
    if (env_knob("B200_HOST_ZEROCOPY", 1) == 0 || n >= mmq_min_batch()) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, x_host) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return at.type == cudaMemoryTypeHost && at.devicePointer ? (const float *) at.devicePointer : nullptr;
}

size_t b200_mul_mat_host_workspace_size(int type, int64_t m, int64_t n, int64_t k) {
    if (!weight_type(type) || m < 0 || n < 0 || k < 0) return 0;
    return align256((size_t) n * k * 4) + align256((size_t) n * m * 4) + align256(b200_mul_mat_workspace_size(type, m, n, k));
}

int b200_mul_mat_host(int type, const void * W, int64_t w_row_stride, const float * x_host, float * dst_host, int64_t m,
                      int64_t n, int64_t k, void * workspace, size_t workspace_bytes, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (!weight_type(type)) return B200_ERR_TYPE;
    if (m < 0 || n < 0 || k < 0) return B200_ERR_SHAPE;
    if (m == 0 || n == 0) return B200_OK;
    const size_t need = b200_mul_mat_host_workspace_size(type, m, n, k);
    if (!workspace || workspace_bytes < need) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
    float * x_dev   = (float *) workspace;
    float * dst_dev = (float *) ((uint8_t *) workspace + align256((size_t) n * k * 4));
    void *  ws      = (uint8_t *) dst_dev + align256((size_t) n * m * 4);
    const float * x_in = k > 0 ? host_x_device_alias(x_host, n) : nullptr;
    if (!x_in) {
        if (k > 0) B200_CUDA_TRY(cudaMemcpyAsync(x_dev, x_host, (size_t) n * k * 4, cudaMemcpyHostToDevice, stream));
        x_in = x_dev;
    }
    tl_x_in_host_memory = x_in != x_dev;
    const int st = b200_mul_mat(type, W, w_row_stride, x_in, k, dst_dev, m, m, n, k, ws,
                                b200_mul_mat_workspace_size(type, m, n, k), stream_);
    tl_x_in_host_memory = false;
    if (st != B200_OK) return st;
    return results_to_host(dst_host, dst_dev, (size_t) n * m, stream);
}

// Several matrices, one HOST activation block: one H2D copy of x, b200_mul_mat_multi on the device, and one D2H copy per
// run of host results that lie back to back (a caller that keeps q/k/v or gate/up results adjacent gets a single copy).
size_t b200_mul_mat_multi_host_workspace_size(int nmat, const int * types, const int64_t * m, int64_t n, int64_t k) {
    if (nmat <= 0 || !types || !m || n < 0 || k < 0) return 0;
    size_t rows = 0;
    for (int i = 0; i < nmat; i++) rows += (size_t) (m[i] > 0 ? m[i] : 0);
    return align256((size_t) n * k * 4) + align256((size_t) n * rows * 4 + 256 * (size_t) nmat) +
           align256(b200_mul_mat_multi_workspace_size(nmat, types, m, n, k));
}

int b200_mul_mat_multi_host(int nmat, const int * types, const void * const * W, const float * x_host,
                            float * const * dst_host, const int64_t * m, int64_t n, int64_t k, void * workspace,
                            size_t workspace_bytes, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (nmat < 0 || nmat > 64 || (nmat > 0 && (!types || !W || !dst_host || !m))) return B200_ERR_SHAPE;
    if (nmat == 0 || n == 0) return B200_OK;
    if (n < 0 || k < 0) return B200_ERR_SHAPE;
    const size_t need = b200_mul_mat_multi_host_workspace_size(nmat, types, m, n, k);
    if (!workspace || workspace_bytes < need) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
    size_t rows = 0;
    for (int i = 0; i < nmat; i++) {
        if (m[i] < 0) return B200_ERR_SHAPE;
        rows += (size_t) m[i];
    }
    float * x_dev   = (float *) workspace;
    float * dst_dev = (float *) ((uint8_t *) workspace + align256((size_t) n * k * 4));
    void *  ws      = (uint8_t *) dst_dev + align256((size_t) n * rows * 4 + 256 * (size_t) nmat);
    const float * x_in = k > 0 ? host_x_device_alias(x_host, n) : nullptr;
    if (!x_in) {
        if (k > 0) B200_CUDA_TRY(cudaMemcpyAsync(x_dev, x_host, (size_t) n * k * 4, cudaMemcpyHostToDevice, stream));
        x_in = x_dev;
    }
    float * dd[64];
    int64_t ds[64];
    size_t  off = 0;
    for (int i = 0; i < nmat; i++) {   // device results packed in list order, exactly like adjacent host results
        dd[i] = dst_dev + off, ds[i] = m[i];
        off += (size_t) n * m[i];
    }
    tl_x_in_host_memory = x_in != x_dev;
    const int st = b200_mul_mat_multi(nmat, types, W, x_in, k, dd, ds, m, n, k, ws,
                                      b200_mul_mat_multi_workspace_size(nmat, types, m, n, k), stream_);
    tl_x_in_host_memory = false;
    if (st != B200_OK) return st;
    for (int i = 0; i < nmat;) {
        int    j     = i + 1;
        size_t count = (size_t) n * m[i];
        while (j < nmat && dst_host[j] == dst_host[i] + count) count += (size_t) n * m[j++];
        const int st2 = results_to_host(dst_host[i], dd[i], count, stream);
        if (st2 != B200_OK) return st2;
        i = j;
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------------- MUL_MAT_ID
// Activations: the n_tokens * ne11 rows of b are quantized once into the workspace (row (t, j) at index t * ne11 + j),
// then one GEMV grid with blockIdx.y = (token, slot) pair reads row (t, slot % ne11).  B200_FUSED_QUANT_MAX_K > 0
// switches rows of k <= that to in-kernel quantization (no workspace).
// group pairs by expert (one pass over an expert's weights per 4 pairs) when experts are shared: at least two pairs per
// expert on average, the 4-column GEMV shape exists for this k; B200_MMID_GROUP=0 never, =2 whenever pairs >= 2
static bool use_mmid_groups(int type, int64_t k, int64_t n_expert, int64_t npairs) {
    const int    v = env_knob("B200_MMID_GROUP", 1);
    if (v == 0 || k <= fused_quant_max_k() || use_inlane_quant()) return false;
    if (n_expert > 1024 || npairs > INT32_MAX / 8 || gemv_max_cols(type, k) < 4) return false;
    return v == 2 ? npairs >= 2 : npairs >= 2 * n_expert;      // measured cross-over (profiles/r01/mmid_bench_v1.log)
}
size_t b200_mul_mat_id_workspace_size(int type, int64_t m, int64_t k, int64_t n_expert, int64_t n_used,
                                      int64_t n_tokens) {
    (void) m;
    if (!weight_type(type) || k <= 0 || n_used <= 0 || n_tokens <= 0 || k <= fused_quant_max_k()) return 0;
    // q8 activations (ne11 <= n_used rows per token) + the grouping plan
    return ws_align((size_t) n_tokens * n_used * q8_soa_layout(k).bytes) + mmid_plan_bytes(n_tokens * n_used, n_expert);
}

int b200_mul_mat_id(int type, const void * as, int64_t expert_stride, const float * b, int64_t b_token_stride,
                    int64_t b_slot_stride, int64_t ne11, const int32_t * ids, int64_t ids_row_stride, float * dst,
                    int64_t dst_token_stride, int64_t dst_slot_stride, int64_t m, int64_t k, int64_t n_expert,
                    int64_t n_used, int64_t n_tokens, void * workspace, size_t workspace_bytes,
                    b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    int64_t row_stride = 0;
    int st = check_weights(type, as, row_stride, k);
    if (st != B200_OK) return st;
    if (m < 0 || n_expert <= 0 || n_used < 0 || n_tokens < 0 || ne11 <= 0 || ne11 > (n_used > 0 ? n_used : 1) ||
        ids_row_stride < n_used)
        return B200_ERR_SHAPE;
    if (m > INT32_MAX || k > INT32_MAX) return B200_ERR_SHAPE;
    if (expert_stride == 0) expert_stride = row_stride * m;
    if (expert_stride % (type == B200_TYPE_Q4_K ? 16 : type == B200_TYPE_IQ4_XS ? 8 : 2) != 0)
        return B200_ERR_ALIGNMENT;
    if (m == 0 || n_used == 0 || n_tokens == 0) return B200_OK;
    if (k == 0) {
        for (int64_t t = 0; t < n_tokens; t++)
            B200_CUDA_TRY(cudaMemset2DAsync(dst + t * dst_token_stride, dst_slot_stride * 4, 0, m * 4, n_used, stream));
        return B200_OK;
    }
    if (gemv_max_cols(type, k) < 1) return B200_ERR_UNSUPPORTED;

    gemv_args g = {};
    g.type = type, g.W = (const uint8_t *) as, g.w_row_stride = row_stride, g.w_expert_stride = expert_stride;
    g.x = b, g.dst = dst, g.m = m, g.k = k, g.ncols = 1, g.world = 1;
    g.x_col_stride = k, g.dst_col_stride = m;
    g.ids = ids, g.ids_row_stride = ids_row_stride, g.n_used = (int) n_used, g.ne11 = (int) ne11;
    g.n_expert = (int) n_expert;
    g.x_token_stride = b_token_stride, g.x_slot_stride = b_slot_stride;
    g.dst_token_stride = dst_token_stride, g.dst_slot_stride = dst_slot_stride;
    const bool      fused = k <= fused_quant_max_k();
    const q8_layout L     = q8_soa_layout(k);
    if (!fused) {
        if (!workspace || workspace_bytes < (size_t) n_tokens * ne11 * L.bytes) return B200_ERR_WORKSPACE;
        if ((uintptr_t) workspace % 16 != 0) return B200_ERR_ALIGNMENT;
        // rows (t, j): one strided quantizer launch per slot j keeps the row stride uniform
        for (int64_t j = 0; j < ne11; j++) {
            st = launch_quantize_q8_soa_strided(b + j * b_slot_stride, b_token_stride,
                                                (uint8_t *) workspace + (size_t) j * L.bytes, (int64_t) ne11 * L.bytes, k,
                                                n_tokens, type == B200_TYPE_IQ4_XS, stream);
            if (st != B200_OK) return st;
        }
    }
    const int64_t npairs_all = n_tokens * n_used;
    if (!fused && use_mmid_groups(type, k, n_expert, npairs_all) && mmid_max_items(npairs_all, n_expert) <= 65535) {
        const size_t q8b = ws_align((size_t) n_tokens * ne11 * L.bytes);
        if (workspace_bytes < q8b + mmid_plan_bytes(npairs_all, n_expert)) return B200_ERR_WORKSPACE;
        int32_t * plan = (int32_t *) ((uint8_t *) workspace + q8b);
        st = launch_mmid_plan(ids, ids_row_stride, n_tokens, n_used, n_expert, plan, stream);
        if (st != B200_OK) return st;
        const int64_t max_items = mmid_max_items(npairs_all, n_expert);
        g.ids = nullptr, g.ncols = 4;
        g.it_hdr = plan, g.it_expert = plan + 4, g.it_pairs = plan + 4 + max_items;
        g.npairs = max_items;
        // how many items to expect when sizing grid.x (the count itself is only known on the device): with c pairs on an
        // expert ceil(c / 4) items, i.e. c / 4 + ~3/8 on average -- rounded up a little, because one item too many
        // costs a second wave of CTAs and one too few only idles a few SMs
        const int64_t hit = n_expert < npairs_all ? n_expert : npairs_all;
        g.items_expected = (npairs_all + 3) / 4 + (hit + 1) / 2;
        if (g.items_expected > max_items) g.items_expected = max_items;
        g.xq = (const uint8_t *) workspace;
        return launch_gemv(g, stream);
    }
    // grid.y is limited to 65535 pairs per launch
    const int64_t tok_per_launch = 65535 / n_used > 0 ? 65535 / n_used : 1;
    for (int64_t t0 = 0; t0 < n_tokens; t0 += tok_per_launch) {
        const int64_t nt = n_tokens - t0 < tok_per_launch ? n_tokens - t0 : tok_per_launch;
        g.npairs = nt * n_used;
        g.x      = b + t0 * b_token_stride;
        g.xq     = fused ? nullptr : (const uint8_t *) workspace + (size_t) t0 * ne11 * L.bytes;
        g.dst    = dst + t0 * dst_token_stride;
        g.ids    = ids + t0 * ids_row_stride;
        st = launch_gemv(g, stream);
        if (st != B200_OK) return st;
    }
    return B200_OK;
}

// MUL_MAT_ID for batches on the tensor cores: the plane is that of ALL experts stacked (b200_mmq_build_plane over
// n_expert * m rows).  Worth it from a few pairs per expert on: the GEMV streams an expert once per 4 pairs.
int b200_mul_mat_id_prefers_mmq(int64_t m, int64_t n_expert, int64_t n_used, int64_t n_tokens) {
    const int64_t avg = env_knob("B200_MMID_MMQ_MIN_AVG", 6);   // pairs per expert on average
    return m > 0 && m % 128 == 0 && n_expert > 0 && n_expert <= 1024 && n_tokens * n_used >= avg * n_expert ? 1 : 0;
}
size_t b200_mul_mat_id_mmq_workspace_size(int type, int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens) {
    if (!weight_type(type) || k <= 0 || n_used <= 0 || n_tokens <= 0) return 0;
    return mmid_mmq_workspace_bytes(m, k, n_expert, n_used * n_tokens);
}
int b200_mul_mat_id_mmq(int type, const void * plane, const float * b, int64_t b_token_stride, int64_t b_slot_stride, int64_t ne11,
                        const int32_t * ids, int64_t ids_row_stride, float * dst, int64_t dst_token_stride,
                        int64_t dst_slot_stride, int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens,
                        void * workspace, size_t workspace_bytes, b200_stream_t stream) {
    if (!weight_type(type)) return B200_ERR_TYPE;
    if (m < 0 || k <= 0 || k % blck_of(type) != 0 || n_expert <= 0 || n_used < 0 || n_tokens < 0 || ne11 <= 0 ||
        ne11 > (n_used > 0 ? n_used : 1) || ids_row_stride < n_used || k > INT32_MAX)
        return B200_ERR_SHAPE;
    if (m == 0 || n_used == 0 || n_tokens == 0) return B200_OK;
    if (m % 128 != 0) return B200_ERR_UNSUPPORTED;
    if (!plane || (uintptr_t) plane % 256 != 0 || (uintptr_t) b % 4 != 0 || (uintptr_t) dst % 4 != 0) return B200_ERR_ALIGNMENT;
    const size_t need = mmid_mmq_workspace_bytes(m, k, n_expert, n_used * n_tokens);
    if (!need) return B200_ERR_UNSUPPORTED;
    if (!workspace || workspace_bytes < need) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
    return launch_mmid_mmq(type, plane, b, b_token_stride, b_slot_stride, ne11, ids, ids_row_stride, dst, dst_token_stride,
                           dst_slot_stride, m, k, n_expert, n_used, n_tokens, workspace, (cudaStream_t) stream);
}

// ---------------------------------------------------------------------------------------------------- row split
int b200_row_split(int64_t m, int world, int rank, const float * tensor_split, int64_t * row_low, int64_t * row_high) {
    if (world <= 0 || rank < 0 || rank >= world || m < 0 || !row_low || !row_high) return B200_ERR_SHAPE;
    // cumulative split points in [0,1), as ggml_backend_cuda_split_buffer_type builds them (ggml-cuda.cu:1012-1052)
    auto frac = [&](int r) -> double {
        if (r <= 0) return 0.0;
        if (r >= world) return 1.0;
        if (!tensor_split) return (double) r / world;
        double tot = 0, acc = 0;
        for (int i = 0; i < world; i++) tot += tensor_split[i];
        for (int i = 0; i < r; i++) acc += tensor_split[i];
        return tot > 0 ? acc / tot : (double) r / world;
    };
    const int64_t rounding = 128;  // get_row_rounding: mmq_y for every int8-MMA capable device (ggml-cuda.cu:727-738)
    int64_t lo = rank == 0 ? 0 : (int64_t) (m * (float) frac(rank));
    lo -= lo % rounding;
    int64_t hi;
    if (rank == world - 1) hi = m;
    else { hi = (int64_t) (m * (float) frac(rank + 1)); hi -= hi % rounding; }
    *row_low = lo, *row_high = hi;
    return B200_OK;
}

int b200_mul_mat_rowsplit_fused(int type, const void * W, const float * x, int64_t x_row_stride,
                                float * const * peer_dst, int32_t * const * peer_sync, int64_t dst_row_stride,
                                int64_t m, int64_t row_low, int64_t row_high, int64_t n, int64_t k, int world, int rank,
                                int32_t * wait_sync, int wait_arrivals, void * workspace, size_t workspace_bytes,
                                b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    int64_t row_stride = 0;
    int st = check_weights(type, W, row_stride, k);
    if (st != B200_OK) return st;
    if (world < 1 || world > 16 || rank < 0 || rank >= world || row_low < 0 || row_high < row_low || row_high > m)
        return B200_ERR_SHAPE;
    if (n < 1 || k <= 0 || x_row_stride < k || dst_row_stride < m) return B200_ERR_SHAPE;
    if (!peer_dst || !peer_sync) return B200_ERR_SHAPE;
    if (gemv_max_cols(type, k) < n) return B200_ERR_UNSUPPORTED;   // one launch = one arrival: n <= 4, k <= 32768
    for (int r = 0; r < world; r++)
        if (!peer_dst[r] || !peer_sync[r] || (uintptr_t) peer_sync[r] % 8 != 0) return B200_ERR_ALIGNMENT;
    if (wait_sync && wait_arrivals < 1) return B200_ERR_SHAPE;
    if (row_high == row_low) {  // nothing to compute on this rank, but the peers still count its arrival
        if (wait_sync) {        // ... and the dependency on the previous result is kept (stream-ordered wait + re-arm)
            st = b200_rowsplit_wait(wait_sync, wait_arrivals, stream_);
            if (st != B200_OK) return st;
        }
        rowsplit_signal_args sa = {};
        sa.world = world;
        for (int r = 0; r < world; r++) sa.flag[r] = peer_sync[r];
        k_rowsplit_signal<<<1, 32, 0, stream>>>(sa);
        B200_CUDA_TRY(cudaGetLastError());
        note_launch();
        return B200_OK;
    }
    const q8_layout L = q8_soa_layout(k);
    if (!workspace || workspace_bytes < (size_t) n * L.bytes) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 16 != 0) return B200_ERR_ALIGNMENT;
    st = launch_quantize_q8_soa(x, x_row_stride, workspace, k, n, type == B200_TYPE_IQ4_XS, stream, wait_sync,
                                wait_arrivals * B200_ARRIVAL_UNIT);
    if (st != B200_OK) return st;
    gemv_args g = {};
    g.type = type, g.W = (const uint8_t *) W, g.w_row_stride = row_stride;
    g.xq = (const uint8_t *) workspace, g.x_col_stride = x_row_stride, g.dst = peer_dst[rank] + row_low;
    g.dst_col_stride = dst_row_stride;
    g.m = row_high - row_low, g.k = k, g.ncols = (int) n;
    g.rank = rank, g.row_offset = row_low;
    if (world == 1) {  // degenerate: a plain local matmul at the row offset, plus the arrival
        g.world = 1;
        st = launch_gemv(g, stream);
        if (st != B200_OK) return st;
        if (wait_sync) B200_CUDA_TRY(cudaMemsetAsync(wait_sync, 0, 4, stream));   // re-arm (no peers to race with)
        rowsplit_signal_args sa = {};
        sa.world = 1, sa.flag[0] = peer_sync[0];
        k_rowsplit_signal<<<1, 32, 0, stream>>>(sa);
        B200_CUDA_TRY(cudaGetLastError());
        note_launch();
        return B200_OK;
    }
    g.world = world;
    g.local_count = peer_sync[rank] + 1;
    g.rearm = wait_sync;
    for (int r = 0; r < world; r++) g.peer_dst[r] = peer_dst[r], g.peer_flag[r] = peer_sync[r];
    return launch_gemv(g, stream);
}

// Prefill under split-mode row: this rank's row slices of up to four matrices that share the activations, ONE persistent
// tensor-core launch whose epilogue stores every result element into all ranks' dst (peer_delta) and counts one arrival on
// every rank's sync block -- GEMM and exchange in one kernel instead of GEMM + all-gather.
int b200_mul_mat_mmq_multi_rowsplit(int nmat, const int * types, const void * const * planes, const float * x, int64_t x_row_stride,
                                    float * const * dst_local, const int64_t * dst_row_stride, const int64_t * m_slice, int64_t n,
                                    int64_t k, int world, int rank, const int64_t * peer_delta, int32_t * const * peer_sync,
                                    void * workspace, size_t workspace_bytes, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (nmat < 1 || nmat > 4 || !types || !planes || !dst_local || !dst_row_stride || !m_slice) return B200_ERR_SHAPE;
    if (world < 1 || world > 16 || rank < 0 || rank >= world || !peer_delta || !peer_sync) return B200_ERR_SHAPE;
    if (n < 1 || k <= 0 || x_row_stride < k || n > INT32_MAX || k > INT32_MAX) return B200_ERR_SHAPE;
    for (int r = 0; r < world; r++)
        if (!peer_sync[r] || (uintptr_t) peer_sync[r] % 8 != 0 || peer_delta[r] % 4 != 0) return B200_ERR_ALIGNMENT;
    if (peer_delta[rank] != 0) return B200_ERR_SHAPE;
    const void * pl[4];
    float *      ds[4];
    int64_t      st_[4], ms[4];
    int cnt = 0, cls_type = 0;
    for (int i = 0; i < nmat; i++) {
        if (!weight_type(types[i])) return B200_ERR_TYPE;
        if (m_slice[i] < 0 || m_slice[i] > INT32_MAX || dst_row_stride[i] < m_slice[i]) return B200_ERR_SHAPE;
        if (m_slice[i] == 0) continue;
        if (!planes[i] || (uintptr_t) planes[i] % 256 != 0 || !dst_local[i] || (uintptr_t) dst_local[i] % 4 != 0) return B200_ERR_ALIGNMENT;
        if (cnt && mmq_nplanes(types[i]) * 4 + mmq_xplanes(types[i]) != mmq_nplanes(cls_type) * 4 + mmq_xplanes(cls_type))
            return B200_ERR_UNSUPPORTED;                          // one launch = one arrival: one precision class per call
        cls_type = types[i];
        pl[cnt] = planes[i], ds[cnt] = dst_local[i], st_[cnt] = dst_row_stride[i], ms[cnt] = m_slice[i], cnt++;
    }
    if (cnt == 0) {   // this rank owns no rows of the group: the peers still count its arrival
        rowsplit_signal_args sa = {};
        sa.world = world;
        for (int r = 0; r < world; r++) sa.flag[r] = peer_sync[r];
        k_rowsplit_signal<<<1, 32, 0, stream>>>(sa);
        B200_CUDA_TRY(cudaGetLastError());
        note_launch();
        return B200_OK;
    }
    if (!workspace || workspace_bytes < mmq_workspace_bytes(n, k)) return B200_ERR_WORKSPACE;
    if ((uintptr_t) workspace % 256 != 0) return B200_ERR_ALIGNMENT;
    mmq_rowsplit rs = {};
    rs.world = world, rs.rank = rank, rs.local_count = peer_sync[rank] + 1;
    for (int r = 0; r < world; r++) rs.peer_delta[r] = (long long) peer_delta[r], rs.peer_flag[r] = peer_sync[r];
    if (world == 1) {   // degenerate: a local matmul plus the arrival
        int st = launch_mmq_multi(cls_type, cnt, pl, x, x_row_stride, ds, st_, ms, n, k, workspace, stream, nullptr);
        if (st != B200_OK) return st;
        rowsplit_signal_args sa = {};
        sa.world = 1, sa.flag[0] = peer_sync[0];
        k_rowsplit_signal<<<1, 32, 0, stream>>>(sa);
        B200_CUDA_TRY(cudaGetLastError());
        note_launch();
        return B200_OK;
    }
    return launch_mmq_multi(cls_type, cnt, pl, x, x_row_stride, ds, st_, ms, n, k, workspace, stream, &rs);
}

size_t b200_mul_mat_rowsplit_workspace_size(int type, int64_t n, int64_t k) {
    if (!weight_type(type) || n <= 0 || k <= 0) return 0;
    return (size_t) n * q8_soa_layout(k).bytes;
}

// stream-ordered: later work on `stream` starts only after all `world` ranks' slices have landed in this rank's dst;
// then the counter is re-armed.  No kernel spins: these are stream memory operations (cuStreamWaitValue32 /
// cuStreamWriteValue32), capturable in CUDA graphs.
int b200_rowsplit_wait(int32_t * sync_local, int arrivals, b200_stream_t stream) {
    typedef CUresult (*memop_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
    static memop_fn wait_fn = nullptr, write_fn = nullptr;
    if (!wait_fn || !write_fn) {
        void * p = nullptr, * q = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &qr) != cudaSuccess || !p) return B200_ERR_CUDA;
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &q, cudaEnableDefault, &qr) != cudaSuccess || !q) return B200_ERR_CUDA;
        wait_fn = (memop_fn) p, write_fn = (memop_fn) q;
    }
    if (!sync_local || arrivals < 1 || arrivals > (INT32_MAX / B200_ARRIVAL_UNIT)) return B200_ERR_SHAPE;
    if (wait_fn((CUstream) stream, (CUdeviceptr) sync_local, (cuuint32_t) arrivals * B200_ARRIVAL_UNIT,
                CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS)
        return B200_ERR_CUDA;
    if (write_fn((CUstream) stream, (CUdeviceptr) sync_local, 0, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS)
        return B200_ERR_CUDA;
    return B200_OK;
}

// ---- result slices pushed to the peers by the COPY ENGINES (prefill under row split): what the reference does with
// cudaMemcpy2DAsync peer copies per device (ggml-cuda.cu:1626-1691), here from one process per GPU into IPC-mapped peers,
// one side stream per peer so that the copies to different ranks run on different engines, arrivals counted like the fused
// kernels do.  Per-device side streams / events are created on first use and live until the process ends.
struct push_state {
    cudaStream_t side[16] = {};
    cudaEvent_t  fork = nullptr, done[16] = {};
    bool         pending[16] = {};
};
static push_state g_push[MAX_DEVICES];
static std::mutex g_push_mutex;   // the side streams / events of a device are shared by all host threads that push on it

static int push_state_get(push_state *& ps, int world) {
    const int dev = current_device();
    if (dev < 0 || dev >= MAX_DEVICES) return B200_ERR_CUDA;
    ps = &g_push[dev];
    if (!ps->fork) B200_CUDA_TRY(cudaEventCreateWithFlags(&ps->fork, cudaEventDisableTiming));
    for (int r = 0; r < world; r++) {
        if (!ps->side[r]) B200_CUDA_TRY(cudaStreamCreateWithFlags(&ps->side[r], cudaStreamNonBlocking));
        if (!ps->done[r]) B200_CUDA_TRY(cudaEventCreateWithFlags(&ps->done[r], cudaEventDisableTiming));
    }
    return B200_OK;
}

static int rowsplit_push_join_locked(b200_stream_t stream_);
int b200_rowsplit_push_join(b200_stream_t stream_) {
    std::lock_guard<std::mutex> lock(g_push_mutex);
    return rowsplit_push_join_locked(stream_);
}
static int rowsplit_push_join_locked(b200_stream_t stream_) {
    push_state * ps = nullptr;
    int st = push_state_get(ps, 0);
    if (st != B200_OK) return st;
    for (int r = 0; r < 16; r++)
        if (ps->pending[r]) {
            B200_CUDA_TRY(cudaStreamWaitEvent((cudaStream_t) stream_, ps->done[r], 0));
            ps->pending[r] = false;
        }
    return B200_OK;
}

int b200_rowsplit_push(int nmat, float * const * src_local, const int64_t * row_stride, const int64_t * cols, int64_t n, int world,
                       int rank, const int64_t * peer_delta, int32_t * const * peer_sync, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (nmat < 0 || world < 1 || world > 16 || rank < 0 || rank >= world || !peer_delta || !peer_sync || n < 0) return B200_ERR_SHAPE;
    if (nmat > 0 && (!src_local || !row_stride || !cols)) return B200_ERR_SHAPE;
    for (int r = 0; r < world; r++)
        if (!peer_sync[r]) return B200_ERR_ALIGNMENT;
    std::lock_guard<std::mutex> lock(g_push_mutex);
    push_state * ps = nullptr;
    int st = push_state_get(ps, world);
    if (st != B200_OK) return st;
    st = rowsplit_push_join_locked(stream_);                    // at most one group of pushes in flight per peer
    if (st != B200_OK) return st;
    B200_CUDA_TRY(cudaEventRecord(ps->fork, stream));           // the slices are complete from here on
    for (int r = 0; r < world; r++) {
        cudaStream_t s = r == rank ? stream : ps->side[r];
        if (r != rank) {
            B200_CUDA_TRY(cudaStreamWaitEvent(s, ps->fork, 0));
            for (int i = 0; i < nmat; i++) {
                if (cols[i] <= 0 || n == 0) continue;
                if (!src_local[i] || row_stride[i] < cols[i]) return B200_ERR_SHAPE;
                B200_CUDA_TRY(cudaMemcpy2DAsync((char *) src_local[i] + peer_delta[r], (size_t) row_stride[i] * 4, src_local[i],
                                                (size_t) row_stride[i] * 4, (size_t) cols[i] * 4, (size_t) n, cudaMemcpyDefault, s));
            }
        }
        rowsplit_signal_args sa = {};                           // this rank's arrival on rank r, behind its copies to r
        sa.world = 1, sa.flag[0] = peer_sync[r];
        k_rowsplit_signal<<<1, 32, 0, s>>>(sa);
        B200_CUDA_TRY(cudaGetLastError());
        note_launch();
        if (r != rank) {
            B200_CUDA_TRY(cudaEventRecord(ps->done[r], s));
            ps->pending[r] = true;
        }
    }
    return B200_OK;
}

int b200_rowsplit_timeout_count(int * count, b200_stream_t stream) {
    if (!count) return B200_ERR_SHAPE;
    return rowsplit_timeouts_fetch(count, (cudaStream_t) stream);
}

int b200_ipc_get_handle(void * device_ptr, void * handle_out, int64_t * offset_out) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    // an IPC handle names the whole ALLOCATION the pointer lies in (a caching allocator hands out interior pointers)
    typedef CUresult (*range_fn)(CUdeviceptr *, size_t *, CUdeviceptr);
    static range_fn fn = nullptr;
    if (!fn) {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &qr) != cudaSuccess || !p) return B200_ERR_CUDA;
        fn = (range_fn) p;
    }
    CUdeviceptr base = 0;
    size_t      size = 0;
    if (fn(&base, &size, (CUdeviceptr) device_ptr) != CUDA_SUCCESS) return B200_ERR_CUDA;
    cudaIpcMemHandle_t h;
    B200_CUDA_TRY(cudaIpcGetMemHandle(&h, (void *) base));
    memcpy(handle_out, &h, sizeof(h));
    if (offset_out) *offset_out = (int64_t) ((CUdeviceptr) device_ptr - base);
    return B200_OK;
}
int b200_ipc_open_handle(const void * handle, int64_t offset, void ** device_ptr_out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void * base = nullptr;
    B200_CUDA_TRY(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    *device_ptr_out = (char *) base + offset;
    return B200_OK;
}
int b200_ipc_close_handle(void * device_ptr, int64_t offset) {
    B200_CUDA_TRY(cudaIpcCloseMemHandle((char *) device_ptr - offset));
    return B200_OK;
}

}  // extern "C"
