// ggml-b200.cpp -- the ggml backend plug-in: binds the B200 quantized MUL_MAT / MUL_MAT_ID path (include/b200_qmm.h)
// behind the reference's own backend interface, so that the unmodified test-backend-ops / llama-bench / llama-cli
// of DreamChaser-luzeyu/llama.cpp.dsp drive it.
//
// Interface implemented (reference: ggml/src/ggml-backend-impl.h): ggml_backend_reg_i (:191-201), ggml_backend_device_i
// (:137-179), ggml_backend_buffer_type_i (:17-29), ggml_backend_buffer_i (:41-58), ggml_backend_i (:87-117), and the
// dynamic-loading entry points ggml_backend_init / ggml_backend_score (:215-251).  It takes the place of the reference's
// CUDA backend for this path (ggml/src/ggml-cuda/ggml-cuda.cu:516-723 buffers, :2374-2831 backend, :2885-3506
// device/reg): same conventions -- 128-byte tensor alignment, synchronous set/get_tensor, asynchronous graph_compute,
// GGML_ABORT on CUDA failures, "not supported" answered through supports_op (no CPU fallback inside the backend).
//
// Loading: build with -DGGML_BACKEND_DL and point GGML_BACKEND_PATH at libggml-b200.so
// (ggml_backend_load_all, ggml/src/ggml-backend-reg.cpp:561-590).  See INTEGRATION.md.
//
// Host code only: every device kernel is reached through the C ABI of include/b200_qmm.h.
#include "ggml-backend-impl.h"
#include "ggml-backend.h"
#include "ggml-impl.h"
#include "ggml.h"

#include "../../include/b200_qmm.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <climits>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#define B200_MAX_DEVICES 16

[[noreturn]] static void b200_cuda_fail(const char * stmt, const char * file, int line, const char * msg) {
    int dev = -1;
    cudaGetDevice(&dev);
    GGML_LOG_ERROR("B200 backend error: %s\n  device %d, %s:%d\n  %s\n", msg, dev, file, line, stmt);
    GGML_ABORT("B200 backend error");   // same convention as ggml_cuda_error, ggml-cuda.cu:66-76
}
#define CUDA_CHECK(expr)                                                                  \
    do {                                                                                  \
        cudaError_t err_ = (expr);                                                        \
        if (err_ != cudaSuccess) b200_cuda_fail(#expr, __FILE__, __LINE__, cudaGetErrorString(err_)); \
    } while (0)
#define B200_CHECK(expr)                                                                  \
    do {                                                                                  \
        int st_ = (expr);                                                                 \
        if (st_ != B200_OK) {                                                             \
            std::string m_ = std::string(b200_status_string(st_)) + ": " + b200_last_cuda_error(); \
            b200_cuda_fail(#expr, __FILE__, __LINE__, m_.c_str());                        \
        }                                                                                 \
    } while (0)

static void set_device(int device) {
    int cur = -1;
    CUDA_CHECK(cudaGetDevice(&cur));
    if (cur != device) CUDA_CHECK(cudaSetDevice(device));
}

static bool b200_type_on_path(ggml_type t) { return b200_type_supported((int) t) == 1; }

// ---------------------------------------------------------------------------------------------------- device info
struct b200_device_context {
    int         device;
    std::string name;          // "B200_0"
    std::string description;   // cudaDeviceProp::name
};

static int b200_device_count() {
    static int n = -1;
    if (n < 0) {
        int total = 0;
        if (cudaGetDeviceCount(&total) != cudaSuccess) total = 0;
        n = 0;
        for (int i = 0; i < total && i < B200_MAX_DEVICES; i++) {
            if (b200_device_ok(i)) n = i + 1;   // sm_100 parts only; a non-B200 device in front ends the list
            else break;
        }
    }
    return n;
}

// ---------------------------------------------------------------------------------------------------- device buffer
// replaces ggml_backend_cuda_buffer_* (ggml-cuda.cu:516-633).  The buffer also owns the int8 "prefill planes" of the
// quantized weights that live in it (built lazily by the first batched matmul, dropped when the tensor is rewritten).
struct b200_plane_key {
    const void * w;          // first byte of the (sub-)matrix: tensor->data + i02*nb02 + i03*nb03
    int64_t      row_stride; // nb01
    int64_t      m, k;
    int          type;
    bool operator<(const b200_plane_key & o) const {
        return std::tie(w, row_stride, m, k, type) < std::tie(o.w, o.row_stride, o.m, o.k, o.type);
    }
};
struct b200_plane {
    void * ptr  = nullptr;   // NULL until the matrix has been seen twice (or the buffer holds weights)
    int    uses = 0;
};

struct b200_buffer_context {
    int    device;
    void * base;
    std::mutex                             mtx;
    std::map<b200_plane_key, b200_plane>   planes;

    b200_buffer_context(int device, void * base) : device(device), base(base) {}
    ~b200_buffer_context() {
        set_device(device);
        for (auto & kv : planes) cudaFree(kv.second.ptr);
        CUDA_CHECK(cudaFree(base));
    }
    // forget every plane whose matrix starts inside [data, data + size): called before those bytes are rewritten
    void drop_planes(const void * data, size_t size) {
        std::lock_guard<std::mutex> lock(mtx);
        if (planes.empty()) return;
        b200_plane_key lo = {data, INT64_MIN, INT64_MIN, INT64_MIN, INT32_MIN};
        auto it  = planes.lower_bound(lo);
        bool synced = false;
        while (it != planes.end() && (const char *) it->first.w < (const char *) data + size) {
            if (it->second.ptr) {
                if (!synced) {
                    set_device(device);
                    CUDA_CHECK(cudaDeviceSynchronize());   // a matmul may still be reading it
                    synced = true;
                }
                cudaFree(it->second.ptr);
            }
            it = planes.erase(it);
        }
    }
    // the plane if one is cached already (never builds): batches of 5..8 columns use it when it is there
    const void * peek_plane(const b200_plane_key & key) {
        std::lock_guard<std::mutex> lock(mtx);
        auto it = planes.find(key);
        return it == planes.end() ? nullptr : it->second.ptr;
    }
    // The int8 plane of a weight matrix for the tensor-core path, or NULL when this call should use a transient one:
    // matrices in a weights buffer get a plane at first use, anything else at its second use without a write between.
    const void * get_plane(const b200_plane_key & key, bool weights, cudaStream_t stream) {
        std::lock_guard<std::mutex> lock(mtx);
        b200_plane & e = planes[key];
        e.uses++;
        if (e.ptr) return e.ptr;
        if (!weights && e.uses < 2) return nullptr;
        const size_t bytes = b200_mmq_plane_size(key.type, key.m, key.k);
        if (bytes == 0) return nullptr;
        set_device(device);
        if (cudaMalloc(&e.ptr, bytes) != cudaSuccess) {   // no room for the cache: keep working with transient planes
            (void) cudaGetLastError();
            e.ptr = nullptr;
            static bool warned = false;
            if (!warned) {
                GGML_LOG_WARN("%s: no room for the %.1f MiB int8 prefill plane of a weight matrix on device %d: batched "
                              "matmuls on it rebuild the plane per call (about 2.5x slower prefill)\n", __func__,
                              bytes / 1024.0 / 1024.0, device);
                warned = true;
            }
            return nullptr;
        }
        B200_CHECK(b200_mmq_build_plane(key.type, key.w, key.row_stride, e.ptr, bytes, key.m, key.k, stream));
        CUDA_CHECK(cudaStreamSynchronize(stream));        // other backend streams may use it next
        return e.ptr;
    }
};

static void b200_buffer_free_buffer(ggml_backend_buffer_t buffer) { delete (b200_buffer_context *) buffer->context; }
static void * b200_buffer_get_base(ggml_backend_buffer_t buffer) { return ((b200_buffer_context *) buffer->context)->base; }

static enum ggml_status b200_buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    // views share their source's storage; nothing to do.  Unlike the reference (ggml-cuda.cu:547-566) no zero fill of
    // row padding is needed: the kernels never let bytes past a row's end reach a result, they only need the 16 bytes
    // after the last row to be addressable, which get_alloc_size guarantees.
    (void) buffer, (void) tensor;
    return GGML_STATUS_SUCCESS;
}

static void b200_buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, uint8_t value, size_t offset,
                                      size_t size) {
    auto * ctx = (b200_buffer_context *) buffer->context;
    ctx->drop_planes(tensor->data, ggml_nbytes(tensor));
    set_device(ctx->device);
    CUDA_CHECK(cudaMemsetAsync((char *) tensor->data + offset, value, size, cudaStreamPerThread));
    CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}

static void b200_buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset,
                                   size_t size) {
    auto * ctx = (b200_buffer_context *) buffer->context;
    ctx->drop_planes(tensor->data, ggml_nbytes(tensor));
    set_device(ctx->device);
    CUDA_CHECK(cudaMemcpyAsync((char *) tensor->data + offset, data, size, cudaMemcpyHostToDevice, cudaStreamPerThread));
    CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}

static void b200_buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset,
                                   size_t size) {
    auto * ctx = (b200_buffer_context *) buffer->context;
    set_device(ctx->device);
    CUDA_CHECK(cudaMemcpyAsync(data, (const char *) tensor->data + offset, size, cudaMemcpyDeviceToHost, cudaStreamPerThread));
    CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}

static bool b200_buffer_is_ours(ggml_backend_buffer_t buffer);

static bool b200_buffer_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    if (!b200_buffer_is_ours(src->buffer)) return false;   // host sources go through set_tensor
    auto * sctx = (b200_buffer_context *) src->buffer->context;
    auto * dctx = (b200_buffer_context *) buffer->context;
    dctx->drop_planes(dst->data, ggml_nbytes(dst));
    if (sctx->device == dctx->device) {
        set_device(dctx->device);
        CUDA_CHECK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(src), cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    } else {
        CUDA_CHECK(cudaMemcpyPeerAsync(dst->data, dctx->device, src->data, sctx->device, ggml_nbytes(src), cudaStreamPerThread));
    }
    CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    return true;
}

static void b200_buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    auto * ctx = (b200_buffer_context *) buffer->context;
    ctx->drop_planes(ctx->base, buffer->size);
    set_device(ctx->device);
    CUDA_CHECK(cudaDeviceSynchronize());
    CUDA_CHECK(cudaMemset(ctx->base, value, buffer->size));
    CUDA_CHECK(cudaDeviceSynchronize());
}

static const ggml_backend_buffer_i b200_buffer_interface = {
    /* .free_buffer   = */ b200_buffer_free_buffer,
    /* .get_base      = */ b200_buffer_get_base,
    /* .init_tensor   = */ b200_buffer_init_tensor,
    /* .memset_tensor = */ b200_buffer_memset_tensor,
    /* .set_tensor    = */ b200_buffer_set_tensor,
    /* .get_tensor    = */ b200_buffer_get_tensor,
    /* .cpy_tensor    = */ b200_buffer_cpy_tensor,
    /* .clear         = */ b200_buffer_clear,
    /* .reset         = */ NULL,
};

// by buffer TYPE, like ggml_backend_buft_is_cuda (ggml-cuda.cu:709-711): llama.cpp probes weight placement with a zero-sized
// dummy buffer of the type, whose interface is empty (ggml_backend_buft_alloc_buffer, src/llama-model.cpp:238-242)
static const char * b200_buft_get_name(ggml_backend_buffer_type_t buft);
static bool b200_buffer_is_ours(ggml_backend_buffer_t buffer) { return buffer && buffer->buft && buffer->buft->iface.get_name == b200_buft_get_name; }

// ---------------------------------------------------------------------------------------------------- buffer type
// replaces ggml_backend_cuda_buffer_type_* (ggml-cuda.cu:635-723)
struct b200_buffer_type_context {
    int         device;
    std::string name;
};

static const char * b200_buft_get_name(ggml_backend_buffer_type_t buft) { return ((b200_buffer_type_context *) buft->context)->name.c_str(); }

static ggml_backend_buffer_t b200_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    auto * bctx = (b200_buffer_type_context *) buft->context;
    set_device(bctx->device);
    void *      ptr = nullptr;
    cudaError_t err = cudaMalloc(&ptr, size > 0 ? size : 1);
    if (err != cudaSuccess) {
        (void) cudaGetLastError();   // out of memory is reported as NULL, not as an abort (ggml-cuda.cu:656-662)
        GGML_LOG_ERROR("%s: allocating %.2f MiB on device %d: cudaMalloc failed: %s\n", __func__, size / 1024.0 / 1024.0,
                       bctx->device, cudaGetErrorString(err));
        return nullptr;
    }
    return ggml_backend_buffer_init(buft, b200_buffer_interface, new b200_buffer_context(bctx->device, ptr), size);
}

static size_t b200_buft_get_alignment(ggml_backend_buffer_type_t) { return 128; }

static size_t b200_buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) {
    size_t size = ggml_nbytes(tensor);
    // the GEMV streams weight rows with 16-byte-granular bulk copies: keep the bytes after the last row addressable
    if (ggml_is_quantized(tensor->type)) size += 32;
    return size;
}

static bool b200_buft_is_host(ggml_backend_buffer_type_t) { return false; }

static const ggml_backend_buffer_type_i b200_buft_interface = {
    /* .get_name       = */ b200_buft_get_name,
    /* .alloc_buffer   = */ b200_buft_alloc_buffer,
    /* .get_alignment  = */ b200_buft_get_alignment,
    /* .get_max_size   = */ NULL,
    /* .get_alloc_size = */ b200_buft_get_alloc_size,
    /* .is_host        = */ b200_buft_is_host,
};

static ggml_backend_dev_t b200_reg_get_device(ggml_backend_reg_t reg, size_t index);
static ggml_backend_reg_t b200_reg();

// ---------------------------------------------------------------------------------------------------- pinned host buffers
// replaces ggml_backend_cuda_host_buffer_type (ggml-cuda.cu:1054-1114): page-locked, device-mapped host memory behind the
// CPU buffer interface, so that the scheduler's input / output copies are asynchronous DMA and the batch <= 8 path can read
// activations from it directly (b200_mul_mat_multi_host).  GGML_CUDA_NO_PINNED in the environment switches it off, as there.
static const char * b200_host_buft_get_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
// our pinned allocations, so that tensor_set / tensor_get can tell "mapped host memory" without a driver query per copy
static std::mutex                 g_pinned_mutex;
static std::map<uintptr_t, size_t> g_pinned;   // base -> size
static bool b200_is_pinned(const void * p, size_t bytes) {
    std::lock_guard<std::mutex> lock(g_pinned_mutex);
    auto it = g_pinned.upper_bound((uintptr_t) p);
    if (it == g_pinned.begin()) return false;
    --it;
    return (uintptr_t) p + bytes <= it->first + it->second;
}
static void b200_host_buffer_free_buffer(ggml_backend_buffer_t buffer) {
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        g_pinned.erase((uintptr_t) buffer->context);
    }
    CUDA_CHECK(cudaFreeHost(buffer->context));
}
static ggml_backend_buffer_t b200_host_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    void * ptr = nullptr;
    if (getenv("GGML_CUDA_NO_PINNED") == nullptr) {
        cudaError_t err = cudaHostAlloc(&ptr, size > 0 ? size : 1, cudaHostAllocPortable | cudaHostAllocMapped);
        if (err != cudaSuccess) {
            (void) cudaGetLastError();
            GGML_LOG_DEBUG("%s: failed to allocate %.2f MiB of pinned memory: %s\n", __func__, size / 1024.0 / 1024.0,
                           cudaGetErrorString(err));
            ptr = nullptr;
        }
    }
    if (ptr == nullptr) return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);   // pageable fallback
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        g_pinned[(uintptr_t) ptr] = size > 0 ? size : 1;
    }
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(ptr, size);
    buffer->buft              = buft;
    buffer->iface.free_buffer = b200_host_buffer_free_buffer;
    return buffer;
}
static ggml_backend_buffer_type_t b200_host_buffer_type() {
    static ggml_backend_buffer_type type = {
        /* .iface   = */ {
            /* .get_name       = */ b200_host_buft_get_name,
            /* .alloc_buffer   = */ b200_host_buft_alloc_buffer,
            /* .get_alignment  = */ ggml_backend_cpu_buffer_type()->iface.get_alignment,
            /* .get_max_size   = */ NULL,
            /* .get_alloc_size = */ ggml_backend_cpu_buffer_type()->iface.get_alloc_size,
            /* .is_host        = */ ggml_backend_cpu_buffer_type()->iface.is_host,
        },
        /* .device  = */ b200_reg_get_device(b200_reg(), 0),
        /* .context = */ nullptr,
    };
    return &type;
}

static ggml_backend_buffer_type_t b200_buffer_type(int device) {
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    static ggml_backend_buffer_type types[B200_MAX_DEVICES];
    static bool                     init = false;
    if (device < 0 || device >= b200_device_count()) return nullptr;
    if (!init) {
        for (int i = 0; i < b200_device_count(); i++) {
            types[i].iface   = b200_buft_interface;
            types[i].device  = b200_reg_get_device(b200_reg(), i);
            types[i].context = new b200_buffer_type_context{i, "B200_" + std::to_string(i)};
        }
        init = true;
    }
    return &types[device];
}

// ---------------------------------------------------------------------------------------------------- split buffers
// replaces ggml_backend_cuda_split_buffer_type and its buffer (ggml-cuda.cu:725-1052): a weight tensor allocated in this
// buffer type is stored as ROW SLICES, one per B200 of the process, under the reference's tensor_split rule
// (b200_row_split = get_row_split / get_row_rounding, ggml-cuda.cu:727-753).  llama.cpp asks for it by name for
// --split-mode row (src/llama-model.cpp:323-341).  Whole-tensor set / get only, like the reference (:851-852).
struct b200_split_extra {                      // tensor->extra of a tensor in a split buffer
    void *  data[B200_MAX_DEVICES] = {};       // the slice on each device (NULL where the device owns no rows)
    int64_t row_low[B200_MAX_DEVICES] = {}, row_high[B200_MAX_DEVICES] = {};
    // int8 prefill planes of the slices, built at the first batched matmul (same cache as b200_buffer_context::planes)
    void *  plane[B200_MAX_DEVICES] = {};
};
struct b200_split_buft_context {
    int         main_device;
    float       tensor_split[B200_MAX_DEVICES];   // per-device proportions (not cumulative); all zero = equal shares
    std::string name;
};
struct b200_split_buffer_context {
    std::vector<b200_split_extra *> extras;
    ~b200_split_buffer_context() {
        for (auto * e : extras) {
            for (int d = 0; d < B200_MAX_DEVICES; d++) {
                if (!e->data[d] && !e->plane[d]) continue;
                set_device(d);
                if (e->data[d]) cudaFree(e->data[d]);
                if (e->plane[d]) cudaFree(e->plane[d]);
            }
            delete e;
        }
    }
};
static const char * b200_split_buft_get_name(ggml_backend_buffer_type_t buft) { return ((b200_split_buft_context *) buft->context)->name.c_str(); }
static bool b200_buft_is_split(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == b200_split_buft_get_name; }
static bool b200_buffer_is_split(ggml_backend_buffer_t buffer) { return buffer && b200_buft_is_split(buffer->buft); }

static void b200_split_buffer_free_buffer(ggml_backend_buffer_t buffer) { delete (b200_split_buffer_context *) buffer->context; }
static void * b200_split_buffer_get_base(ggml_backend_buffer_t) { return (void *) 0x1000; }   // slices are allocated per tensor

static enum ggml_status b200_split_buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    GGML_ASSERT(tensor->view_src == nullptr && "views of split tensors are not supported");
    GGML_ASSERT(ggml_is_contiguous(tensor) && tensor->ne[2] == 1 && tensor->ne[3] == 1 && "split tensors are plain matrices");
    auto * bctx = (b200_split_buffer_context *) buffer->context;
    auto * tctx = (b200_split_buft_context *) buffer->buft->context;
    const int     ndev = b200_device_count();
    const size_t  row  = ggml_row_size(tensor->type, tensor->ne[0]);
    auto * extra = new b200_split_extra;
    bctx->extras.push_back(extra);
    bool any = false;
    for (int d = 0; d < ndev; d++) any = any || tctx->tensor_split[d] > 0.f;
    for (int d = 0; d < ndev; d++) {
        B200_CHECK(b200_row_split(tensor->ne[1], ndev, d, any ? tctx->tensor_split : nullptr, &extra->row_low[d], &extra->row_high[d]));
        const int64_t rows = extra->row_high[d] - extra->row_low[d];
        if (rows == 0) continue;
        set_device(d);
        // + 32 bytes: the GEMV's 16-byte-granular bulk copies may touch the bytes after the last row (get_alloc_size rule)
        CUDA_CHECK(cudaMalloc(&extra->data[d], (size_t) rows * row + 32));
        CUDA_CHECK(cudaMemset((char *) extra->data[d] + (size_t) rows * row, 0, 32));
    }
    tensor->extra = extra;
    return GGML_STATUS_SUCCESS;
}
static void b200_split_drop_planes(b200_split_extra * extra) {
    for (int d = 0; d < B200_MAX_DEVICES; d++) {
        if (!extra->plane[d]) continue;
        set_device(d);
        CUDA_CHECK(cudaDeviceSynchronize());
        cudaFree(extra->plane[d]);
        extra->plane[d] = nullptr;
    }
}
static void b200_split_buffer_set_tensor(ggml_backend_buffer_t, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor) && "split tensors must be set in their entirety");
    auto * extra = (b200_split_extra *) tensor->extra;
    b200_split_drop_planes(extra);
    const size_t row = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int d = 0; d < b200_device_count(); d++) {
        if (!extra->data[d]) continue;
        set_device(d);
        CUDA_CHECK(cudaMemcpyAsync(extra->data[d], (const char *) data + extra->row_low[d] * row,
                                   (size_t) (extra->row_high[d] - extra->row_low[d]) * row, cudaMemcpyHostToDevice, cudaStreamPerThread));
    }
    for (int d = 0; d < b200_device_count(); d++) {
        if (!extra->data[d]) continue;
        set_device(d);
        CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    }
}
static void b200_split_buffer_get_tensor(ggml_backend_buffer_t, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor) && "split tensors must be read in their entirety");
    auto * extra = (b200_split_extra *) tensor->extra;
    const size_t row = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int d = 0; d < b200_device_count(); d++) {
        if (!extra->data[d]) continue;
        set_device(d);
        CUDA_CHECK(cudaMemcpyAsync((char *) data + extra->row_low[d] * row, extra->data[d],
                                   (size_t) (extra->row_high[d] - extra->row_low[d]) * row, cudaMemcpyDeviceToHost, cudaStreamPerThread));
        CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    }
}
static void b200_split_buffer_clear(ggml_backend_buffer_t, uint8_t) {}
static const ggml_backend_buffer_i b200_split_buffer_interface = {
    /* .free_buffer   = */ b200_split_buffer_free_buffer,
    /* .get_base      = */ b200_split_buffer_get_base,
    /* .init_tensor   = */ b200_split_buffer_init_tensor,
    /* .memset_tensor = */ NULL,
    /* .set_tensor    = */ b200_split_buffer_set_tensor,
    /* .get_tensor    = */ b200_split_buffer_get_tensor,
    /* .cpy_tensor    = */ NULL,
    /* .clear         = */ b200_split_buffer_clear,
    /* .reset         = */ NULL,
};
static ggml_backend_buffer_t b200_split_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    // nothing is allocated here: init_tensor allocates each tensor's slices on their devices (ggml-cuda.cu:948-957)
    return ggml_backend_buffer_init(buft, b200_split_buffer_interface, new b200_split_buffer_context, size);
}
static size_t b200_split_buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) {
    return ggml_nbytes(tensor) + 32 * (size_t) b200_device_count();
}
static const ggml_backend_buffer_type_i b200_split_buft_interface = {
    /* .get_name       = */ b200_split_buft_get_name,
    /* .alloc_buffer   = */ b200_split_buft_alloc_buffer,
    /* .get_alignment  = */ b200_buft_get_alignment,
    /* .get_max_size   = */ NULL,
    /* .get_alloc_size = */ b200_split_buft_get_alloc_size,
    /* .is_host        = */ b200_buft_is_host,
};
// the entry llama.cpp looks up as "ggml_backend_split_buffer_type" (ggml/include/ggml-backend.h:188)
static ggml_backend_buffer_type_t b200_split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    static std::map<std::pair<int, std::vector<float>>, ggml_backend_buffer_type *> types;
    const int ndev = b200_device_count();
    if (ndev == 0 || main_device < 0 || main_device >= ndev) return nullptr;
    std::vector<float> ts(B200_MAX_DEVICES, 0.f);
    if (tensor_split)
        for (int d = 0; d < ndev; d++) ts[d] = tensor_split[d];
    auto key = std::make_pair(main_device, ts);
    auto it  = types.find(key);
    if (it != types.end()) return it->second;
    auto * ctx = new b200_split_buft_context{main_device, {}, "B200_Split"};
    for (int d = 0; d < B200_MAX_DEVICES; d++) ctx->tensor_split[d] = ts[d];
    auto * type = new ggml_backend_buffer_type{
        /* .iface   = */ b200_split_buft_interface,
        /* .device  = */ b200_reg_get_device(b200_reg(), main_device),
        /* .context = */ ctx,
    };
    types[key] = type;
    return type;
}

// ---------------------------------------------------------------------------------------------------- backend (stream)
// replaces ggml_backend_cuda_* (ggml-cuda.cu:2374-2831).  One CUDA stream and one grow-only scratch allocation per
// backend instance; everything graph_compute launches is ordered on that stream, so the scratch is reused node to node.
struct b200_backend_context {
    int          device;
    std::string  name;
    cudaStream_t stream    = nullptr;
    cudaEvent_t  copy_event = nullptr;   // orders a cross-backend copy issued on this stream before the consumer's stream
    void *       workspace = nullptr;
    size_t       workspace_bytes = 0;
    // row-split matmuls driven from this backend: one stream, scratch and event pair on every OTHER device
    struct peer_state {
        cudaStream_t stream = nullptr;
        cudaEvent_t  done   = nullptr;
        void *       ws = nullptr;
        size_t       ws_bytes = 0;
    } peers[B200_MAX_DEVICES];
    cudaEvent_t split_start = nullptr;
    bool        peer_access_on = false;

    peer_state & peer(int d) {
        peer_state & p = peers[d];
        if (!p.stream) {
            set_device(d);
            CUDA_CHECK(cudaStreamCreateWithFlags(&p.stream, cudaStreamNonBlocking));
            CUDA_CHECK(cudaEventCreateWithFlags(&p.done, cudaEventDisableTiming));
        }
        return p;
    }
    void * peer_scratch(int d, size_t bytes) {
        peer_state & p = peer(d);
        if (bytes <= p.ws_bytes) return p.ws;
        set_device(d);
        if (p.ws) {
            CUDA_CHECK(cudaStreamSynchronize(p.stream));
            CUDA_CHECK(cudaFree(p.ws));
        }
        p.ws_bytes = (bytes + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
        CUDA_CHECK(cudaMalloc(&p.ws, p.ws_bytes));
        return p.ws;
    }
    // every device reads x and writes its rows of dst in this backend's device memory: peer access both ways
    // (the reference toggles it per batch size, ggml_cuda_set_peer_access ggml-cuda.cu:1315-1371; NVSwitch needs no such care)
    void enable_peer_access() {
        if (peer_access_on) return;
        const int ndev = b200_device_count();
        for (int a = 0; a < ndev; a++) {
            for (int b = 0; b < ndev; b++) {
                if (a == b) continue;
                int can = 0;
                CUDA_CHECK(cudaDeviceCanAccessPeer(&can, a, b));
                if (!can) GGML_ABORT("B200 backend: devices %d and %d have no peer access; --split-mode row needs it", a, b);
                set_device(a);
                cudaError_t err = cudaDeviceEnablePeerAccess(b, 0);
                if (err != cudaSuccess && err != cudaErrorPeerAccessAlreadyEnabled) CUDA_CHECK(err);
                (void) cudaGetLastError();
            }
        }
        peer_access_on = true;
    }

    explicit b200_backend_context(int device) : device(device), name("B200_" + std::to_string(device)) {
        set_device(device);
        CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        CUDA_CHECK(cudaEventCreateWithFlags(&copy_event, cudaEventDisableTiming));
    }
    ~b200_backend_context() {
        set_device(device);
        cudaStreamSynchronize(stream);
        if (workspace) cudaFree(workspace);
        cudaEventDestroy(copy_event);
        if (split_start) cudaEventDestroy(split_start);
        cudaStreamDestroy(stream);
        for (int d = 0; d < B200_MAX_DEVICES; d++) {
            if (!peers[d].stream) continue;
            cudaSetDevice(d);
            cudaStreamSynchronize(peers[d].stream);
            if (peers[d].ws) cudaFree(peers[d].ws);
            cudaEventDestroy(peers[d].done);
            cudaStreamDestroy(peers[d].stream);
        }
        cudaSetDevice(device);
    }
    void * scratch(size_t bytes) {
        if (bytes <= workspace_bytes) return workspace;
        if (workspace) {
            CUDA_CHECK(cudaStreamSynchronize(stream));   // kernels in flight still read the old one
            CUDA_CHECK(cudaFree(workspace));
        }
        workspace_bytes = (bytes + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
        CUDA_CHECK(cudaMalloc(&workspace, workspace_bytes));
        return workspace;
    }
};

static const char * b200_backend_get_name(ggml_backend_t backend) { return ((b200_backend_context *) backend->context)->name.c_str(); }

static void b200_backend_free(ggml_backend_t backend) {
    delete (b200_backend_context *) backend->context;
    delete backend;
}

static void b200_backend_set_tensor_async(ggml_backend_t backend, ggml_tensor * tensor, const void * data, size_t offset,
                                          size_t size) {
    auto * ctx = (b200_backend_context *) backend->context;
    ggml_backend_buffer_t buf = tensor->view_src ? tensor->view_src->buffer : tensor->buffer;
    GGML_ASSERT(b200_buffer_is_ours(buf) && "unsupported buffer type");
    ((b200_buffer_context *) buf->context)->drop_planes(tensor->data, ggml_nbytes(tensor));
    set_device(ctx->device);
    // (small copies from pinned memory -- activations from the B200_Host buffer type -- go out as a copy kernel)
    if (b200_is_pinned(data, size)) B200_CHECK(b200_copy_mapped_async((char *) tensor->data + offset, data, size, ctx->stream));
    else B200_CHECK(b200_copy_async((char *) tensor->data + offset, data, size, ctx->stream));
}

static void b200_backend_get_tensor_async(ggml_backend_t backend, const ggml_tensor * tensor, void * data, size_t offset,
                                          size_t size) {
    auto * ctx = (b200_backend_context *) backend->context;
    ggml_backend_buffer_t buf = tensor->view_src ? tensor->view_src->buffer : tensor->buffer;
    GGML_ASSERT(b200_buffer_is_ours(buf) && "unsupported buffer type");
    set_device(ctx->device);
    if (b200_is_pinned(data, size)) B200_CHECK(b200_copy_mapped_async(data, (const char *) tensor->data + offset, size, ctx->stream));
    else B200_CHECK(b200_copy_async(data, (const char *) tensor->data + offset, size, ctx->stream));
}

static bool b200_backend_is_ours(ggml_backend_t backend);

// replaces ggml_backend_cuda_cpy_tensor_async (ggml-cuda.cu:2407-2468): device-to-device copies between two B200 backends
// (same or different GPU) without a host round trip; anything else is declined and the scheduler copies through the host.
static bool b200_backend_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const ggml_tensor * src,
                                          ggml_tensor * dst) {
    if (!b200_backend_is_ours(backend_src) || !b200_backend_is_ours(backend_dst)) return false;
    ggml_backend_buffer_t sbuf = src->view_src ? src->view_src->buffer : src->buffer;
    ggml_backend_buffer_t dbuf = dst->view_src ? dst->view_src->buffer : dst->buffer;
    if (!b200_buffer_is_ours(sbuf) || !b200_buffer_is_ours(dbuf)) return false;
    auto * sctx = (b200_backend_context *) backend_src->context;
    auto * dctx = (b200_backend_context *) backend_dst->context;
    auto * sb   = (b200_buffer_context *) sbuf->context;
    auto * db   = (b200_buffer_context *) dbuf->context;
    if (sb->device != sctx->device || db->device != dctx->device) return false;
    db->drop_planes(dst->data, ggml_nbytes(dst));
    if (backend_src == backend_dst) {
        set_device(sctx->device);
        CUDA_CHECK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(dst), cudaMemcpyDeviceToDevice, sctx->stream));
        return true;
    }
    // copy on the source stream (it is ordered after the producer of src), then let the destination stream wait for it
    set_device(sctx->device);
    if (sctx->device == dctx->device) {
        CUDA_CHECK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(dst), cudaMemcpyDeviceToDevice, sctx->stream));
    } else {
        CUDA_CHECK(cudaMemcpyPeerAsync(dst->data, dctx->device, src->data, sctx->device, ggml_nbytes(dst), sctx->stream));
    }
    CUDA_CHECK(cudaEventRecord(sctx->copy_event, sctx->stream));
    set_device(dctx->device);
    CUDA_CHECK(cudaStreamWaitEvent(dctx->stream, sctx->copy_event, 0));
    return true;
}

static void b200_backend_synchronize(ggml_backend_t backend) {
    auto * ctx = (b200_backend_context *) backend->context;
    set_device(ctx->device);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}

// GGML_OP_MUL_MAT: dst[i13][i12][j][i] = sum_c src0[i13/r3][i12/r2][i][c] * src1[i13][i12][j][c]
// (ggml_compute_forward_mul_mat's broadcast rule, ggml/src/ggml-cpu/ggml-cpu.c:1266-1300; the CUDA backend's dispatcher
// for the same node is ggml_cuda_mul_mat, ggml-cuda.cu:1877-1938)
static void b200_compute_mul_mat(b200_backend_context * ctx, ggml_tensor * dst) {
    const ggml_tensor * src0 = dst->src[0];
    const ggml_tensor * src1 = dst->src[1];
    GGML_TENSOR_BINARY_OP_LOCALS

    const int     type = (int) src0->type;
    const int64_t m = ne01, k = ne00;
    const int64_t r2 = ne12 / ne02, r3 = ne13 / ne03;
    if (ggml_nelements(dst) == 0) return;

    // a weight broadcast over dim 2 of contiguous activations is one matmul with r2 * ne11 columns
    const bool    merge = r2 > 1 && nb12 == (size_t) ne11 * nb11 && nb2 == (size_t) ne1 * nb1;
    const int64_t n     = merge ? ne11 * r2 : ne11;
    const int64_t n12   = merge ? ne02 : ne12;

    ggml_backend_buffer_t wbuf = src0->view_src ? src0->view_src->buffer : src0->buffer;
    const bool mmq     = n >= b200_mmq_min_batch();
    const bool weights = wbuf->usage == GGML_BACKEND_BUFFER_USAGE_WEIGHTS;

    const size_t ws_bytes = std::max(b200_mul_mat_workspace_size(type, m, n, k), b200_mul_mat_mmq_workspace_size(m, n, k));
    void *       ws       = ws_bytes ? ctx->scratch(ws_bytes) : nullptr;

    for (int64_t i13 = 0; i13 < ne13; i13++) {
        for (int64_t i12 = 0; i12 < n12; i12++) {
            const int64_t i03 = i13 / r3;
            const int64_t i02 = merge ? i12 : i12 / r2;
            const int64_t j12 = merge ? i12 * r2 : i12;   // first dim-2 index of src1 / dst covered by this matmul
            const char *  W   = (const char *) src0->data + i02 * nb02 + i03 * nb03;
            const float * x   = (const float *) ((const char *) src1->data + j12 * nb12 + i13 * nb13);
            float *       y   = (float *) ((char *) dst->data + j12 * nb2 + i13 * nb3);

            const void * plane = nullptr;
            if (mmq) {
                b200_plane_key key = {W, (int64_t) nb01, m, k, type};
                plane = ((b200_buffer_context *) wbuf->context)->get_plane(key, weights, ctx->stream);
            } else if (b200_mmq_use_cached_plane(type, n)) {     // 5..8 columns: two GEMV passes, or one pass over a plane
                b200_plane_key key = {W, (int64_t) nb01, m, k, type};   // that an earlier prefill left behind
                plane = ((b200_buffer_context *) wbuf->context)->peek_plane(key);
            }
            if (plane) {
                B200_CHECK(b200_mul_mat_mmq(type, plane, x, nb11 / 4, y, nb1 / 4, m, n, k, ws, ctx->workspace_bytes, ctx->stream));
            } else {
                B200_CHECK(b200_mul_mat(type, W, nb01, x, nb11 / 4, y, nb1 / 4, m, n, k, ws, ctx->workspace_bytes, ctx->stream));
            }
        }
    }
}

// GGML_OP_MUL_MAT with src0 in a split buffer -- replaces the row-split half of ggml_cuda_op_mul_mat (ggml-cuda.cu:1393-1701:
// per-device row range, broadcast of the quantized activations by peer copies, 2-D peer copies of the result slices back
// to the main device, events).  Here every device computes its row slice with the SAME kernels on its own stream and
// reads x / writes its rows of dst DIRECTLY in this backend's device memory through NVLink peer access: no staging
// copies, one event pair per device.  Results are bit-identical to the unsplit matmul (rows are independent).
static void b200_compute_mul_mat_split(b200_backend_context * ctx, ggml_tensor * dst) {
    const ggml_tensor * src0 = dst->src[0];
    const ggml_tensor * src1 = dst->src[1];
    auto * extra = (b200_split_extra *) src0->extra;
    GGML_ASSERT(extra && "split tensor without slices (init_tensor was not called)");
    const int     type = (int) src0->type, ndev = b200_device_count();
    const int64_t k = src0->ne[0], n = src1->ne[1];
    if (ggml_nelements(dst) == 0) return;
    const float * x = (const float *) src1->data;
    float *       y = (float *) dst->data;
    const int64_t xs = (int64_t) (src1->nb[1] / sizeof(float)), ys = (int64_t) (dst->nb[1] / sizeof(float));
    const bool    mmq = n >= b200_mmq_min_batch();
    ctx->enable_peer_access();
    set_device(ctx->device);                                        // (the peer-access loop leaves another device current)
    if (!ctx->split_start) CUDA_CHECK(cudaEventCreateWithFlags(&ctx->split_start, cudaEventDisableTiming));
    CUDA_CHECK(cudaEventRecord(ctx->split_start, ctx->stream));    // x is ready / dst may be overwritten from here on
    for (int d = 0; d < ndev; d++) {
        const int64_t rows = extra->row_high[d] - extra->row_low[d];
        if (rows == 0) continue;
        cudaStream_t stream = ctx->stream;
        void *       ws;
        size_t       ws_bytes;
        const size_t need = std::max(b200_mul_mat_workspace_size(type, rows, n, k), b200_mul_mat_mmq_workspace_size(rows, n, k));
        if (d == ctx->device) {
            ws = need ? ctx->scratch(need) : nullptr, ws_bytes = ctx->workspace_bytes;
        } else {
            ws = need ? ctx->peer_scratch(d, need) : nullptr, ws_bytes = ctx->peers[d].ws_bytes;
            stream = ctx->peer(d).stream;
            set_device(d);
            CUDA_CHECK(cudaStreamWaitEvent(stream, ctx->split_start, 0));
        }
        set_device(d);
        float * yd = y + extra->row_low[d];                         // this device's rows of every column of dst
        if (mmq) {
            if (!extra->plane[d]) {                                 // int8 prefill plane of the slice, built once
                const size_t pb = b200_mmq_plane_size(type, rows, k);
                if (cudaMalloc(&extra->plane[d], pb) == cudaSuccess) {
                    B200_CHECK(b200_mmq_build_plane(type, extra->data[d], 0, extra->plane[d], pb, rows, k, stream));
                } else {
                    (void) cudaGetLastError();
                    extra->plane[d] = nullptr;
                }
            }
            if (extra->plane[d]) B200_CHECK(b200_mul_mat_mmq(type, extra->plane[d], x, xs, yd, ys, rows, n, k, ws, ws_bytes, stream));
            else B200_CHECK(b200_mul_mat(type, extra->data[d], 0, x, xs, yd, ys, rows, n, k, ws, ws_bytes, stream));
        } else {
            B200_CHECK(b200_mul_mat(type, extra->data[d], 0, x, xs, yd, ys, rows, n, k, ws, ws_bytes, stream));
        }
        if (d != ctx->device) CUDA_CHECK(cudaEventRecord(ctx->peers[d].done, stream));
    }
    set_device(ctx->device);
    for (int d = 0; d < ndev; d++)
        if (d != ctx->device && extra->row_high[d] > extra->row_low[d]) CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->peers[d].done, 0));
}

// GGML_OP_MUL_MAT_ID: dst[t][s] = as[ids[t][s]] * b[t][s % ne11]  (ggml.c:2771-2799).  One call, ids stay on the device.
static void b200_compute_mul_mat_id(b200_backend_context * ctx, ggml_tensor * dst) {
    const ggml_tensor * src0 = dst->src[0];
    const ggml_tensor * src1 = dst->src[1];
    const ggml_tensor * ids  = dst->src[2];
    GGML_TENSOR_BINARY_OP_LOCALS

    const int     type     = (int) src0->type;
    const int64_t m = ne01, k = ne00, n_expert = ne02;
    const int64_t n_used   = ids->ne[0];
    const int64_t n_tokens = ids->ne[1];
    if (ggml_nelements(dst) == 0) return;

    // batches: the tensor cores over the stacked int8 plane of all experts (cached with the weight buffer like the
    // MUL_MAT planes); the GEMV would stream an expert once per 4 pairs
    if (b200_mul_mat_id_prefers_mmq(m, n_expert, n_used, n_tokens) && nb02 == (size_t) m * ggml_row_size(src0->type, k)) {
        ggml_backend_buffer_t wbuf = src0->view_src ? src0->view_src->buffer : src0->buffer;
        b200_plane_key key = {src0->data, (int64_t) nb01, n_expert * m, k, type};
        const void * plane = ((b200_buffer_context *) wbuf->context)->get_plane(key, wbuf->usage == GGML_BACKEND_BUFFER_USAGE_WEIGHTS,
                                                                              ctx->stream);
        if (plane) {
            const size_t wsb = b200_mul_mat_id_mmq_workspace_size(type, m, k, n_expert, n_used, n_tokens);
            void *       wsp = ctx->scratch(wsb);
            B200_CHECK(b200_mul_mat_id_mmq(type, plane, (const float *) src1->data, nb12 / 4, nb11 / 4, ne11,
                                           (const int32_t *) ids->data, ids->nb[1] / 4, (float *) dst->data, nb2 / 4, nb1 / 4, m, k,
                                           n_expert, n_used, n_tokens, wsp, ctx->workspace_bytes, ctx->stream));
            return;
        }
    }
    const size_t ws_bytes = b200_mul_mat_id_workspace_size(type, m, k, n_expert, n_used, n_tokens);
    void *       ws       = ws_bytes ? ctx->scratch(ws_bytes) : nullptr;
    B200_CHECK(b200_mul_mat_id(type, src0->data, nb02, (const float *) src1->data, nb12 / 4, nb11 / 4, ne11,
                               (const int32_t *) ids->data, ids->nb[1] / 4, (float *) dst->data, nb2 / 4, nb1 / 4, m, k,
                               n_expert, n_used, n_tokens, ws, ctx->workspace_bytes, ctx->stream));
}

// A plain 2-D MUL_MAT (no broadcast, packed weight rows, contiguous results): the shape the grouped entry points take.
static bool b200_mul_mat_is_plain(const ggml_tensor * node) {
    if (node->op != GGML_OP_MUL_MAT || ggml_is_empty(node)) return false;
    const ggml_tensor * a = node->src[0];
    if (b200_buffer_is_split(a->buffer)) return false;
    const ggml_tensor * b = node->src[1];
    return a->ne[2] == 1 && a->ne[3] == 1 && b->ne[2] == 1 && b->ne[3] == 1 && a->nb[1] == ggml_row_size(a->type, a->ne[0]) &&
           node->nb[1] == (size_t) node->ne[0] * sizeof(float);
}
static bool b200_is_noop(const ggml_tensor * node) {
    return ggml_is_empty(node) || node->op == GGML_OP_NONE || node->op == GGML_OP_RESHAPE || node->op == GGML_OP_VIEW ||
           node->op == GGML_OP_PERMUTE || node->op == GGML_OP_TRANSPOSE;
}

// Several MUL_MAT nodes that read the SAME activations (q/k/v of a layer, gate/up): one activation quantizer and shared
// kernel launches instead of one quantizer + one launch per node (b200_mul_mat_multi / b200_mul_mat_mmq_multi).  The
// reference gets a similar effect from CUDA-graph replay of its per-node launches (ggml-cuda.cu:2646-2706); here the
// work itself shrinks (decode: 450 -> 274 launches per 8B token, 570 -> 650 tok/s in bench.py).
// Returns false if the group has to go node by node (a weight plane is not cached yet).
static bool b200_compute_mul_mat_group(b200_backend_context * ctx, ggml_tensor * const * nodes, int count) {
    const ggml_tensor * x = nodes[0]->src[1];
    const int64_t       k = x->ne[0], n = x->ne[1];
    int           types[8];
    const void *  W[8];
    float *       dst[8];
    int64_t       dstride[8], m[8], mmax = 0;
    for (int i = 0; i < count; i++) {
        const ggml_tensor * a = nodes[i]->src[0];
        types[i] = (int) a->type, W[i] = a->data, dst[i] = (float *) nodes[i]->data;
        dstride[i] = (int64_t) (nodes[i]->nb[1] / sizeof(float)), m[i] = a->ne[1];
        mmax = std::max(mmax, m[i]);
    }
    const void * planes[8];
    bool cached = n < b200_mmq_min_batch();                      // small batch: the planes only if all of them are there already
    for (int i = 0; cached && i < count; i++) {
        const ggml_tensor *   a    = nodes[i]->src[0];
        ggml_backend_buffer_t wbuf = a->view_src ? a->view_src->buffer : a->buffer;
        b200_plane_key key = {a->data, (int64_t) a->nb[1], m[i], k, types[i]};
        planes[i] = b200_mmq_use_cached_plane(types[i], n) ? ((b200_buffer_context *) wbuf->context)->peek_plane(key) : nullptr;
        cached = planes[i] != nullptr;
    }
    if (n < b200_mmq_min_batch() && !cached) {
        const size_t ws_bytes = b200_mul_mat_multi_workspace_size(count, types, m, n, k);
        void *       ws       = ws_bytes ? ctx->scratch(ws_bytes) : nullptr;
        B200_CHECK(b200_mul_mat_multi(count, types, W, (const float *) x->data, (int64_t) (x->nb[1] / sizeof(float)), dst, dstride,
                                      m, n, k, ws, ctx->workspace_bytes, ctx->stream));
        return true;
    }
    for (int i = 0; !cached && i < count; i++) {
        const ggml_tensor *   a    = nodes[i]->src[0];
        ggml_backend_buffer_t wbuf = a->view_src ? a->view_src->buffer : a->buffer;
        b200_plane_key key = {a->data, (int64_t) a->nb[1], m[i], k, types[i]};
        planes[i] = ((b200_buffer_context *) wbuf->context)->get_plane(key, wbuf->usage == GGML_BACKEND_BUFFER_USAGE_WEIGHTS,
                                                                      ctx->stream);
        if (!planes[i]) return false;
    }
    const size_t ws_bytes = b200_mul_mat_mmq_workspace_size(mmax, n, k);
    void *       ws       = ctx->scratch(ws_bytes);
    B200_CHECK(b200_mul_mat_mmq_multi(count, types, planes, (const float *) x->data, (int64_t) (x->nb[1] / sizeof(float)), dst,
                                      dstride, m, n, k, ws, ctx->workspace_bytes, ctx->stream));
    return true;
}

static enum ggml_status b200_backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    auto * ctx = (b200_backend_context *) backend->context;
    set_device(ctx->device);
    static const bool group_nodes = getenv("B200_NO_GROUPING") == nullptr;
    for (int i = 0; i < cgraph->n_nodes; i++) {
        ggml_tensor * node = cgraph->nodes[i];
        if (ggml_is_empty(node)) continue;
        if (node->op == GGML_OP_MUL_MAT && b200_buffer_is_split(node->src[0]->buffer)) {
            b200_compute_mul_mat_split(ctx, node);
            continue;
        }
        if (group_nodes && b200_mul_mat_is_plain(node)) {
            // the run of plain MUL_MATs over the same src1 that starts here (no-op nodes in between do not matter)
            ggml_tensor * run[8] = {node};
            int           count = 1, last = i;
            for (int j = i + 1; j < cgraph->n_nodes && count < 8; j++) {
                ggml_tensor * nj = cgraph->nodes[j];
                if (b200_is_noop(nj)) continue;
                if (!b200_mul_mat_is_plain(nj) || nj->src[1] != node->src[1]) break;
                run[count++] = nj, last = j;
            }
            if (count > 1 && b200_compute_mul_mat_group(ctx, run, count)) {
                i = last;
                continue;
            }
        }
        switch (node->op) {
            case GGML_OP_NONE:
            case GGML_OP_RESHAPE:
            case GGML_OP_VIEW:
            case GGML_OP_PERMUTE:
            case GGML_OP_TRANSPOSE:
                break;
            case GGML_OP_MUL_MAT:
                b200_compute_mul_mat(ctx, node);
                break;
            case GGML_OP_MUL_MAT_ID:
                b200_compute_mul_mat_id(ctx, node);
                break;
            default:
                // supports_op said no; the scheduler never sends such a node here (same abort as ggml-cuda.cu:2652-2655)
                GGML_LOG_ERROR("%s: op not supported %s (%s)\n", __func__, node->name, ggml_op_name(node->op));
                GGML_ABORT("B200 backend: unsupported op");
        }
    }
    return GGML_STATUS_SUCCESS;
}

struct b200_event_context {
    int         device;
    cudaEvent_t event;
};

static void b200_backend_event_record(ggml_backend_t backend, ggml_backend_event_t event) {
    auto * ctx = (b200_backend_context *) backend->context;
    set_device(ctx->device);
    CUDA_CHECK(cudaEventRecord(((b200_event_context *) event->context)->event, ctx->stream));
}

static void b200_backend_event_wait(ggml_backend_t backend, ggml_backend_event_t event) {
    auto * ctx = (b200_backend_context *) backend->context;
    set_device(ctx->device);
    CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ((b200_event_context *) event->context)->event, 0));
}

static const ggml_backend_i b200_backend_interface = {
    /* .get_name           = */ b200_backend_get_name,
    /* .free               = */ b200_backend_free,
    /* .set_tensor_async   = */ b200_backend_set_tensor_async,
    /* .get_tensor_async   = */ b200_backend_get_tensor_async,
    /* .cpy_tensor_async   = */ b200_backend_cpy_tensor_async,
    /* .synchronize        = */ b200_backend_synchronize,
    /* .graph_plan_create  = */ NULL,
    /* .graph_plan_free    = */ NULL,
    /* .graph_plan_update  = */ NULL,
    /* .graph_plan_compute = */ NULL,
    /* .graph_compute      = */ b200_backend_graph_compute,
    /* .event_record       = */ b200_backend_event_record,
    /* .event_wait         = */ b200_backend_event_wait,
};

static ggml_guid_t b200_backend_guid() {
    static ggml_guid guid = {0xb2, 0x00, 0x51, 0x0a, 0x74, 0xc6, 0x05, 0x7d, 0x3e, 0x6d, 0x71, 0x38, 0x6d, 0x6d, 0x51, 0x01};
    return &guid;
}
static bool b200_backend_is_ours(ggml_backend_t backend) { return backend && ggml_guid_matches(backend->guid, b200_backend_guid()); }

// ---------------------------------------------------------------------------------------------------- device
// replaces ggml_backend_cuda_device_* (ggml-cuda.cu:2885-3339)
static const char * b200_device_get_name(ggml_backend_dev_t dev) { return ((b200_device_context *) dev->context)->name.c_str(); }
static const char * b200_device_get_description(ggml_backend_dev_t dev) { return ((b200_device_context *) dev->context)->description.c_str(); }

static void b200_device_get_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    set_device(((b200_device_context *) dev->context)->device);
    CUDA_CHECK(cudaMemGetInfo(free, total));
}

static enum ggml_backend_dev_type b200_device_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }

static void b200_device_get_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    props->name        = b200_device_get_name(dev);
    props->description = b200_device_get_description(dev);
    props->type        = b200_device_get_type(dev);
    b200_device_get_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = {
        /* .async                 = */ true,
        /* .host_buffer           = */ getenv("GGML_CUDA_NO_PINNED") == nullptr,
        /* .buffer_from_host_ptr  = */ false,
        /* .events                = */ true,
    };
}

static ggml_backend_t b200_device_init_backend(ggml_backend_dev_t dev, const char * params) {
    (void) params;
    auto * dctx = (b200_device_context *) dev->context;
    return new ggml_backend{
        /* .guid      = */ b200_backend_guid(),
        /* .interface = */ b200_backend_interface,
        /* .device    = */ dev,
        /* .context   = */ new b200_backend_context(dctx->device),
    };
}

static ggml_backend_buffer_type_t b200_device_get_buffer_type(ggml_backend_dev_t dev) {
    return b200_buffer_type(((b200_device_context *) dev->context)->device);
}

static ggml_backend_buffer_type_t b200_device_get_host_buffer_type(ggml_backend_dev_t) { return b200_host_buffer_type(); }

// Which nodes this backend computes.  Everything else is answered "no" and stays with the scheduler's other backends
// (ggml_backend_sched assigns by supports_op, ggml/src/ggml-backend.cpp:701-742) -- there is no CPU code in this plug-in.
static bool b200_device_supports_op(ggml_backend_dev_t dev, const ggml_tensor * op) {
    // every source that already has storage must live in one of this device's buffers (ggml-cuda.cu:2964-2976)
    for (int i = 0; i < GGML_MAX_SRC; i++) {
        const ggml_tensor * s = op->src[i];
        if (!s || !s->buffer) continue;
        if (b200_buffer_is_split(s->buffer)) {
            // a row-split weight: only as src0 of a plain MUL_MAT (the reference: ggml-cuda.cu:2964-2976, and no split
            // buffers for MUL_MAT_ID, :2001)
            if (i != 0 || op->op != GGML_OP_MUL_MAT || s->ne[2] != 1 || s->ne[3] != 1 || !ggml_is_contiguous(s)) return false;
            if (op->src[1] && (op->src[1]->ne[2] != 1 || op->src[1]->ne[3] != 1)) return false;
            continue;
        }
        if (!b200_buffer_is_ours(s->buffer) || s->buffer->buft->device != dev) return false;
    }
    switch (op->op) {
        case GGML_OP_NONE:
        case GGML_OP_RESHAPE:
        case GGML_OP_VIEW:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT:
        case GGML_OP_MUL_MAT_ID: {
            const ggml_tensor * a = op->src[0];
            const ggml_tensor * b = op->src[1];
            if (!a || !b || !b200_type_on_path(a->type)) return false;
            if (b->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
            // rows of k contiguous elements on both sides; row / matrix strides are free (permuted and viewed operands)
            if (a->nb[0] != ggml_type_size(a->type) || b->nb[0] != sizeof(float) || op->nb[0] != sizeof(float)) return false;
            if (b->nb[1] % 16 != 0 || b->nb[2] % 16 != 0 || b->nb[3] % 16 != 0) return false;   // float4 loads of x
            if (op->nb[1] % sizeof(float) != 0 || op->nb[1] < (size_t) op->ne[0] * sizeof(float)) return false;
            if (a->ne[0] > INT32_MAX || a->ne[1] > INT32_MAX) return false;
            if (op->op == GGML_OP_MUL_MAT_ID) {
                const ggml_tensor * ids = op->src[2];
                if (!ids || ids->type != GGML_TYPE_I32 || ids->nb[0] != sizeof(int32_t)) return false;
                if (a->ne[0] > 32768) return false;   // one GEMV CTA spans a row; longer rows only exist for MUL_MAT
                // b200_mul_mat_id takes packed expert rows (the reference keeps nb01 in src0_row, ggml-cuda.cu:2037-2056;
                // a viewed / permuted expert tensor stays with another backend)
                if (a->nb[1] != ggml_row_size(a->type, a->ne[0])) return false;
                if (a->ne[3] != 1 || b->ne[3] != 1) return false;
            }
            return true;
        }
        default:
            return false;
    }
}

static bool b200_device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    if (b200_buft_is_split(buft)) return true;   // every device takes part in a split matmul (ggml-cuda.cu:3266-3268)
    return buft->iface.get_name == b200_buft_get_name && buft->device == dev;
}

static int64_t b200_op_batch_size(const ggml_tensor * op) {
    switch (op->op) {
        case GGML_OP_MUL_MAT: return op->ne[1];
        case GGML_OP_MUL_MAT_ID: return op->ne[2];
        default: return 0;
    }
}

// same threshold as ggml_backend_cuda_device_offload_op (ggml-cuda.cu:3285-3291): worth copying CPU-resident weights over
// for batches of 32 tokens and more
static bool b200_device_offload_op(ggml_backend_dev_t, const ggml_tensor * op) { return b200_op_batch_size(op) >= 32; }

static ggml_backend_event_t b200_device_event_new(ggml_backend_dev_t dev) {
    auto * dctx = (b200_device_context *) dev->context;
    set_device(dctx->device);
    auto * e = new b200_event_context{dctx->device, nullptr};
    CUDA_CHECK(cudaEventCreateWithFlags(&e->event, cudaEventDisableTiming));
    return new ggml_backend_event{
        /* .device  = */ dev,
        /* .context = */ e,
    };
}

static void b200_device_event_free(ggml_backend_dev_t, ggml_backend_event_t event) {
    auto * e = (b200_event_context *) event->context;
    set_device(e->device);
    CUDA_CHECK(cudaEventDestroy(e->event));
    delete e;
    delete event;
}

static void b200_device_event_synchronize(ggml_backend_dev_t, ggml_backend_event_t event) {
    auto * e = (b200_event_context *) event->context;
    set_device(e->device);
    CUDA_CHECK(cudaEventSynchronize(e->event));
}

static const ggml_backend_device_i b200_device_interface = {
    /* .get_name             = */ b200_device_get_name,
    /* .get_description      = */ b200_device_get_description,
    /* .get_memory           = */ b200_device_get_memory,
    /* .get_type             = */ b200_device_get_type,
    /* .get_props            = */ b200_device_get_props,
    /* .init_backend         = */ b200_device_init_backend,
    /* .get_buffer_type      = */ b200_device_get_buffer_type,
    /* .get_host_buffer_type = */ b200_device_get_host_buffer_type,
    /* .buffer_from_host_ptr = */ NULL,
    /* .supports_op          = */ b200_device_supports_op,
    /* .supports_buft        = */ b200_device_supports_buft,
    /* .offload_op           = */ b200_device_offload_op,
    /* .event_new            = */ b200_device_event_new,
    /* .event_free           = */ b200_device_event_free,
    /* .event_synchronize    = */ b200_device_event_synchronize,
};

// ---------------------------------------------------------------------------------------------------- registry
// replaces ggml_backend_cuda_reg_* (ggml-cuda.cu:3341-3506)
struct b200_reg_context {
    std::vector<ggml_backend_dev_t> devices;
};

static const char * b200_reg_get_name(ggml_backend_reg_t) { return "B200"; }
static size_t       b200_reg_get_device_count(ggml_backend_reg_t reg) { return ((b200_reg_context *) reg->context)->devices.size(); }

static ggml_backend_dev_t b200_reg_get_device(ggml_backend_reg_t reg, size_t index) {
    auto * ctx = (b200_reg_context *) reg->context;
    GGML_ASSERT(index < ctx->devices.size());
    return ctx->devices[index];
}

static ggml_backend_feature * b200_get_features(ggml_backend_reg_t) {
    static ggml_backend_feature features[] = {
        {"ARCH", "sm_100a"},
        {"GEMV", "TMA bulk + dp4a"},
        {"MMQ", "tcgen05 i8 + TMEM"},
        {nullptr, nullptr},
    };
    return features;
}

static void * b200_reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_get_features") == 0) return (void *) b200_get_features;
    // --split-mode row: llama.cpp fetches the split buffer type by this name (src/llama-model.cpp:323-341)
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *) b200_split_buffer_type;
    return nullptr;
}

static const ggml_backend_reg_i b200_reg_interface = {
    /* .get_name         = */ b200_reg_get_name,
    /* .get_device_count = */ b200_reg_get_device_count,
    /* .get_device       = */ b200_reg_get_device,
    /* .get_proc_address = */ b200_reg_get_proc_address,
};

static ggml_backend_reg_t b200_reg() {
    static ggml_backend_reg reg;
    static std::once_flag   once;
    std::call_once(once, [] {
        auto * ctx = new b200_reg_context;
        reg = ggml_backend_reg{
            /* .api_version = */ GGML_BACKEND_API_VERSION,
            /* .iface       = */ b200_reg_interface,
            /* .context     = */ ctx,
        };
        for (int i = 0; i < b200_device_count(); i++) {
            cudaDeviceProp prop;
            CUDA_CHECK(cudaGetDeviceProperties(&prop, i));
            auto * dctx = new b200_device_context{i, "B200_" + std::to_string(i), prop.name};
            ctx->devices.push_back(new ggml_backend_device{
                /* .iface   = */ b200_device_interface,
                /* .reg     = */ &reg,
                /* .context = */ dctx,
            });
        }
    });
    return &reg;
}

// public entry points, named after the reference's per-backend headers (ggml/include/ggml-cuda.h:26-47)
extern "C" {
GGML_BACKEND_API ggml_backend_reg_t ggml_backend_b200_reg(void) { return b200_reg(); }
GGML_BACKEND_API int                ggml_backend_b200_get_device_count(void) { return b200_device_count(); }
GGML_BACKEND_API bool ggml_backend_is_b200(ggml_backend_t backend) { return backend && ggml_guid_matches(backend->guid, b200_backend_guid()); }
GGML_BACKEND_API ggml_backend_t ggml_backend_b200_init(int device) {
    if (device < 0 || device >= b200_device_count()) {
        GGML_LOG_ERROR("%s: invalid device %d\n", __func__, device);
        return nullptr;
    }
    return b200_device_init_backend(b200_reg_get_device(b200_reg(), device), nullptr);
}
GGML_BACKEND_API ggml_backend_buffer_type_t ggml_backend_b200_buffer_type(int device) { return b200_buffer_type(device); }
}

static int b200_score() { return b200_device_count() > 0 ? 100 : 0; }

GGML_BACKEND_DL_IMPL(ggml_backend_b200_reg)
GGML_BACKEND_DL_SCORE_IMPL(b200_score)
