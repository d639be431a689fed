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

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
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
struct b200_plane {
    void * ptr;
    size_t bytes;
};

struct b200_buffer_context {
    int    device;
    void * base;
    std::mutex                               mtx;
    std::unordered_map<const void *, b200_plane> planes;   // key: tensor->data

    b200_buffer_context(int device, void * base) : device(device), base(base) {}
    ~b200_buffer_context() {
        set_device(device);
        for (auto & kv : planes) cudaFree(kv.second.ptr);
        CUDA_CHECK(cudaFree(base));
    }
    void drop_plane(const void * data) {
        std::lock_guard<std::mutex> lock(mtx);
        auto it = planes.find(data);
        if (it != planes.end()) {
            set_device(device);
            CUDA_CHECK(cudaDeviceSynchronize());   // a matmul may still be reading it
            cudaFree(it->second.ptr);
            planes.erase(it);
        }
    }
    void drop_all_planes() {
        std::lock_guard<std::mutex> lock(mtx);
        if (planes.empty()) return;
        set_device(device);
        CUDA_CHECK(cudaDeviceSynchronize());
        for (auto & kv : planes) cudaFree(kv.second.ptr);
        planes.clear();
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
    ctx->drop_plane(tensor->data);
    set_device(ctx->device);
    CUDA_CHECK(cudaMemsetAsync((char *) tensor->data + offset, value, size, cudaStreamPerThread));
    CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
}

static void b200_buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset,
                                   size_t size) {
    auto * ctx = (b200_buffer_context *) buffer->context;
    ctx->drop_plane(tensor->data);
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
    dctx->drop_plane(dst->data);
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
    ctx->drop_all_planes();
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

static bool b200_buffer_is_ours(ggml_backend_buffer_t buffer) { return buffer && buffer->iface.free_buffer == b200_buffer_free_buffer; }

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

