// common.cuh -- block formats, PTX helpers and launch bookkeeping shared by the sm_100a kernels.
//
// Block formats are ggml's on-disk = in-memory weight formats (reference: ggml/src/ggml-common.h:167-172, 209-227,
// 285-296, 320-326, 412-418).  Kernels treat weight rows as raw byte streams staged through shared memory with 1-D
// bulk TMA copies (cp.async.bulk), so the structs below are used only for sizes and field offsets.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/b200_qmm.h"

namespace b200 {

constexpr int QK  = 32;   // elements per q4_0 / q8_0 / q8_1 block
constexpr int QKK = 256;  // elements per K-quant / iq4_xs super-block

template <int TYPE> struct qtraits;
template <> struct qtraits<B200_TYPE_Q4_0>   { static constexpr int blck = 32,  bytes = 18,  align = 2;  };
template <> struct qtraits<B200_TYPE_Q8_0>   { static constexpr int blck = 32,  bytes = 34,  align = 2;  };
template <> struct qtraits<B200_TYPE_Q4_K>   { static constexpr int blck = 256, bytes = 144, align = 16; };
template <> struct qtraits<B200_TYPE_Q6_K>   { static constexpr int blck = 256, bytes = 210, align = 2;  };
template <> struct qtraits<B200_TYPE_IQ4_XS> { static constexpr int blck = 256, bytes = 136, align = 8;  };

// field offsets inside a block
constexpr int Q4K_OFF_SCALES = 4,  Q4K_OFF_QS = 16;                       // {half d; half dmin; u8 scales[12]; u8 qs[128]}
constexpr int Q6K_OFF_QH = 128, Q6K_OFF_SCALES = 192, Q6K_OFF_D = 208;    // {u8 ql[128]; u8 qh[64]; i8 scales[16]; half d}
constexpr int IQ4XS_OFF_SCALES_H = 2, IQ4XS_OFF_SCALES_L = 4, IQ4XS_OFF_QS = 8;  // {half d; u16 sh; u8 sl[4]; u8 qs[128]}

// non-linear 4-bit codebook of IQ4_NL / IQ4_XS (reference: ggml/src/ggml-quants.c:2434, ggml-cuda/common.cuh:455),
// packed little-endian four entries per word for prmt-based lookup
constexpr uint32_t IQ4NL_T0 = 0xBFAD9881u;  // -127 -104  -83  -65
constexpr uint32_t IQ4NL_T1 = 0xF6EADDCFu;  //  -49  -35  -22  -10
constexpr uint32_t IQ4NL_T2 = 0x26190D01u;  //    1   13   25   38
constexpr uint32_t IQ4NL_T3 = 0x71594535u;  //   53   69   89  113

extern thread_local char g_last_cuda_error[256];
void note_launch();
int  check_cuda(cudaError_t e, const char * what);
// per-DEVICE launch state: a kernel's dynamic shared-memory limit is an attribute of the function in one device's
// context, and one process may drive several B200s (the ggml plug-in registers B200_0..7), so both are keyed by the
// current device (the reference keeps shmem_limit_raised[GGML_CUDA_MAX_DEVICES] for the same reason, mmq.cuh:2793)
constexpr int MAX_DEVICES = 64;
int current_device();                 // cudaGetDevice, -1 on failure
int sm_count_current_device();        // cached per device; 0 on failure
int env_knob(const char * name, int dflt);   // integer environment knob, read ONCE per process per name

// raise `kern`'s dynamic shared-memory limit once per device (thread-safe: setting it twice is harmless)
#define B200_SMEM_ATTR_ONCE(kern, bytes)                                                                          \
    do {                                                                                                          \
        static std::atomic<uint64_t> done_{0};                                                                    \
        const int dev_ = ::b200::current_device();                                                                \
        if (dev_ < 0 || dev_ >= ::b200::MAX_DEVICES) return B200_ERR_CUDA;                                        \
        if (!(done_.load(std::memory_order_acquire) >> dev_ & 1)) {                                               \
            B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (bytes))); \
            done_.fetch_or(uint64_t(1) << dev_, std::memory_order_release);                                       \
        }                                                                                                         \
    } while (0)

#define B200_CUDA_TRY(expr)                                        \
    do {                                                           \
        int _st = ::b200::check_cuda((expr), #expr);               \
        if (_st != B200_OK) return _st;                            \
    } while (0)

// ---------------------------------------------------------------------------------------------------- device helpers
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// try_wait with a suspend-time hint: a warp whose barrier phase is not complete sleeps in hardware (up to ~the hint, it is
// woken by the arrival) instead of spinning.  Without it the polling loops of the waiting roles (TMA producer, MMA issuer,
// epilogue warps at a chunk start) were 19 % of all issued instructions of the prefill GEMM and competed with the
// epilogue math for issue slots (profiles/r02/ncu_mmq_p_*_source_summary.txt: BRA / SYNCS / NOP / YIELD).
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(0x989680u)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}

// 1-D bulk TMA copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP).
// dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void tma_load_1d(uint32_t dst_smem, const void * src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     dst_smem),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// same with an L2 eviction-priority hint (weights are streamed once per token: evict_first)
__device__ __forceinline__ void tma_load_1d_hint(uint32_t dst_smem, const void * src, uint32_t bytes, uint32_t bar,
                                                 uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(dst_smem),
        "l"(src), "r"(bytes), "r"(bar), "l"(policy)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}

// programmatic dependent launch: let the next grid start early / wait for the previous grid's results
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds16(uint32_t a) {
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t a) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
// 16 bytes from a shared address that is only 2-byte aligned: five aligned words + clamped funnel shifts
__device__ __forceinline__ uint4 lds128_align2(uint32_t a) {
    const uint32_t a4 = a & ~3u;
    const uint32_t sh = (a & 2u) * 8u;
    const uint32_t w0 = lds32(a4), w1 = lds32(a4 + 4), w2 = lds32(a4 + 8), w3 = lds32(a4 + 12), w4 = lds32(a4 + 16);
    uint4 v;
    v.x = __funnelshift_rc(w0, w1, sh);
    v.y = __funnelshift_rc(w1, w2, sh);
    v.z = __funnelshift_rc(w2, w3, sh);
    v.w = __funnelshift_rc(w3, w4, sh);
    return v;
}

__device__ __forceinline__ float half_bits_to_float(uint32_t h) { return __half2float(__ushort_as_half((unsigned short) h)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

#endif  // __CUDACC__

}  // namespace b200
