// kernels.h -- internal launcher interface between api.cu (the C ABI) and the kernel translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// one fused row-split launch of one rank adds exactly this much to every rank's arrival counter (its CTAs add 1 each,
// the last one the remainder), so counters compare against arrivals * UNIT whatever the grids were
#define B200_ARRIVAL_UNIT 4096

int launch_quantize_q8_1_ggml(const float * x, int64_t x_stride, void * y, int64_t k, int64_t nrows,
                              cudaStream_t stream);
// wait_sync != nullptr: the kernel first waits (bounded spin) until wait_sync[0] >= wait_target -- row-split arrivals
int launch_quantize_q8_soa(const float * x, int64_t x_stride, void * y, int64_t k, int64_t nrows, bool interleave,
                           cudaStream_t stream, int32_t * wait_sync = nullptr, int wait_target = 0);
// same, with an explicit output row pitch in bytes (>= q8_soa_layout(k).bytes, multiple of 16)
int launch_quantize_q8_soa_strided(const float * x, int64_t x_stride, void * y, int64_t y_pitch, int64_t k, int64_t nrows,
                                   bool interleave, cudaStream_t stream, int32_t * wait_sync = nullptr,
                                   int wait_target = 0);
int launch_dequantize_f32(int type, const void * w, int64_t w_stride, float * y, int64_t k, int64_t nrows,
                          cudaStream_t stream);

// Batched-GEMV ("mmvq") problem: up to 8 activation columns against a streamed weight matrix, optionally with
// per-(token,slot) expert indirection (MUL_MAT_ID) and with the row-split all-gather fused into the epilogue.
struct gemv_args {
    int            type;
    const uint8_t * W;
    int64_t        w_row_stride;     // bytes
    int64_t        w_expert_stride;  // bytes (MUL_MAT_ID)
    const float *  x;                // f32 activations (quantized in the kernel prologue) ...
    int64_t        x_col_stride;     // elements
    const uint8_t * xq;              // ... or activations already in the SoA q8 layout (column pitch = layout.bytes)
    float *        dst;
    int64_t        dst_col_stride;   // elements
    int64_t        m, k;
    int            ncols;            // 1..8
    bool           inlane;           // xq == nullptr: every lane quantizes its own 64 activations from x (no scratch)
    // MUL_MAT_ID (ids != nullptr): grid.y = n_tokens * n_used pairs, ncols must be 1
    const int32_t * ids;
    int64_t        ids_row_stride;
    int            n_used, ne11;
    int            n_expert;         // ids outside [0, n_expert) leave their dst row untouched (0 = no check)
    int64_t        npairs;
    int64_t        x_token_stride, x_slot_stride, dst_token_stride, dst_slot_stride;
    // grouped MUL_MAT_ID (it_hdr != nullptr, ids == nullptr): grid.y = npairs = upper bound of the item count,
    // items_expected sizes grid.x; the plan arrays are written by launch_mmid_plan
    const int32_t * it_hdr;
    const int32_t * it_expert;
    const int32_t * it_pairs;
    int64_t         items_expected;
    // several matrices sharing x in one launch (nmat > 1): m = total rows; matrix 0 is (W, dst, dst_col_stride),
    // matrix i >= 1 starts at virtual row mstart[i-1]; all contiguous, same type and k, every m_i % 8 == 0
    int             nmat;
    int64_t         mstart[3];
    const uint8_t * Wm[3];
    float *         dstm[3];
    int64_t         dcsm[3];
    // fused row-split all-gather (world > 1): dst rows are offset by row_offset and mirrored into every peer
    int            world, rank;
    int64_t        row_offset;
    float *        peer_dst[16];
    int32_t *      peer_flag[16];   // every rank's arrival counter
    int32_t *      local_count;     // this rank's CTA-completion counter (zero between launches)
    int32_t *      rearm;           // arrival counter the quantizer of this launch waited on: reset by the last CTA
};
// MUL_MAT_ID plan: group the (token, slot) pairs by the expert they chose, 4 pairs per work item (one GEMV column each)
size_t  mmid_plan_bytes(int64_t npairs, int64_t n_expert);
int64_t mmid_max_items(int64_t npairs, int64_t n_expert);
int     rowsplit_timeouts_fetch(int * count, cudaStream_t stream);   // quantize.cu: waits that gave up since the last call
int     launch_mmid_plan(const int32_t * ids, int64_t ids_row_stride, int64_t n_tokens, int64_t n_used, int64_t n_expert,
                         void * plan, cudaStream_t stream);
void   gemv_trace_set(void * base, long long launches);   // profiling aid (tools/gemv_trace.py)
int    gemv_max_cols(int type, int64_t k);   // how many columns fit in shared memory next to a minimal ring
int    launch_gemv(const gemv_args & a, cudaStream_t stream);

// ---- batched matmul on tcgen05 (mmq.cu): int8 "prefill plane(s)" of a weight tensor + the GEMM that consumes them
int    mmq_nplanes(int type);   // int8 planes per weight tensor (2 for Q8_0: value + residual)
int    mmq_xplanes(int type);   // int8 planes of the activations multiplied with this type (2 for Q4_0 / Q8_0)
size_t mmq_plane_bytes(int type, int64_t m, int64_t k);   // all planes of one tensor + their scales
int    mmq_chunk();   // elements of k per scale (build-time: 256 or 512)
size_t mmq_workspace_bytes(int64_t n, int64_t k);
int    launch_plane_build(int type, const void * W, int64_t w_stride, void * plane, int64_t m, int64_t k,
                          cudaStream_t stream);
// split-mode row for the tensor-core path: where the peers' copies of dst are and how the launch reports completion
struct mmq_rowsplit {
    int       world, rank;
    long long peer_delta[16];   // bytes from this rank's symmetric dst allocation to rank r's, as mapped here ([rank] = 0)
    int32_t * peer_flag[16];    // every rank's arrival counter
    int32_t * local_count;      // this rank's CTA counter (its sync block + 1)
};
// several matrices of one precision class (mmq_nplanes, mmq_xplanes of `type`) and one k against the same activations:
// one activation quantizer, shared GEMM launches.  rs != null: the row-split form (<= 4 matrices, one persistent launch
// whose epilogue stores into every rank and counts one arrival)
int    launch_mmq_multi(int type, int nmat, const void * const * planes, const float * x, int64_t x_stride,
                        float * const * dst, const int64_t * dst_stride, const int64_t * m, int64_t n, int64_t k,
                        void * workspace, cudaStream_t stream, const mmq_rowsplit * rs = nullptr);
int    launch_mmq(int type, const void * plane, const float * x, int64_t x_stride, float * dst, int64_t dst_stride, int64_t m,
                  int64_t n, int64_t k, void * workspace, cudaStream_t stream);

// MUL_MAT_ID on the tensor cores over the stacked plane of all experts (m % 128 == 0, packed experts)
size_t mmid_mmq_workspace_bytes(int64_t m, int64_t k, int64_t n_expert, int64_t npairs);
int    launch_mmid_mmq(int type, const void * plane, const float * b, int64_t b_token_stride, int64_t b_slot_stride, int64_t ne11,
                       const int32_t * ids, int64_t ids_row_stride, float * dst, int64_t dst_token_stride, int64_t dst_slot_stride,
                       int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens, void * workspace,
                       cudaStream_t stream);

}  // namespace b200
