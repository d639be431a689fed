/*
 * b200_qmm.h -- C ABI of the B200-native (sm_100a) quantized MUL_MAT / MUL_MAT_ID hot path.
 *
 * This is the drop-in boundary: plain pointers, sizes and a CUDA stream handle; no C++ or torch types.  Every pointer
 * named "device" is a CUDA device pointer valid on the current device; every call is asynchronous on `stream` and
 * returns a b200_status (0 = success).  There is NO CPU fallback: when no sm_100 device is present the calls fail with
 * B200_ERR_NO_DEVICE / B200_ERR_CUDA.
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference tree,
 * DreamChaser-luzeyu/llama.cpp.dsp @ 9561c161).  The ggml backend plug-in that binds these into
 * ggml_backend_i / supports_op / graph_compute lives in llama.cpp.dsp_b200/ggml_backend/ (see INTEGRATION.md).
 */
#ifndef B200_QMM_H
#define B200_QMM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define B200_API
#else
#define B200_API __attribute__((visibility("default")))
#endif

/* cudaStream_t, kept opaque so that C callers need no CUDA headers */
typedef void * b200_stream_t;

typedef enum {
    B200_OK              = 0,
    B200_ERR_TYPE        = 1, /* ggml type not on the path */
    B200_ERR_SHAPE       = 2, /* k not a multiple of the block size, negative sizes, ... */
    B200_ERR_ALIGNMENT   = 3, /* pointer / stride alignment the kernels require is violated */
    B200_ERR_WORKSPACE   = 4, /* workspace missing or too small */
    B200_ERR_NO_DEVICE   = 5, /* no CUDA device of compute capability 10.x */
    B200_ERR_CUDA        = 6, /* a CUDA runtime call failed: see b200_last_cuda_error() */
    B200_ERR_UNSUPPORTED = 7,
    B200_ERR_TOO_LARGE   = 8  /* b200_mul_mat_noalloc: the problem needs more than its built-in scratch */
} b200_status;

/* weight types, numerically equal to enum ggml_type (ggml/include/ggml.h:350-392) */
typedef enum {
    B200_TYPE_Q4_0   = 2,
    B200_TYPE_Q8_0   = 8,
    B200_TYPE_Q8_1   = 9,
    B200_TYPE_Q4_K   = 12,
    B200_TYPE_Q6_K   = 14,
    B200_TYPE_IQ4_XS = 23
} b200_type;

/* ---- introspection ------------------------------------------------------------------------------------------- */

B200_API int          b200_qmm_abi_version(void);          /* bumped on any signature change; currently 1 */
B200_API const char * b200_status_string(int status);
B200_API const char * b200_last_cuda_error(void);          /* text of the last failing CUDA call on this thread */
B200_API int          b200_device_ok(int device);          /* 1 if `device` is an sm_100 part this library runs on */
B200_API int          b200_sm_count(int device);
/* number of kernels this library has launched since load (all threads); bench.py reports the delta as gpu_launches */
B200_API int64_t      b200_launch_count(void);
/* Tuning knobs (B200_* environment variables) are read ONCE per process at first use; a caller that
 * changes them afterwards (tests, sweep tools) calls this to have them read again. */
B200_API void         b200_reload_env(void);
/* profiling aid (tools/gemv_trace.py): the next `launches` GEMV launches write 4096 x 8 globaltimer stamps each to `base` */
B200_API void         b200_debug_gemv_trace(void * device_base, int64_t launches);

/* is (type) one of Q4_0, Q8_0, Q4_K, Q6_K, IQ4_XS?      -- replaces the type switch in ggml_cuda_mul_mat,
 * ggml/src/ggml-cuda/ggml-cuda.cu:1877-1938 and ggml_cuda_should_use_mmq, ggml/src/ggml-cuda/mmq.cu:99-153 */
B200_API int    b200_type_supported(int type);
B200_API int    b200_block_size(int type);                  /* elements per block   (ggml_blck_size)  */
B200_API size_t b200_type_size(int type);                   /* bytes per block      (ggml_type_size)  */
B200_API size_t b200_row_size(int type, int64_t k);         /* bytes per row of k   (ggml_row_size)   */
/*
 * Bytes a device allocation holding `nrows` rows must have so that the kernels' 16-byte-granular bulk loads stay
 * inside it -- replaces ggml_backend_cuda_buffer_type_get_alloc_size, ggml/src/ggml-cuda/ggml-cuda.cu:675-688.
 * (Any cudaMalloc'ed block already satisfies this: the driver's granularity is >= 256 B.)
 */
B200_API size_t b200_weight_alloc_size(int type, int64_t k, int64_t nrows);

/* ---- activation quantizer ------------------------------------------------------------------------------------ */

/*
 * f32 rows -> ggml block_q8_1 {half d; half s; int8 qs[32]} (ggml/src/ggml-common.h:216-227), BIT-EXACT with
 * quantize_row_q8_1_ref (ggml/src/ggml-quants.c:220-253).
 * Replaces quantize_row_q8_1_cuda, ggml/src/ggml-cuda/quantize.cu:129-141 (whose own arithmetic is NOT bit-exact with
 * the CPU reference; this one is).
 *   x : device, nrows rows of k f32, row stride x_row_stride elements (>= k)
 *   y : device, nrows * (k/32) blocks of 36 bytes, rows packed back to back
 */
B200_API int b200_quantize_q8_1(const float * x, int64_t x_row_stride, void * y, int64_t k, int64_t nrows,
                                b200_stream_t stream);

/* ---- dequantizer ---------------------------------------------------------------------------------------------- */

/*
 * quantized rows -> f32, BIT-EXACT with dequantize_row_{q4_0,q8_0,q4_K,q6_K,iq4_xs}
 * (ggml/src/ggml-quants.c:255-273, 349-363, 1280-1302, 1690-1722, 2454-2476).
 * Replaces the to_fp32 table ggml_get_to_fp32_cuda, ggml/src/ggml-cuda/convert.cu:654 (kernels :6-25,70-96,189-218,
 * 254-278,440-455).
 *   w : device, nrows rows, row stride w_row_stride bytes (0 => b200_row_size(type,k))
 *   y : device, nrows rows of k f32, packed
 */
B200_API int b200_dequantize_f32(int type, const void * w, int64_t w_row_stride, float * y, int64_t k, int64_t nrows,
                                 b200_stream_t stream);

/* ---- MUL_MAT --------------------------------------------------------------------------------------------------- */

/*
 * dst[j*dst_row_stride + i] = sum_c W[i][c] * x[j*x_row_stride + c],   i < m, j < n
 * (ggml_mul_mat semantics, ggml/src/ggml.c:2722-2747; weights "src0" = m rows of k, activations "src1" = n rows of k,
 * dst = n rows of m).  Activations are quantized on the fly to q8_1 exactly as b200_quantize_q8_1 does.
 * Replaces ggml_cuda_mul_mat for quantized src0 (ggml/src/ggml-cuda/ggml-cuda.cu:1877-1938) and what it dispatches
 * to: quantize_row_q8_1_cuda + mul_mat_vec_q (mmvq.cu:130-206,423-511) for small n, quantize_mmq_q8_1_cuda +
 * mul_mat_q (+ stream-k fixup) (mmq.cuh:2516-2762, mmq.cu:3-97) for large n.
 *
 *   W            device, quantized, m rows; row stride w_row_stride bytes (0 => packed). Must be 16-byte aligned for
 *                Q4_K, 8 for IQ4_XS, 2 otherwise.
 *   x            device f32, n rows of k, row stride x_row_stride elements
 *   dst          device f32, n rows of m, row stride dst_row_stride elements
 *   workspace    device scratch of at least b200_mul_mat_workspace_size(type, m, n, k) bytes (may be NULL if that is 0)
 */
B200_API size_t b200_mul_mat_workspace_size(int type, int64_t m, int64_t n, int64_t k);
B200_API int    b200_mul_mat(int type, const void * W, int64_t w_row_stride, const float * x, int64_t x_row_stride,
                             float * dst, int64_t dst_row_stride, int64_t m, int64_t n, int64_t k, void * workspace,
                             size_t workspace_bytes, b200_stream_t stream);

/*
 * b200_mul_mat for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as b200_mul_mat's batch <= 8 path (activation quantizer + GEMV kernel), minus the
 * workspace: the q8_1 activation scratch is a __device__ array inside the library of B200_NOALLOC_SCRATCH_BYTES, which
 * every GPU that loads the library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree).
 * Limits: n <= 8, k <= 32768 and n * (k + 12 * k/32, sections rounded to 16) <= B200_NOALLOC_SCRATCH_BYTES; anything
 * larger returns B200_ERR_TOO_LARGE (text in b200_last_cuda_error()).
 * All calls on one GPU share the scratch and it keeps the last call's contents (nothing in it needs to start at zero:
 * each call's quantizer overwrites every byte its GEMV reads).  Calls that OVERLAP on one GPU race on it: serialize
 * them on one stream, or use b200_mul_mat with one workspace per stream.
 */
#define B200_NOALLOC_SCRATCH_BYTES (384 * 1024)
B200_API size_t b200_noalloc_scratch_size(void);            /* = B200_NOALLOC_SCRATCH_BYTES */
B200_API int b200_mul_mat_noalloc(int type, const void * W_device, int64_t w_row_stride, const float * x_device,
                                  int64_t x_row_stride, float * dst_device, int64_t dst_row_stride, int64_t m, int64_t n,
                                  int64_t k, b200_stream_t stream);

/*
 * nmat MUL_MATs that read the SAME activations x (q/k/v of one attention block, gate/up of one FFN): dst[i] = W[i] x^T.
 * x is quantized once instead of nmat times, and consecutive matrices of one type whose row counts are multiples of 8
 * are multiplied by ONE kernel launch over their concatenated rows (up to 4 per launch) -- fewer, longer launches is
 * what batch-1 decode is short of.  The reference issues quantize_row_q8_1_cuda + mul_mat_vec_q per node
 * (ggml/src/ggml-cuda/ggml-cuda.cu:1877-1938); results equal nmat separate b200_mul_mat calls up to fp32 summation order.
 * All W[i] are packed (row stride = b200_row_size); types[i] may differ; n >= 9 falls back to one b200_mul_mat per matrix.
 */
B200_API size_t b200_mul_mat_multi_workspace_size(int nmat, const int * types, const int64_t * m, int64_t n, int64_t k);
B200_API int    b200_mul_mat_multi(int nmat, const int * types, const void * const * W, const float * x,
                                   int64_t x_row_stride, float * const * dst, const int64_t * dst_row_stride,
                                   const int64_t * m, int64_t n, int64_t k, void * workspace, size_t workspace_bytes,
                                   b200_stream_t stream);

/* ---- batched MUL_MAT on the tensor cores (n > 8) ---------------------------------------------------------------- */

/*
 * For n > 8 b200_mul_mat runs an int8 tcgen05 GEMM with TMEM accumulators instead of the GEMV.  It consumes an int8
 * "prefill plane" of the weights: P[i][k] int8 with one fp32 scale per (row, 256-element chunk of k), and quantizes the
 * activations to int8 with one scale per (column, chunk) -- ggml-cpu's q8_K granularity (ggml/src/ggml-common.h:329-334).
 * Replaces quantize_mmq_q8_1_cuda + mul_mat_q + mul_mat_q_stream_k_fixup (ggml/src/ggml-cuda/quantize.cu:143-169,
 * mmq.cuh:2516-2762, mmq.cu:3-97), whose per-32 scale application after every K=32 MMA cannot keep a B200 tensor core
 * busy.  Results differ from the q8_1 path by rounding only and stay inside the harness bound (NMSE <= 5e-4 vs ggml-cpu).
 *
 * The plane depends only on the weights.  Callers that keep weights resident (the ggml plug-in's buffers) build it once
 * per tensor with b200_mmq_build_plane and pass it to b200_mul_mat_mmq; b200_mul_mat without a plane builds a transient
 * one in its workspace on every call (correct, slower).  Q8_0 keeps TWO planes (value + residual, ~15 bits) and runs the
 * GEMM twice, because one int8 per-256 plane would discard the dynamic range of its per-32 scales.
 */
B200_API size_t b200_mmq_plane_size(int type, int64_t m, int64_t k);
B200_API int    b200_mmq_build_plane(int type, const void * W, int64_t w_row_stride, void * plane, size_t plane_bytes,
                                     int64_t m, int64_t k, b200_stream_t stream);
B200_API size_t b200_mul_mat_mmq_workspace_size(int64_t m, int64_t n, int64_t k);
B200_API int    b200_mul_mat_mmq(int type, const void * plane, const float * x, int64_t x_row_stride, float * dst,
                                 int64_t dst_row_stride, int64_t m, int64_t n, int64_t k, void * workspace,
                                 size_t workspace_bytes, b200_stream_t stream);
/* Several prepared planes of one k against the SAME activations (q/k/v, gate/up): x is quantized to int8 once and the
 * m-tiles of up to four single-plane matrices share one GEMM grid, so a k/v projection's eight tiles do not hold a launch
 * of their own.  Same results as one b200_mul_mat_mmq per matrix (bit-identical: tiles are independent). */
B200_API int    b200_mul_mat_mmq_multi(int nmat, const int * types, const void * const * planes, const float * x,
                                       int64_t x_row_stride, float * const * dst, const int64_t * dst_row_stride,
                                       const int64_t * m, int64_t n, int64_t k, void * workspace,
                                       size_t workspace_bytes, b200_stream_t stream);
/* smallest n that b200_mul_mat sends to the tensor-core path (MMVQ_MAX_BATCH_SIZE + 1, ggml-cuda/mmvq.cuh:3) */
B200_API int    b200_mmq_min_batch(void);
/* 1 when a caller that already HOLDS a prepared plane of this type should use it for a batch of n columns: from
 * b200_mmq_min_batch() on for every type, and from 5 columns on for Q4_K / Q6_K / IQ4_XS (the GEMV needs two passes above
 * 4 columns).  The reference switches kernels at a fixed batch of 8 (MMVQ_MAX_BATCH_SIZE, mmvq.cuh:3). */
B200_API int    b200_mmq_use_cached_plane(int type, int64_t n);
/* elements of k that share one scale on the tensor-core path (plane and activation quantization granularity; a
 * build-time constant, 256 or 512): what a restatement of the path's arithmetic has to use (tests/qref.py:mul_mat_plane) */
B200_API int    b200_mmq_chunk(void);

/*
 * Same product with HOST activations and a HOST result: the activations are copied host->device, the matmul runs, and
 * dst is copied device->host, all stream-ordered on `stream` (the caller synchronizes the stream before reading dst).
 * This is what the ggml scheduler does around a MUL_MAT node whose inputs live in a CPU buffer: copy the split inputs
 * to the backend (ggml_backend_sched_compute_splits, ggml/src/ggml-backend.cpp:1363-1396, landing in
 * ggml_backend_cuda_set_tensor_async, ggml/src/ggml-cuda/ggml-cuda.cu:2386-2393), graph_compute, then
 * ggml_backend_cuda_get_tensor_async (:2395-2402).  Weights stay device-resident as in a ggml weight buffer
 * (set once through buffer.set_tensor, ggml-cuda.cu:576-582).  bench.py's `e2e` figure is measured through this call.
 *   x_host / dst_host   host memory, packed [n][k] / [n][m].  PINNED (page-locked, UVA-mapped) buffers are not staged
 *                       at all for n <= 8: the activation quantizer reads x_host over PCIe and a copy kernel stores the
 *                       result into dst_host -- a kernel boundary instead of a copy node each way (bench.py e2e:
 *                       3.90 -> 1.87 ms per token).  Pageable buffers, and B200_HOST_ZEROCOPY=0, use cudaMemcpyAsync.
 *   workspace           device scratch of at least b200_mul_mat_host_workspace_size() bytes, 256-byte aligned
 */
B200_API size_t b200_mul_mat_host_workspace_size(int type, int64_t m, int64_t n, int64_t k);
B200_API int    b200_mul_mat_host(int type, const void * W, int64_t w_row_stride, const float * x_host,
                                  float * dst_host, int64_t m, int64_t n, int64_t k, void * workspace,
                                  size_t workspace_bytes, b200_stream_t stream);

/* Several matrices against one HOST activation block (b200_mul_mat_multi with the copies of b200_mul_mat_host): x is
 * copied to the device once, and host results that lie back to back (dst_host[i+1] == dst_host[i] + n * m[i]) come back
 * in one copy.  nmat <= 64. */
B200_API size_t b200_mul_mat_multi_host_workspace_size(int nmat, const int * types, const int64_t * m, int64_t n,
                                                       int64_t k);
B200_API int    b200_mul_mat_multi_host(int nmat, const int * types, const void * const * W, const float * x_host,
                                        float * const * dst_host, const int64_t * m, int64_t n, int64_t k,
                                        void * workspace, size_t workspace_bytes, b200_stream_t stream);

/* ---- MUL_MAT_ID ------------------------------------------------------------------------------------------------ */

/*
 * A SEQUENCE of MUL_MAT groups at batch 1 (the decode token) as ONE persistent kernel launch -- replaces the per-node
 * launch loop of the decode graph (ggml_cuda_compute_forward -> ggml_cuda_mul_mat -> quantize_row_q8_1_cuda +
 * mul_mat_vec_q per node, ggml/src/ggml-cuda/ggml-cuda.cu:2646-2706, 1877-1938), which the reference amortizes with
 * CUDA-graph capture (ggml-cuda.cu:2470-2781).
 * A group = up to 4 weight matrices (any mix of the five types, packed rows) multiplied with the same activation vector
 * x[k] (q/k/v; gate/up), results dst[i][m[i]].  Groups run in list order.  `dependent` != 0: the group starts only after
 * every earlier group of the sequence is complete (what stream order gives consecutive b200_mul_mat calls) -- set it
 * when x, or a dst the group overwrites, is touched by an earlier group.  Groups with dependent == 0 may overlap the
 * tail of earlier groups and must not alias their dst.  Results are bit-identical to b200_mul_mat.
 * The PLAN (group list + geometry + arrival counters) lives in caller-owned device memory of b200_sequence_plan_size()
 * bytes, 16-byte aligned; build it once per set of pointers (outside graph capture), run it every token.  A plan must
 * not be run concurrently with itself.  Limits: n = 1, k <= 32768; B200_ERR_UNSUPPORTED otherwise (use b200_mul_mat_multi).
 */
typedef struct {
    const float * x_device;      /* [k] f32 */
    int64_t       k;
    int           nmat;          /* 1..4 */
    int           dependent;
    int           type[4];
    const void *  W_device[4];   /* m[i] packed rows */
    float *       dst_device[4]; /* [m[i]] f32 */
    int64_t       m[4];
} b200_seq_group;
B200_API size_t b200_sequence_plan_size(int ngroups);
B200_API int b200_sequence_plan_build(const b200_seq_group * groups, int ngroups, void * plan_device, size_t plan_bytes,
                                      b200_stream_t stream);
/*
 * The plan of ONE RANK of a split-mode-row sequence over `world` GPUs (one process per GPU) -- replaces the row-split
 * choreography of ggml_cuda_op_mul_mat for a whole decode token (ggml/src/ggml-cuda/ggml-cuda.cu:1560-1700: per matmul a
 * broadcast of the activations, peer copies of the result slices, events).  `groups` describe the WHOLE matrices (m = all
 * rows, dst = where row 0 goes) except that W points at THIS RANK's rows: rows [lo, hi) = b200_row_split(m, world, rank),
 * the reference's get_row_split rule.  Every dst lies inside this rank's symmetric allocation [peer_base[rank],
 * + span_bytes); peer_base[r], peer_sync[r] (>= 2 * ngroups ints, zeroed once) and peer_ll[r] are rank r's buffers mapped
 * into this process (b200_ipc_*).  Every group is ordered after the one before it.
 *   peer_ll != NULL ("flag in the data"): every rank has an exchange buffer of 2 * span_bytes used by one plan at a time;
 *     building a plan clears this rank's buffer (stream-ordered), and the ranks must meet at a barrier between building
 *     their plans and the first run.  A computed row is stored in the peers as 8 bytes {value, tag of (launch,
 *     group)}; each rank polls the tags of the rows it expects, copies the values into its dst and publishes the group on a
 *     local counter -- one NVLink store latency per group, no system-scope fence, no remote atomics.  Consecutive groups
 *     may not write the same dst rows (B200_ERR_SHAPE).
 *   peer_ll == NULL: rows are stored straight into the peers' dst, groups are published with a system-scope fence and one
 *     remote atomic per rank (a round trip per group).
 * Either way each rank holds the whole dst when its b200_sequence_run returns.  All ranks run their plans the same number
 * of times.
 */
B200_API int b200_sequence_plan_build_rowsplit(const b200_seq_group * groups, int ngroups, void * plan_device,
                                               size_t plan_bytes, int world, int rank, void * const * peer_base,
                                               size_t span_bytes, int32_t * const * peer_sync, void * const * peer_ll,
                                               b200_stream_t stream);
B200_API int b200_sequence_run(void * plan_device, b200_stream_t stream);
/* *error_group = 0: fine; g > 0: group g gave up waiting (~1 s) for earlier groups or for another rank's rows; every later
 * wait of such a launch gives up at once, so its results are invalid.  Synchronizes the stream and clears the mark. */
B200_API int b200_sequence_status(void * plan_device, int * error_group, b200_stream_t stream);
/* profiling aid (tools/seq_trace.py): later runs write 8 globaltimer stamps per (group, CTA) to trace_device; NULL = off */
B200_API int b200_debug_sequence_trace(void * plan_device, void * trace_device, b200_stream_t stream);

/*
 * dst[t][s][:] = as[ids[t][s]] * b[t][s % ne11][:]        (ggml_mul_mat_id, ggml/src/ggml.c:2771-2799)
 * Entirely on the device: the expert ids are read by the kernel, there is NO device->host copy or stream sync.
 * Replaces ggml_cuda_mul_mat_id + k_copy_src1_to_contiguous + k_copy_dst_from_contiguous,
 * ggml/src/ggml-cuda/ggml-cuda.cu:1940-2130 (which copies ids to the host and synchronizes, :2008-2011).
 *
 *   as   device, n_expert matrices of m rows x k, expert stride expert_stride bytes (0 => m * row_size)
 *   b    device f32 [n_tokens][ne11][k]; token stride b_token_stride, slot stride b_slot_stride (elements)
 *   ids  device int32 [n_tokens][>= n_used]; row stride ids_row_stride elements
 *   dst  device f32 [n_tokens][n_used][m]; token stride dst_token_stride, slot stride dst_slot_stride (elements)
 */
B200_API size_t b200_mul_mat_id_workspace_size(int type, int64_t m, int64_t k, int64_t n_expert, int64_t n_used,
                                               int64_t n_tokens);
B200_API int    b200_mul_mat_id(int type, const void * as, int64_t expert_stride, const float * b,
                                int64_t b_token_stride, int64_t b_slot_stride, int64_t ne11, const int32_t * ids,
                                int64_t ids_row_stride, float * dst, int64_t dst_token_stride,
                                int64_t dst_slot_stride, int64_t m, int64_t k, int64_t n_expert, int64_t n_used,
                                int64_t n_tokens, void * workspace, size_t workspace_bytes, b200_stream_t stream);

/* ---- multi-GPU row split ---------------------------------------------------------------------------------------- */

/*
 * MUL_MAT_ID for BATCHES on the tensor cores -- replaces the host loop of ggml_cuda_mul_mat_id for ne12 > 1
 * (ggml/src/ggml-cuda/ggml-cuda.cu:2008-2011: ids copied to the host + stream synchronize; :2059-2118: per expert gather
 * (k_copy_src1_to_contiguous), GEMM, scatter).  Here a device-side plan sorts the (token, slot) pairs by expert, the
 * activation quantizer gathers rows in that order, ONE persistent tcgen05 GEMM walks the plan's work items over the stacked
 * planes of all experts and scatters each column to its dst row.  No host synchronization; ids stay on the device.
 * `plane` = b200_mmq_build_plane over the n_expert * m rows of the packed expert tensor (b200_mmq_plane_size(type,
 * n_expert * m, k) bytes).  Needs m % 128 == 0.  Same argument meaning as b200_mul_mat_id otherwise.
 * b200_mul_mat_id_prefers_mmq: 1 when the batch is large enough for this path to win over the GEMV (a few pairs per expert).
 */
B200_API int    b200_mul_mat_id_prefers_mmq(int64_t m, int64_t n_expert, int64_t n_used, int64_t n_tokens);
B200_API size_t b200_mul_mat_id_mmq_workspace_size(int type, int64_t m, int64_t k, int64_t n_expert, int64_t n_used,
                                                   int64_t n_tokens);
B200_API int b200_mul_mat_id_mmq(int type, const void * plane_device, const float * b_device, int64_t b_token_stride,
                                 int64_t b_slot_stride, int64_t ne11, const int32_t * ids_device, int64_t ids_row_stride,
                                 float * dst_device, int64_t dst_token_stride, int64_t dst_slot_stride, int64_t m, int64_t k,
                                 int64_t n_expert, int64_t n_used, int64_t n_tokens, void * workspace_device,
                                 size_t workspace_bytes, b200_stream_t stream);

/*
 * Row range [*row_low, *row_high) of an m-row weight owned by `rank` of `world` under the reference's tensor_split
 * rule: cumulative fractions, boundaries rounded down to a multiple of 128 rows, last rank takes the remainder
 * (get_row_split / get_row_rounding, ggml/src/ggml-cuda/ggml-cuda.cu:727-753).  tensor_split == NULL => equal split.
 */
B200_API int b200_row_split(int64_t m, int world, int rank, const float * tensor_split, int64_t * row_low,
                            int64_t * row_high);

/*
 * Row-split MUL_MAT with the all-gather FUSED into the GEMV epilogue: every rank computes rows [row_low,row_high) of
 * dst and stores them straight into every rank's dst buffer over NVLink (peer_dst[r] = rank r's dst, mapped through
 * CUDA IPC / peer access); the last CTA to finish bumps every rank's arrival counter once.  No collective launch.
 * Replaces the broadcast + strided peer-copy + event choreography of ggml_cuda_op_mul_mat in split mode,
 * ggml/src/ggml-cuda/ggml-cuda.cu:1560-1700.  A plain NCCL all-gather of the slices produced by b200_mul_mat
 * (dst = local slice) is the non-fused alternative.
 *
 *   W          this rank's row slice (rows row_low..row_high of the full matrix), packed
 *   x          device f32 [n][k], identical on every rank; n <= 4, k <= 32768
 *   peer_dst   host array of `world` device pointers; entry r is rank r's full [n][m] dst (entry `rank` is local)
 *   peer_sync  host array of `world` device pointers; entry r is rank r's sync block: two int32, 8-byte aligned,
 *              zero-initialised -- [0] arrival counter, [1] scratch owned by rank r's kernels
 *   wait_sync, wait_arrivals
 *              NULL, or THIS rank's sync block of the earlier fused call(s) whose result x was computed from, and the
 *              number of arrivals that makes that result whole: the activation quantizer of this call then waits
 *              in-kernel (bounded spin on the arrival counter, the
 *              producers run on other GPUs) until that result is whole, and this call re-arms the block -- the chained
 *              form of b200_rowsplit_wait without a stream operation between the two matmuls.  A wait that gives up
 *              (~2 s) is counted per device instead of hanging the stream: b200_rowsplit_timeout_count.
 *   workspace  device scratch of b200_mul_mat_rowsplit_workspace_size(type, n, k) bytes
 *
 * Completion: every fused call of every rank counts one ARRIVAL on each rank's block, so dst on THIS rank is whole once
 * `world` arrivals are in (`world * c` when c fused calls -- independent matmuls such as q/k/v -- share the block).
 * b200_rowsplit_wait(peer_sync[rank], arrivals, stream) makes later work on the stream wait for that (stream memory
 * operations -- no spinning kernel) and re-arms the counter.  Reuse rule for a sync block: between two fused calls on the same block every rank must have issued a
 * fused call + wait on a DIFFERENT block (alternate between two blocks, or use one per matmul of the graph).
 */
B200_API size_t b200_mul_mat_rowsplit_workspace_size(int type, int64_t n, int64_t k);
B200_API int b200_mul_mat_rowsplit_fused(int type, const void * W, const float * x, int64_t x_row_stride,
                                         float * const * peer_dst, int32_t * const * peer_sync,
                                         int64_t dst_row_stride, int64_t m, int64_t row_low, int64_t row_high,
                                         int64_t n, int64_t k, int world, int rank, int32_t * wait_sync,
                                         int wait_arrivals, void * workspace, size_t workspace_bytes,
                                         b200_stream_t stream);
/*
 * Prefill under split-mode row, GEMM and exchange in ONE kernel -- replaces ggml_cuda_op_mul_mat's row-split choreography
 * for batches (ggml-cuda.cu:1560-1700: GEMM per device, then 2-D peer copies of the result slices).  This rank multiplies
 * its row slices (planes[i] = b200_mmq_build_plane over its rows, m_slice[i] of them) of up to four matrices that share
 * the activations; the persistent tensor-core kernel's epilogue stores every result element into ALL ranks' dst:
 * dst_local[i] = this rank's copy of dst + row_low (row stride dst_row_stride[i] floats, the full m), and rank r's copy
 * lies peer_delta[r] bytes from it (symmetric allocations mapped with b200_ipc_*; peer_delta[rank] = 0).  When its last
 * CTA is done the launch counts ONE arrival on every rank's sync block (peer_sync as in b200_mul_mat_rowsplit_fused), so
 * b200_rowsplit_wait(peer_sync[rank], world, stream) orders later work behind the whole dst.  All matrices of a call
 * must belong to one precision class (K-quants and IQ4_XS / Q4_0 / Q8_0): B200_ERR_UNSUPPORTED otherwise.
 * workspace: b200_mul_mat_mmq_workspace_size(max m_slice, n, k).
 */
B200_API int b200_mul_mat_mmq_multi_rowsplit(int nmat, const int * types, const void * const * planes, const float * x,
                                             int64_t x_row_stride, float * const * dst_local,
                                             const int64_t * dst_row_stride, const int64_t * m_slice, int64_t n, int64_t k,
                                             int world, int rank, const int64_t * peer_delta,
                                             int32_t * const * peer_sync, void * workspace, size_t workspace_bytes,
                                             b200_stream_t stream);
/*
 * The same exchange by the COPY ENGINES -- the reference's own way (cudaMemcpy2DAsync peer copies per device,
 * ggml-cuda.cu:1626-1691), from one process per GPU: after a local batched matmul has written this rank's row slices
 * (src_local[i] = its copy of dst + row_low, `cols[i]` result rows wide, `n` tokens high, row stride row_stride[i] floats),
 * b200_rowsplit_push copies every slice to the same place in every peer's dst (peer_delta) on one side stream per peer,
 * counts this rank's arrival on each rank's sync block behind the copies, and returns; b200_rowsplit_wait(peer_sync[rank],
 * world, stream) orders later work behind the whole dst as usual.  At most one group of pushes per peer is in flight: the
 * next push (or b200_rowsplit_push_join) makes `stream` wait for the previous one -- call the join before the slices are
 * overwritten and before ending a stream capture.  Bytes leave the SMs untouched; measured against the fused epilogue
 * stores in DESIGN.md 7.
 */
B200_API int b200_rowsplit_push(int nmat, float * const * src_local, const int64_t * row_stride, const int64_t * cols,
                                int64_t n, int world, int rank, const int64_t * peer_delta,
                                int32_t * const * peer_sync, b200_stream_t stream);
B200_API int b200_rowsplit_push_join(b200_stream_t stream);
/*
 * Stream-ordered copy of `bytes` between device memory and host memory (either direction, or device to device) --
 * what ggml_backend_cuda_set_tensor_async / _get_tensor_async do with cudaMemcpyAsync (ggml-cuda.cu:2374-2405).  When the
 * host side is pinned (mapped under UVA) and the copy is small (<= B200_COPY_KERNEL_MAX, default 1 MiB) a copy KERNEL does
 * it with 16-byte accesses over PCIe, chained to its neighbours by programmatic dependent launch: ~2 us per copy in a
 * stream of launches instead of ~8-10 us per copy node.  Anything else falls back to cudaMemcpyAsync.
 */
B200_API int b200_copy_async(void * dst, const void * src, size_t bytes, b200_stream_t stream);
/* same for a caller that KNOWS both pointers are device-accessible as they are (device memory, or host memory from
 * cudaHostAlloc under unified addressing): no pointer-attribute query per call */
B200_API int b200_copy_mapped_async(void * dst, const void * src, size_t bytes, b200_stream_t stream);
B200_API int b200_rowsplit_wait(int32_t * sync_local, int arrivals, b200_stream_t stream);
/* *count = in-kernel row-split waits on the current device that gave up since the last call (then reset to 0); the
 * matmuls behind such a wait ran on incomplete activations.  Synchronizes the stream.  No counterpart in the reference,
 * whose peer copies are ordered by events (ggml-cuda.cu:1676-1691). */
B200_API int b200_rowsplit_timeout_count(int * count, b200_stream_t stream);

/* CUDA IPC helpers so that one-process-per-GPU hosts (torchrun) can exchange peer buffers without torch types.
 * The handle names the allocation device_ptr lies in; *offset_out is device_ptr's offset inside it. */
B200_API int b200_ipc_get_handle(void * device_ptr, void * handle_out_64_bytes, int64_t * offset_out);
B200_API int b200_ipc_open_handle(const void * handle_64_bytes, int64_t offset, void ** device_ptr_out);
B200_API int b200_ipc_close_handle(void * device_ptr, int64_t offset);

#ifdef __cplusplus
}
#endif
#endif /* B200_QMM_H */
