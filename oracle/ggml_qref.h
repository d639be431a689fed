/*
 * oracle/ggml_qref.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, no SIMD, no FMA contraction) of the reference's algorithm for the quantized
 * MUL_MAT / MUL_MAT_ID hot path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this.  The product (libb200qmm.so, libggml-b200.so) never links or calls it.
 *
 * Parity status: PINNED.  Every function here is checked bit-for-bit (quantize/dequantize) or to float-summation-order
 * tolerance (dot products) against the UNMODIFIED reference built from /root/reference by oracle/Makefile.ref
 * (oracle/_ref/libggml-base.so, libggml-cpu-*.so) in tests/test_oracle_vs_ref.py, and against the committed golden
 * vectors under tests/golden/ (generated from that reference build by tests/golden/make_golden.py).
 *
 * All file:line citations are relative to /root/reference/.
 */
#ifndef B200_ORACLE_GGML_QREF_H
#define B200_ORACLE_GGML_QREF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ggml type ids used on the path (ggml/include/ggml.h:350-392, enum ggml_type) */
enum {
    QREF_TYPE_F32    = 0,
    QREF_TYPE_Q4_0   = 2,
    QREF_TYPE_Q8_0   = 8,
    QREF_TYPE_Q8_1   = 9,
    QREF_TYPE_Q4_K   = 12,
    QREF_TYPE_Q6_K   = 14,
    QREF_TYPE_Q8_K   = 15,
    QREF_TYPE_IQ4_XS = 23,
};

/* elements per block / bytes per block (ggml/src/ggml-common.h:167-172,209-227,285-296,320-334,412-418) */
int    qref_blck_size(int type);
size_t qref_type_size(int type);
size_t qref_row_size(int type, int64_t k);

/* fp16 <-> fp32, round-to-nearest-even (ggml/src/ggml-impl.h:339-490) */
uint16_t qref_fp32_to_fp16(float f);
float    qref_fp16_to_fp32(uint16_t h);

/* activation quantizers: bit-exact targets */
void qref_quantize_row_q8_1(const float * x, void * y, int64_t k); /* ggml-quants.c:220-253 */
void qref_quantize_row_q8_0(const float * x, void * y, int64_t k); /* ggml-quants.c:194-218 */
void qref_quantize_row_q8_K(const float * x, void * y, int64_t k); /* ggml-quants.c:2479-2516 */
/* simple weight quantizers used only to make realistic test inputs */
void qref_quantize_row_q4_0(const float * x, void * y, int64_t k); /* ggml-quants.c:31-66 */

/* dequantizers: bit-exact targets (ggml-quants.c:255-273, 349-363, 1280-1302, 1690-1722, 2454-2476) */
void qref_dequantize_row(int type, const void * x, float * y, int64_t k);

/* y[r*k + c] for nrows rows laid out back to back (row stride = qref_row_size) */
void qref_dequantize_rows(int type, const void * x, float * y, int64_t nrows, int64_t k);

/*
 * MUL_MAT, ggml-cpu semantics (ggml/src/ggml-cpu/ggml-cpu.c:1266-1538): activations are quantized row by row to
 * the weight type's vec_dot_type (q8_0 for Q4_0/Q8_0, q8_K for Q4_K/Q6_K/IQ4_XS; table ggml-cpu.c:223-352), then
 * dst[j*m + i] = vec_dot(W row i, Xq row j) with the scalar vec_dot kernels
 * (ggml-cpu/ggml-cpu-quants.c:1912, 3597, 6667, 8415, 12665).
 *   W   : m rows of k elements, quantized, row stride row_stride_bytes (0 => tightly packed)
 *   x   : n rows of k f32 (row stride k)
 *   dst : n rows of m f32
 * nthreads <= 0 : use all hardware threads (OpenMP).
 */
void qref_mul_mat(int type, const void * W, size_t row_stride_bytes, const float * x, float * dst,
                  int64_t m, int64_t n, int64_t k, int nthreads);

/*
 * Same product with the reference's CUDA-path semantics: activations quantized to q8_1, per-32 integer dot, float
 * scaling (ggml/src/ggml-cuda/vecdotq.cuh:23-42,151-163,357-380,476-497,527-545,604-619,671-715,763-787,1069-1133);
 * accumulated in double so that it is a tight (not NMSE-loose) check of the device kernels' bit decoding.
 */
void qref_mul_mat_q8_1(int type, const void * W, size_t row_stride_bytes, const float * x, float * dst,
                       int64_t m, int64_t n, int64_t k, int nthreads);

/* exact product of the DEQUANTIZED weights with the raw f32 activations, double accumulation ("ground truth") */
void qref_mul_mat_f64(int type, const void * W, size_t row_stride_bytes, const float * x, float * dst,
                      int64_t m, int64_t n, int64_t k, int nthreads);

/*
 * MUL_MAT_ID (ggml/src/ggml.c:2771-2799; CPU reference ggml-cpu.c:1540-1718):
 *   as  : n_expert matrices, each m rows x k, expert stride expert_stride_bytes
 *   b   : [n_tokens][ne11][k] f32, ne11 in {1, n_used}  (slot s reads b[t][s % ne11])
 *   ids : [n_tokens][ids_stride] int32, first n_used entries of each row used
 *   dst : [n_tokens][n_used][m] f32
 * semantics 0 = ggml-cpu (vec_dot_type), 1 = q8_1, 2 = f64 ground truth
 */
void qref_mul_mat_id(int type, const void * as, size_t expert_stride_bytes, const float * b, const int32_t * ids,
                     int64_t ids_stride, float * dst, int64_t m, int64_t k, int64_t n_expert, int64_t n_used,
                     int64_t ne11, int64_t n_tokens, int semantics, int nthreads);

/* normalized mean squared error as tests/test-backend-ops.cpp:262-273 computes it */
double qref_nmse(const float * a, const float * b, size_t n);

#ifdef __cplusplus
}
#endif
#endif
