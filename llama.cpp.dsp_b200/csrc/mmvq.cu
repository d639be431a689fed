// mmvq.cu -- batch 1..4 quantized GEMV for sm_100a: the decode hot path.
//
// Replaces (reference): mul_mat_vec_q<type,ncols_y> + vec_dot_*_q8_1 + quantize_q8_1 as used by
// ggml_cuda_op_mul_mat_vec_q (ggml/src/ggml-cuda/mmvq.cu:130-206,423-511; vecdotq.cuh:23-42,151-163,357-380,476-497,
// 527-545,604-619,671-715,763-787,1069-1133; quantize.cu:4-38), and for MUL_MAT_ID the host loop of
// ggml_cuda_mul_mat_id (ggml-cuda.cu:1994-2130).
//
// Design (B200-first, not a port):
//  * HBM-bound streaming.  One persistent CTA per SM (148 x 544 threads: 16 compute warps + 1 producer warp, 226 KB of
//    shared memory) owns a contiguous slab of weight rows and pulls it through a ring of stages (V = 2/4/8 whole rows each)
//    with 1-D bulk TMA copies (cp.async.bulk, SASS UBLKCP) that complete on per-stage "full" mbarriers; compute warps hand
//    a stage back through an "empty" mbarrier.  Raw ggml block bytes (18/34/144/210/136-byte blocks) are re-sliced out of
//    shared memory; nothing is repacked.
//  * k is split ACROSS WARPS, rows across lane groups: every lane owns one fixed 64-element unit of k for the whole
//    kernel, so its q8_1 activations live in REGISTERS (16 words + scales) and shared-memory traffic is weights only.
//    Activations are quantized once per distinct activation vector by k_quantize_q8_soa (quantize.cu, launched with
//    programmatic dependent launch in front of the GEMV); in-kernel quantizers exist behind knobs and measured slower.
//  * A team of T = ceil(k / 2048) warps covers a row; team t works ring slots s = t (mod nteams).  The last warp of a
//    team to finish a stage sums the T partials in a fixed order (deterministic results, one writer per output).
//  * int8 dot products with dp4a, fp32 scale math (gemv_core.cuh).
//  * gemvseq.cu runs the same stages as ONE launch over a whole sequence of matmul groups.
#include "gemv_core.cuh"
#include "kernels.h"

#include <cstdlib>

namespace b200 {

// ------------------------------------------------------------------------------------------------ the kernel
// Work decomposition inside a CTA:
//   lane  = one 64-element unit of k (fixed for the whole kernel: its q8_1 activations stay in registers)
//   warp  = one "chunk" of 32 consecutive units (2048 elements of k)
//   team  = the T = ceil(units_per_row / 32) warps that together cover a whole row
//   stage = V consecutive rows (V = 4 or 8), one TMA transaction, owned by team (stage % nteams)
// so teams work on different stages concurrently and every compute warp runs V unit-dots back to back, then ONE
// transposing butterfly (V values x 32 lanes -> one row sum per lane group) instead of a shuffle tree per row.
struct gemv_kargs {
    const uint8_t * W;
    int64_t         w_row_stride, w_expert_stride;
    const float *   x;
    int64_t         x_col_stride;
    const uint8_t * xq;
    float *         dst;
    int64_t         dst_col_stride;
    int             m, k;
    int             nst;          // ring depth (>= nteams)
    int             T;            // warps per team
    int             nteams;       // teams per CTA
    int             row_pitch;    // bytes between rows inside a stage
    int             stage_bytes;  // V * row_pitch + slack, multiple of 128
    int             act_bytes;    // > 0: the CTA quantizes x into this much shared memory (fused quantizer), 128-aligned
    const int32_t * ids;
    int64_t         ids_row_stride;
    int             n_used, ne11, n_expert;
    int64_t         x_token_stride, x_slot_stride, dst_token_stride, dst_slot_stride;
    // grouped MUL_MAT_ID (it_hdr != nullptr): blockIdx.y = work item = one expert x up to NCOLS (token, slot) pairs that
    // chose it (k_mmid_plan); it_hdr[0] = number of items, it_pairs[item][c] = pair index or -1
    const int32_t * it_hdr;
    const int32_t * it_expert;
    const int32_t * it_pairs;
    int             world, rank;
    int64_t         row_offset;
    int             evict_first;  // weight loads carry the L2 evict_first policy: weights are streamed once per token, the
                                  // activations and the q8 scratch they would push out of L2 are re-read by every CTA
                                  // (+1.5-2 % on the 8B token, profiles/r01/decode_ab_v2.log; B200_GEMV_EVICT_FIRST=0: off)
    int             dbg;          // profiling aid: 2 = skip the TMA loads, 8 = skip the dot products
    unsigned long long * trace;   // profiling aid: 8 globaltimer stamps per CTA (tools/gemv_trace.py), else nullptr
};
__device__ __forceinline__ void trace_stamp(const gemv_kargs & a, int i) {
    if (a.trace) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.trace[((size_t) blockIdx.y * gridDim.x + blockIdx.x) * 8 + i] = t;
    }
}
struct gemv_peers {
    float *   dst[16];
    int32_t * flag[16];
    int32_t * local;
    int32_t * rearm;
};

// Several weight matrices of one type and row length multiplied with the SAME activations in one launch (q/k/v, gate/up):
// the launch treats them as one matrix of a.m = sum(m_i) rows; matrices 1.. are described here, matrix 0 by gemv_kargs.
struct gemv_multi {
    int             nmat;         // 1 = plain launch
    int             mstart[3];    // first virtual row of matrix 1, 2, 3
    const uint8_t * W[3];
    float *         dst[3];
    int64_t         dst_col_stride[3];
};
__device__ __forceinline__ void gemv_locate(const gemv_multi & mm, int & r, const uint8_t *& W, float *& dst,
                                            int64_t & dcs) {
    if (mm.nmat > 1) {
        int sub = 0;
#pragma unroll
        for (int i = 0; i < 3; i++)                             // compile-time indices: the arrays stay in param space
            if (i + 1 < mm.nmat && r >= mm.mstart[i]) W = mm.W[i], dst = mm.dst[i], dcs = mm.dst_col_stride[i], sub = mm.mstart[i];
        r -= sub;
    }
}

// Launch shapes: MINB = 2 -> 8 compute warps, grid of two CTAs per SM; 1-2 columns, rows of <= 16384 elements.
//                            A single column is compiled for THREE resident CTAs per SM (<= 72 registers, <= 74 KB of
//                            shared memory): the third slot is where the NEXT matmul's CTAs land under programmatic
//                            dependent launch, so their barrier set-up and first weight stages overlap this kernel.
//                MINB = 1 -> 3-4 columns: 8 compute warps, one CTA per SM (registers for 4 x 24 activation words);
//                            1-2 columns: 16 compute warps, one CTA per SM (rows longer than 16384 elements).
template <int NCOLS, int MINB> struct gemv_shape {
    static constexpr int NW = (MINB == 2 || NCOLS > 2) ? 8 : 16;
    static constexpr int RESIDENT = MINB == 2 ? (NCOLS == 1 ? 3 : 2) : 1;
    static constexpr int THREADS = (NW + 1) * 32;
};

template <int TYPE, int NCOLS, int V, int MINB>
__global__ void __launch_bounds__(gemv_shape<NCOLS, MINB>::THREADS, gemv_shape<NCOLS, MINB>::RESIDENT)
    k_gemv(const gemv_kargs a, const gemv_peers peers, const gemv_multi mm) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int BYTES = qtraits<TYPE>::bytes;
    constexpr int BLCK  = qtraits<TYPE>::blck;
    constexpr int NW    = gemv_shape<NCOLS, MINB>::NW;          // compute warps; warp NW is the producer/epilogue warp

    const int  lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int  nst = a.nst, T = a.T, nteams = a.nteams;

    // shared memory: [q8 activations: act_bytes][stages: nst * stage_bytes][partials: nst * 2 * NCOLS * V * T floats]
    //                [full bars][empty bars]
    const uint32_t sm_stage  = smem_u32(smem) + a.act_bytes;
    float *        partial   = reinterpret_cast<float *>(smem + a.act_bytes + (size_t) nst * a.stage_bytes);
    const int      part_sz   = NCOLS * V * T;                   // floats per (slot, parity)
    const uint32_t bar_full  = smem_u32(partial) + (uint32_t) nst * 2 * part_sz * 4;
    const uint32_t bar_empty = bar_full + 8 * nst;
    int * const    arrived   = reinterpret_cast<int *>(reinterpret_cast<uint8_t *>(partial) + (size_t) nst * 2 * part_sz * 4 + 16 * nst);

    const uint8_t * W   = a.W;
    const float *   x   = a.x;
    float *         dst = a.dst;
    __shared__ float * s_dcol[4];                               // grouped MUL_MAT_ID: dst row of each column (or null)
    int pair_of[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; c++) pair_of[c] = 0;
    if (a.it_hdr) {  // grouped MUL_MAT_ID: the plan comes from k_mmid_plan, the kernel before us
        pdl_wait();
        if ((int) blockIdx.y >= a.it_hdr[0]) return;            // the grid is sized for the worst case
        W += (int64_t) a.it_expert[blockIdx.y] * a.w_expert_stride;
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const int p = a.it_pairs[blockIdx.y * 4 + c];
            pair_of[c] = p >= 0 ? p : a.it_pairs[blockIdx.y * 4];   // an empty column repeats column 0 and stores nothing
            if (threadIdx.x == c) {
                const int64_t tok = pair_of[c] / a.n_used, slot = pair_of[c] % a.n_used;
                s_dcol[c] = p >= 0 ? dst + tok * a.dst_token_stride + slot * a.dst_slot_stride : nullptr;
            }
        }
    } else if (a.ids) {  // MUL_MAT_ID: pair p = (token, slot); the ids come from the previous kernel (router top-k)
        pdl_wait();
        const int64_t p = blockIdx.y, tok = p / a.n_used, slot = p % a.n_used;
        const int     e = a.ids[tok * a.ids_row_stride + slot];
        // an id outside the expert range leaves its dst row untouched, as in the grouped plan (the reference asserts
        // 0 <= id < n_as, ggml-cuda.cu:2071); CTA-uniform, taken before any barrier is initialised
        if (a.n_expert > 0 && (e < 0 || e >= a.n_expert)) return;
        W   += (int64_t) e * a.w_expert_stride;
        x   += tok * a.x_token_stride + (slot % a.ne11) * a.x_slot_stride;
        dst += tok * a.dst_token_stride + slot * a.dst_slot_stride;
    }

    // this CTA's rows and stages
    // (several matrices: whole stages only, so that no stage straddles two matrices -- the host guarantees m_i % V == 0)
    const int  gran      = mm.nmat > 1 ? V : 1;
    const int  r_begin   = (int) ((int64_t) blockIdx.x * (a.m / gran) / gridDim.x) * gran;
    const int  r_end     = (int) ((int64_t) (blockIdx.x + 1) * (a.m / gran) / gridDim.x) * gran;
    const int  nstages   = (r_end - r_begin + V - 1) / V;
    const int  row_bytes = a.k / BLCK * BYTES;
    const bool contiguous = a.w_row_stride == row_bytes;

    if (threadIdx.x == 0) {
        trace_stamp(a, 0);                                      // CTA entry
        for (int s = 0; s < nst; s++) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
            arrived[s] = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_launch_dependents();

    if (warp == NW) {
        // ---------------------------------------------------------------------------------------- TMA + epilogue warp
        const uint64_t policy = a.evict_first ? l2_policy_evict_first() : 0;
        auto issue = [&](int st, int slot) {  // all lanes call; stage st -> ring slot `slot` (= st % nst)
            const int      r0 = r_begin + st * V, nr = min(V, r_end - r0);
            const uint32_t dstb = sm_stage + slot * a.stage_bytes;
            const uint32_t bar  = bar_full + 8 * slot;
            if (a.dbg == 2) {
                if (lane == 0) mbar_arrive(bar);
            } else if (contiguous) {
                if (lane == 0) {
                    const uint8_t * Wl = W;
                    float *         dl = nullptr;
                    int64_t         sl = 0;
                    int             rl = r0;
                    gemv_locate(mm, rl, Wl, dl, sl);
                    const uint64_t g    = (uint64_t) (Wl + (int64_t) rl * row_bytes);
                    const uint64_t g0   = g & ~(uint64_t) 15;
                    const uint32_t span = (uint32_t) (((g + (uint64_t) nr * row_bytes + 15) & ~(uint64_t) 15) - g0);
                    mbar_expect_tx(bar, span);
                    if (a.evict_first) tma_load_1d_hint(dstb, (const void *) g0, span, bar, policy);
                    else tma_load_1d(dstb, (const void *) g0, span, bar);
                }
            } else {  // strided rows: one bulk copy per row, lane i copies row i
                uint32_t span = 0;
                uint64_t g0   = 0;
                if (lane < nr) {
                    const uint64_t g = (uint64_t) (W + (int64_t) (r0 + lane) * a.w_row_stride);
                    g0   = g & ~(uint64_t) 15;
                    span = (uint32_t) (((g + (uint64_t) row_bytes + 15) & ~(uint64_t) 15) - g0);
                }
                uint32_t total = span;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
                if (lane == 0) mbar_expect_tx(bar, total);
                __syncwarp();
                if (lane < nr) tma_load_1d(dstb + lane * a.row_pitch, (const void *) g0, span, bar);
            }
        };
        const int pre = min(nst, nstages);
        for (int st = 0; st < pre; st++) issue(st, st);         // weights only: legal before the dependency wait
        pdl_wait();                                             // dst may still be read by the previous kernel
        if (lane == 0) trace_stamp(a, 5);                       // producer released by the previous grid
        int slot = 0, par = 0;
        for (int st = 0; st < nstages; st++) {
            mbar_wait(bar_empty + 8 * slot, par);               // stage st is consumed and its results are written
            if (st + nst < nstages) issue(st + nst, slot);
            if (++slot == nst) slot = 0, par ^= 1;
        }
        if (a.world > 1) {
            __syncwarp();
            if (lane == 0) {
                // every row of this CTA is stored in every peer; the "empty" arrivals (release.cta) ordered those
                // stores before this lane, so ONE system-scope fence per CTA publishes them.  Each CTA then adds 1 to
                // every rank's arrival counter and the last one tops the launch up to B200_ARRIVAL_UNIT, so a rank's
                // launch always counts UNIT whatever its grid size; the consumer side is the next call's quantizer
                // (wait_sync) or a stream wait-value (b200_rowsplit_wait).
                asm volatile("fence.acq_rel.sys;" ::: "memory");   // (__threadfence_system is fence.sc.sys: MEMBAR.SC + ERRBAR + CCTL.IVALL)
                const int total = (int) (gridDim.x * gridDim.y);
                int add = 1;
                if (atomicAdd(peers.local, 1) == total - 1) {
                    atomicExch(peers.local, 0);
                    if (peers.rearm) {
                        atomicExch(peers.rearm, 0);             // every CTA of this rank is past the wait on it
                        // ... and the re-arm must be in place before any peer can see this launch's signal and start
                        // sending the next arrivals to that counter
                        asm volatile("fence.acq_rel.sys;" ::: "memory");
                    }
                    add += B200_ARRIVAL_UNIT - total;
                }
                for (int r = 0; r < a.world; r++)
                    asm volatile("red.relaxed.sys.global.add.s32 [%0], %1;" ::"l"(peers.flag[r]), "r"(add) : "memory");
            }
        }
        if (lane == 0) trace_stamp(a, 6);                       // last result written
        return;
    }

    // -------------------------------------------------------------------------------------------- compute warps
    const int k32 = a.k / 32;
    const int upr = (a.k + 63) / 64;                            // units per row
    const q8_layout L = q8_soa_layout(a.k);

    pdl_wait();                                                 // x is produced by the previous kernel
    if (threadIdx.x == 0) trace_stamp(a, 1);
    if (a.act_bytes) {
        // fused quantizer: all NW compute warps (spare ones included) quantize the row into shared memory once,
        // eight lanes per 32-block, with exactly the arithmetic of the standalone kernel (actquant.cuh)
        quantize_row_q8_soa<TYPE == B200_TYPE_IQ4_XS, 8, false, true>(x, a.k, smem, L, warp, NW, lane, (uintptr_t) x % 16 == 0);
        asm volatile("bar.sync 1, %0;" ::"n"(NW * 32) : "memory");
    }
    const int team = warp / T, wt = warp - team * T;
    if (team >= nteams) return;                                 // spare warps (NW not a multiple of T)
    const int  ug   = wt * 32 + lane;                           // this lane's unit of k, fixed for the whole kernel
    const bool have = ug < upr;
    const int  ugc  = have ? ug : upr - 1;                      // lanes past the row end mirror the last unit; their
                                                                // products are dropped by a select, never multiplied
    q8_unit u[NCOLS];
    if (a.act_bytes) {
        make_unit_q8<TYPE, true>(u[0], smem, L, a.k, ugc);      // the host plans the fused quantizer for one column only
    } else if (a.xq == nullptr) {
        // in-lane quantizer: the lane quantizes exactly the 64 activations of its own unit from f32 (two q8_1 blocks;
        // Q6_K: four 16-element halves shared with the neighbouring lane) -- no quantizer kernel, no scratch, and the
        // only dependency left on the previous kernel is x itself
        const float * xc = x;
        if (a.ids) xc = x;                                       // (x was already offset to this pair's row above)
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const float * xr = xc + (int64_t) c * a.x_col_stride;
            make_unit_f32<TYPE>(u[c], xr, a.k, ugc, ((uintptr_t) xr & 15) == 0);
        }
    } else {
        const uint8_t * xq = a.xq;
        if (a.it_hdr) {
#pragma unroll
            for (int c = 0; c < NCOLS; c++) {                   // column c = activation row (token, slot % ne11) of its pair
                const size_t r = (size_t) (pair_of[c] / a.n_used) * a.ne11 + (pair_of[c] % a.n_used) % a.ne11;
                make_unit_q8<TYPE>(u[c], xq + r * L.bytes, L, a.k, ugc);
            }
        } else {
            if (a.ids) xq += (size_t) ((blockIdx.y / a.n_used) * a.ne11 + (blockIdx.y % a.n_used) % a.ne11) * L.bytes;
#pragma unroll
            for (int c = 0; c < NCOLS; c++) make_unit_q8<TYPE>(u[c], xq + (size_t) c * L.bytes, L, a.k, ugc);
        }
    }

    if (threadIdx.x == 0) trace_stamp(a, 2);                    // activations in registers
    const int      brow     = butterfly_row<V>(lane);
    const bool     bwrite   = (lane & (32 / V - 1)) == 0;
    const uint32_t row_step = contiguous ? (uint32_t) row_bytes : (uint32_t) a.row_pitch;
    const uint32_t c1       = contiguous ? 0u : (uint32_t) (a.w_row_stride & 15);
    int slot = team, par = 0;                                   // slot s is only ever used by team s % nteams
    for (int st = team; st < nstages; st += nteams) {
        const int r0 = r_begin + st * V, nr = min(V, r_end - r0);
        mbar_wait(bar_full + 8 * slot, par);
        if (threadIdx.x == 0 && st == team) trace_stamp(a, 3);  // first weights landed
        // row v starts at base + v * row_step + (its global address & 15): every bulk copy began at the 16-byte
        // boundary below its first byte (one copy per stage when rows are contiguous, one per row otherwise)
        const uint32_t base = sm_stage + slot * a.stage_bytes;
        const uint8_t * Wl  = W;
        float *         dl  = dst;
        int64_t         dcs = a.dst_col_stride;
        int             rl  = r0;
        gemv_locate(mm, rl, Wl, dl, dcs);
        const uint32_t c0   = (uint32_t) ((uint64_t) (Wl + (int64_t) rl * a.w_row_stride) & 15);
        float acc[NCOLS][V];
        if (nr == V && !(a.dbg & 8)) {
            // full stage (all but a CTA's last one): no per-row branch, so the V independent dot products are one
            // basic block and their shared-memory loads and integer chains interleave (the kernel is bound by the
            // issue latency of its 4 warps per scheduler, profiles/r02/ncu_gemv_seq_70b_*)
#pragma unroll
            for (int v = 0; v < V; v++) {
                float out[NCOLS];
                const uint32_t ra = base + v * row_step + ((c0 + v * c1) & 15u);
                unit_dot<TYPE, NCOLS>(ra, ugc, k32, u, out);
#pragma unroll
                for (int c = 0; c < NCOLS; c++) acc[c][v] = have ? out[c] : 0.f;
            }
        } else {
#pragma unroll
            for (int v = 0; v < V; v++) {
                float out[NCOLS];
#pragma unroll
                for (int c = 0; c < NCOLS; c++) out[c] = 0.f;
                if (v < nr && !(a.dbg & 8)) {                   // warp-uniform (dbg 8: loads only)
                    const uint32_t ra = base + v * row_step + ((c0 + v * c1) & 15u);
                    unit_dot<TYPE, NCOLS>(ra, ugc, k32, u, out);
                }
#pragma unroll
                for (int c = 0; c < NCOLS; c++) acc[c][v] = have ? out[c] : 0.f;
            }
        }
        float * p = partial + (size_t) (slot * 2 + par) * part_sz;
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            butterfly_rows<V>(acc[c], lane);
            if (bwrite) p[(c * V + brow) * T + wt] = acc[c][0];
        }
        // the LAST warp of the team to get here sums the T partials of every (column, row) in a fixed order
        // (deterministic results), stores them and hands the slot back to the producer
        __syncwarp();
        int last = 0;
        if (lane == 0) {
            __threadfence_block();
            last = atomicAdd(&arrived[slot], 1) == T - 1;
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last) {
            if (lane == 0) arrived[slot] = 0;
            __threadfence_block();
            const int fv = lane % V, fc = lane / V;
            if (fc < NCOLS && fv < nr) {
                const volatile float * q = p + (fc * V + fv) * T;
                float v = 0.f;
                for (int w = 0; w < T; w++) v += q[w];
                if (a.world > 1) {
                    // no fence here: the stage's "empty" arrival (release.cta) orders these stores before the
                    // producer lane's single system-scope fence at the end of the CTA (a fence per stage costs an
                    // NVLink round trip each: measured 49 us per matmul instead of ~10)
                    for (int r = 0; r < a.world; r++) peers.dst[r][fc * a.dst_col_stride + a.row_offset + r0 + fv] = v;
                    if (a.dbg & 4) __threadfence_system();      // experiment: fence per stage
                } else if (a.it_hdr) {
                    float * dp = s_dcol[fc];
                    if (dp) dp[rl + fv] = v;
                } else {
                    dl[fc * dcs + rl + fv] = v;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty + 8 * slot);
        }
        slot += nteams;
        if (slot >= nst) slot -= nst, par ^= 1;
    }
    if (threadIdx.x == 0) trace_stamp(a, 4);                    // team 0 finished its last stage
}

// ------------------------------------------------------------------------------------------------ host side
static int env_int(const char * name, int dflt) { return env_knob(name, dflt); }

// tools/gemv_trace.py, tools/layer_trace.py (b200_debug_gemv_trace): base pointer + number of slabs; one slab of
// 4096 x 8 stamps per GEMV launch, in launch order across all kernel variants
static unsigned long long * g_trace_base = nullptr;
static long long            g_trace_cap = 0, g_trace_seq = 0;
void gemv_trace_set(void * base, long long launches) {
    g_trace_base = (unsigned long long *) base, g_trace_cap = launches, g_trace_seq = 0;
}
static unsigned long long * next_trace_slab() {
    if (!g_trace_base || g_trace_seq >= g_trace_cap) return nullptr;
    return g_trace_base + (size_t) (g_trace_seq++) * 4096 * 8;
}

struct gemv_plan {
    int T, nteams, V, nst, row_pitch, stage_bytes, minb, act_bytes;
    size_t smem;
};

static int type_block_bytes(int type) {
    switch (type) {
        case B200_TYPE_Q4_0: return 18;
        case B200_TYPE_Q8_0: return 34;
        case B200_TYPE_Q4_K: return 144;
        case B200_TYPE_Q6_K: return 210;
        case B200_TYPE_IQ4_XS: return 136;
    }
    return 0;
}
static int type_block_elems(int type) { return type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 ? 32 : 256; }

static constexpr size_t SMEM_3CTA = 74 * 1024;    // three CTAs (+1 KB reserved each) fit in 228 KB
static constexpr size_t SMEM_2CTA = 112 * 1024;   // two CTAs
static constexpr size_t SMEM_1CTA = 226 * 1024;
static constexpr int    WARPS_2CTA = 8, WARPS_1CTA = 16;   // compute warps (one more warp is the producer)

// Choose team size, rows per stage and ring depth.  `rows_per_cta` steers V: short slabs take V = 4 so that every
// team of the CTA gets a stage.  Returns false if the shape cannot be planned (k > 32768: T would exceed 16 warps).
static bool make_plan(int type, int64_t k, int ncols, bool contiguous, int64_t rows_per_cta, bool fused, gemv_plan & P) {
    P.act_bytes = fused ? (int) ((q8_soa_layout(k).bytes + 127) & ~127u) : 0;
    const int upr = (int) ((k + 63) / 64);
    const int T   = (upr + 31) / 32;
    if (T > (ncols > 2 ? WARPS_2CTA : WARPS_1CTA)) return false;
    P.T = T;
    // Measured (profiles/r01/gemv_sweep_*.log): one 16-warp CTA per SM with a 226 KB ring beats two or three 8-warp
    // CTAs at every Llama shape (fewer redundant activation loads, more teams per CTA), so it is the default;
    // B200_GEMV_MINB=2 selects the small-CTA shape.
    P.minb = env_int("B200_GEMV_MINB", 1);
    if (T > WARPS_2CTA || ncols > 2) P.minb = 1;
    const int row_bytes = (int) (k / type_block_elems(type)) * type_block_bytes(type);
    P.row_pitch = contiguous ? row_bytes : ((row_bytes + 15 + 15) & ~15);
    const int want = env_int("B200_GEMV_STAGES", 0);
    for (;;) {
        const int    nw     = (P.minb == 2 || ncols > 2) ? WARPS_2CTA : WARPS_1CTA;
        const size_t total  = P.minb == 2 ? ((ncols == 1 && env_int("B200_GEMV_RESIDENT", 3) == 3) ? SMEM_3CTA : SMEM_2CTA)
                                          : SMEM_1CTA;
        if (total <= (size_t) P.act_bytes + 4096) {
            if (P.minb == 1) return false;
            P.minb = 1;
            continue;
        }
        const size_t budget = total - P.act_bytes;
        P.nteams = nw / T;
        if (const int cap_t = env_int("B200_GEMV_TEAMS", 0)) P.nteams = P.nteams < cap_t ? P.nteams : cap_t;   // tuning knob
        // rows per stage: every team of the CTA should get a stage, and the ring should hold >= 3 stages
        int vfirst = 8;
        while (vfirst > 2 && (rows_per_cta < (int64_t) P.nteams * vfirst || (size_t) vfirst * P.row_pitch * 3 > budget))
            vfirst >>= 1;
        // long rows (k >= 6144: four or more warps per row, at most four teams): 4 rows per stage.  8-row stages leave each
        // team a single ring slot, 2-row stages pay the per-stage costs twice; measured 10-17 % faster on the 70B row
        // lengths (k = 8192, 28672) for Q4_K and Q6_K alike, equal at worst (profiles/r02/gemv_sweep_70b.log)
        if (T >= 4 && (size_t) 4 * P.row_pitch * 2 + 64 <= budget) vfirst = 4;
        vfirst = env_int("B200_GEMV_V", vfirst);
        for (int V = vfirst; V >= 2; V >>= 1) {
            // payload + 15 bytes of leading misalignment + 16 so word reads of a 2-byte-aligned tail stay inside
            P.stage_bytes = (V * P.row_pitch + 15 + 16 + 127) & ~127;
            const size_t per_stage = (size_t) P.stage_bytes + 2 * (size_t) ncols * V * T * 4 + 24;
            int nst = (int) (budget / per_stage);
            const int cap = want > 0 ? want : (P.nteams * 2 > 4 ? P.nteams * 2 : 4);
            if (nst > cap) nst = cap;
            if (nst >= 2 || (V == 2 && nst >= 1)) {
                int nteams = P.nteams < nst ? P.nteams : nst;   // a team needs a slot of its own ...
                nst -= nst % nteams;                            // ... and keeps it: slot s is always used by team
                                                                // s % nteams, so no warp can run a full mbarrier
                                                                // phase ahead of the slot's previous user
                P.nteams = nteams, P.nst = nst, P.V = V, P.smem = (size_t) P.act_bytes + (size_t) nst * per_stage;
                return true;
            }
        }
        if (P.minb == 1) return false;
        P.minb = 1;                                             // 2 rows do not fit in half an SM: take all of it
    }
}

int gemv_max_cols(int type, int64_t k) {
    gemv_plan P;
    for (int n = 4; n >= 1; n--)
        if (make_plan(type, k, n, true, 1 << 20, false, P)) return n;
    return 0;
}

template <int TYPE, int NCOLS, int V, int MINB>
static int launch_gemv_k(const gemv_args & g, const gemv_plan & P, int gx, cudaStream_t stream) {
    auto kern = k_gemv<TYPE, NCOLS, V, MINB>;
    B200_SMEM_ATTR_ONCE(kern, MINB == 2 ? SMEM_2CTA : SMEM_1CTA);
    gemv_kargs a;
    a.W = g.W, a.w_row_stride = g.w_row_stride, a.w_expert_stride = g.w_expert_stride;
    a.x = g.x, a.x_col_stride = g.x_col_stride, a.xq = g.xq;
    a.dst = g.dst, a.dst_col_stride = g.dst_col_stride;
    a.m = (int) g.m, a.k = (int) g.k;
    a.nst = P.nst, a.T = P.T, a.nteams = P.nteams, a.row_pitch = P.row_pitch, a.stage_bytes = P.stage_bytes;
    a.act_bytes = P.act_bytes;
    a.ids = g.ids, a.ids_row_stride = g.ids_row_stride, a.n_used = g.n_used > 0 ? g.n_used : 1;
    a.n_expert = g.n_expert;
    a.it_hdr = g.it_hdr, a.it_expert = g.it_expert, a.it_pairs = g.it_pairs;
    a.ne11 = g.ne11 > 0 ? g.ne11 : 1;
    a.x_token_stride = g.x_token_stride, a.x_slot_stride = g.x_slot_stride;
    a.dst_token_stride = g.dst_token_stride, a.dst_slot_stride = g.dst_slot_stride;
    a.world = g.world, a.rank = g.rank, a.row_offset = g.row_offset;
    a.dbg = env_int("B200_GEMV_DEBUG", 0);
    a.evict_first = env_int("B200_GEMV_EVICT_FIRST", 1);
    a.trace = nullptr;
    a.trace = next_trace_slab();
    gemv_peers peers = {};
    if (g.world > 1) {
        for (int r = 0; r < g.world && r < 16; r++) peers.dst[r] = g.peer_dst[r], peers.flag[r] = g.peer_flag[r];
        peers.local = g.local_count, peers.rearm = g.rearm;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) gx, (unsigned) ((g.ids || g.it_hdr) ? g.npairs : 1), 1);
    cfg.blockDim         = dim3((unsigned) gemv_shape<NCOLS, MINB>::THREADS, 1, 1);
    cfg.dynamicSmemBytes = P.smem;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = env_int("B200_PDL", 1);
    cfg.attrs = attr, cfg.numAttrs = 1;
    gemv_multi mm = {};
    mm.nmat = g.nmat > 1 ? g.nmat : 1;
    for (int i = 1; i < mm.nmat; i++) {
        mm.mstart[i - 1] = (int) g.mstart[i - 1], mm.W[i - 1] = g.Wm[i - 1], mm.dst[i - 1] = g.dstm[i - 1];
        mm.dst_col_stride[i - 1] = g.dcsm[i - 1];
    }
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, a, peers, mm));
    note_launch();
    return B200_OK;
}

template <int TYPE, int NCOLS>
static int launch_gemv_t(const gemv_args & g, cudaStream_t stream) {
    const int g_sm_count = sm_count_current_device();
    if (!g_sm_count) return B200_ERR_CUDA;
    const int     row_bytes  = (int) (g.k / type_block_elems(TYPE)) * type_block_bytes(TYPE);
    const bool    contiguous = g.w_row_stride == row_bytes;
    const int64_t pairs      = g.it_hdr ? g.items_expected : g.ids ? g.npairs : 1;
    // first guess of the grid (two CTAs per SM) to learn the slab length, then the plan fixes CTAs per SM
    gemv_plan P;
    int64_t gx0 = (int64_t) g_sm_count * 2;
    if (pairs > 1) gx0 = (gx0 + pairs - 1) / pairs;             // MUL_MAT_ID: the pairs share the machine
    // f32 activations (xq == nullptr): quantize inside the kernel, cooperatively into shared memory (one column) or
    // in-lane (g.inlane: any column count, no shared memory)
    const bool fused = g.xq == nullptr && !g.inlane;
    if (fused && NCOLS != 1) return B200_ERR_UNSUPPORTED;
    if (g.nmat > 1 && (!contiguous || g.ids || g.it_hdr || g.world > 1)) return B200_ERR_UNSUPPORTED;
    if (g.it_hdr && (fused || g.inlane || g.world > 1)) return B200_ERR_UNSUPPORTED;
    if (!make_plan(TYPE, g.k, NCOLS, contiguous, (g.m + gx0 - 1) / gx0, fused, P)) return B200_ERR_UNSUPPORTED;
    int64_t gx = (int64_t) g_sm_count * P.minb * env_int("B200_GEMV_WAVES", 1);
    if (pairs > 1) gx = g.it_hdr ? gx / pairs : (gx + pairs - 1) / pairs;   // grouped: stay inside one wave of CTAs
    const int64_t max_ctas = (g.m + P.V - 1) / P.V;             // at least one stage each
    if (gx > max_ctas) gx = max_ctas;
    if (gx < 1) gx = 1;
    // shapes that are never planned (two CTAs per SM with 3-4 columns) are not instantiated
    if constexpr (NCOLS <= 2) {
        if (P.minb == 2) {
            switch (P.V) {
                case 8: return launch_gemv_k<TYPE, NCOLS, 8, 2>(g, P, (int) gx, stream);
                case 4: return launch_gemv_k<TYPE, NCOLS, 4, 2>(g, P, (int) gx, stream);
                default: return launch_gemv_k<TYPE, NCOLS, 2, 2>(g, P, (int) gx, stream);
            }
        }
    }
    switch (P.V) {
        case 8: return launch_gemv_k<TYPE, NCOLS, 8, 1>(g, P, (int) gx, stream);
        case 4: return launch_gemv_k<TYPE, NCOLS, 4, 1>(g, P, (int) gx, stream);
        default: return launch_gemv_k<TYPE, NCOLS, 2, 1>(g, P, (int) gx, stream);
    }
}

template <int TYPE>
static int launch_gemv_n(const gemv_args & g, cudaStream_t stream) {
    switch (g.ncols) {
        case 1: return launch_gemv_t<TYPE, 1>(g, stream);
        case 2: return launch_gemv_t<TYPE, 2>(g, stream);
        case 3: return launch_gemv_t<TYPE, 3>(g, stream);
        case 4: return launch_gemv_t<TYPE, 4>(g, stream);
    }
    return B200_ERR_SHAPE;
}

// ------------------------------------------------------------------------------------------------ MUL_MAT_ID plan
// Groups the (token, slot) pairs of one MUL_MAT_ID by the expert they chose: item = one expert x up to 4 pairs, so an
// expert's weights are streamed once per 4 pairs instead of once per pair (the reference gathers rows per expert on the
// host after a D2H copy of ids, ggml-cuda.cu:2008-2011,2059-2118; here ids never leave the device).
// plan = [n_items, 0, 0, 0][item_expert: max_items][item_pairs: max_items x 4].  Which pairs share an item depends on
// atomic order, results do not: every column of the GEMV is computed on its own.
constexpr int MMID_MAX_EXPERTS = 1024;
int64_t mmid_max_items(int64_t npairs, int64_t n_expert) {
    const int64_t a = npairs / 4 + (n_expert < npairs ? n_expert : npairs);
    return a < npairs ? a : npairs;
}
size_t mmid_plan_bytes(int64_t npairs, int64_t n_expert) {
    if (n_expert > MMID_MAX_EXPERTS || npairs <= 0) return 0;
    return 16 + (size_t) mmid_max_items(npairs, n_expert) * 20;
}
__global__ void __launch_bounds__(1024) k_mmid_plan(const int32_t * __restrict__ ids, int64_t ids_row_stride, int npairs,
                                                    int n_used, int n_expert, int max_items, int32_t * __restrict__ hdr,
                                                    int32_t * __restrict__ item_expert, int32_t * __restrict__ item_pairs) {
    __shared__ int cnt[MMID_MAX_EXPERTS], off[MMID_MAX_EXPERTS], cur[MMID_MAX_EXPERTS];
    pdl_launch_dependents();
    pdl_wait();                                                 // ids come from the router kernels before us
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) cnt[e] = 0, cur[e] = 0;
    for (int i = threadIdx.x; i < max_items; i += blockDim.x) {
        item_expert[i] = 0;
        item_pairs[4 * i] = item_pairs[4 * i + 1] = item_pairs[4 * i + 2] = item_pairs[4 * i + 3] = -1;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
        const int e = ids[(int64_t) (p / n_used) * ids_row_stride + p % n_used];
        if (e >= 0 && e < n_expert) atomicAdd(&cnt[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
        for (int e = 0; e < n_expert; e++) off[e] = run, run += (cnt[e] + 3) / 4;
        hdr[0] = run, hdr[1] = hdr[2] = hdr[3] = 0;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
        const int e = ids[(int64_t) (p / n_used) * ids_row_stride + p % n_used];
        if (e < 0 || e >= n_expert) continue;                   // (an invalid id leaves its dst row untouched)
        const int pos = atomicAdd(&cur[e], 1), item = off[e] + pos / 4;
        item_pairs[4 * item + (pos & 3)] = p;
        item_expert[item] = e;
    }
}
int launch_mmid_plan(const int32_t * ids, int64_t ids_row_stride, int64_t n_tokens, int64_t n_used, int64_t n_expert,
                     void * plan, cudaStream_t stream) {
    const int64_t npairs = n_tokens * n_used, max_items = mmid_max_items(npairs, n_expert);
    int32_t * hdr = (int32_t *) plan;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim  = dim3(1, 1, 1);
    cfg.blockDim = dim3(1024, 1, 1);
    cfg.stream   = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_mmid_plan, ids, ids_row_stride, (int) npairs, (int) n_used, (int) n_expert,
                                     (int) max_items, hdr, hdr + 4, hdr + 4 + max_items));
    note_launch();
    return B200_OK;
}

int launch_gemv(const gemv_args & g, cudaStream_t stream) {
    if (g.m == 0 || g.ncols == 0) return B200_OK;
    switch (g.type) {
        case B200_TYPE_Q4_0:   return launch_gemv_n<B200_TYPE_Q4_0>(g, stream);
        case B200_TYPE_Q8_0:   return launch_gemv_n<B200_TYPE_Q8_0>(g, stream);
        case B200_TYPE_Q4_K:   return launch_gemv_n<B200_TYPE_Q4_K>(g, stream);
        case B200_TYPE_Q6_K:   return launch_gemv_n<B200_TYPE_Q6_K>(g, stream);
        case B200_TYPE_IQ4_XS: return launch_gemv_n<B200_TYPE_IQ4_XS>(g, stream);
    }
    return B200_ERR_TYPE;
}

}  // namespace b200
