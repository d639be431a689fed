// gemvseq.cu -- batch-1 decode as ONE persistent launch over a SEQUENCE of matmul groups (b200_sequence_*).
//
// Replaces (reference): the per-node launch loop of the decode graph -- ggml_cuda_compute_forward -> ggml_cuda_mul_mat ->
// quantize_row_q8_1_cuda + mul_mat_vec_q per MUL_MAT node (ggml/src/ggml-cuda/ggml-cuda.cu:2646-2706, 1877-1938;
// mmvq.cu:130-206; quantize.cu:4-38), which the reference amortizes with CUDA-graph capture (ggml-cuda.cu:2470-2781).
//
// Why (profiles/r01/layer_trace_v1.log): with one launch per group, 3-5 us of every 4-16 us group are a
// dependency chain (previous grid drains -> quantizer grid -> GEMV grid released -> activations loaded) during which HBM
// is idle because the next GEMV's CTAs cannot even become resident before the previous ones leave (both need the whole
// 226 KB of shared memory).  Here one CTA per SM stays resident for the whole sequence:
//  * a "group" = up to 4 weight matrices that read the same activation vector (q/k/v, gate/up), exactly the unit of
//    b200_mul_mat_multi; groups are executed in list order;
//  * warp 16 walks the (group, stage) list of its CTA and keeps issuing bulk TMA copies of weight rows into ONE byte ring
//    shared by all groups (stage number q -> mbarrier q % 32, phase q / 32; its bytes are allocated circularly, each group
//    with the stage size that suits its row length), so the next group's first
//    stages are requested while the compute warps finish the current one, wait for the other CTAs and quantize the next
//    activations: weights never depend on earlier results.  32 stage barriers are shared by all groups; a parity wait
//    is safe because the stage that used the barrier before was consumed by the same team or by an earlier group;
//  * every CTA quantizes the group's activation vector itself into shared memory (16 warps, one pass over <= 56 KB from
//    L2, the arithmetic of quantize_row_q8_1_ref -- actquant.cuh), so there is no quantizer kernel and no q8 scratch;
//  * a group marked DEPENDENT starts only after every CTA has finished all earlier groups (a release/acquire counter per
//    group in the plan: same ordering a stream gives consecutive launches).  Groups not marked dependent only need their
//    own inputs and may overlap the tail of earlier groups;
//  * with 226 KB of shared memory carved out L1 is a few KB, so the descriptor of the group in work is staged in shared
//    memory (compute warps: at the boundary; issue warp: cp.async one group ahead) and all per-stage bookkeeping is
//    incremental -- per-stage reads of the plan from global memory cost an L2 round trip each and made the first version
//    2.6x slower (profiles/r02/seq_bench_*.log);
//  * per-stage compute is the code of mmvq.cu (gemv_core.cuh): one 64-element unit of k per lane in registers, dp4a,
//    butterfly over V rows, deterministic team reduction.  Results are bit-identical to b200_mul_mat.
//  * split-mode row (b200_sequence_plan_build_rowsplit): one such launch per rank; a computed row leaves the epilogue as
//    8-byte {value, tag} stores into every peer's exchange buffer, each rank polls the tags of the rows it expects, copies
//    them into its dst and publishes the group on a LOCAL counter ("flag in the data": no system-scope fence, no remote
//    atomics); the older protocol (rows stored straight into the peers' dst, fence.acq_rel.sys + one remote add per rank)
//    is kept behind peer_ll == NULL.
// Measured (profiles/r02/seq_*.log, bench_n{2,4,8}_final.json): on ONE GPU a CUDA-graph replay of grouped
// calls is still faster (8B token 1.59 ms as one launch against 1.46 ms; 70B 10.0 against 8.2 ms) -- a group boundary
// costs 3-7 us (publish, grid-wide arrival, quantizer, first stage), every round trip of it queues behind the weight
// prefetch, and the run-time-dispatched stage loop executes ~25 % more instructions than the templated per-call kernel
// while both are issue-latency-bound.  Under row split the launches are what hurts, and this kernel is the fastest path:
// 156-159 / 219-223 / 268-274 tok/s on 2 / 4 / 8 GPUs against 118 / 142 / 157 with one fused launch per matmul.
// n = 1 only (the decode token); larger batches keep the per-group entry points.
#include "gemv_core.cuh"
#include "kernels.h"

#include <cstddef>
#include <cstdio>
#include <cstring>
#include <vector>

namespace b200 {

constexpr int SEQ_NW       = 16;                 // compute warps; warp SEQ_NW issues the TMA copies
constexpr int SEQ_THREADS  = (SEQ_NW + 1) * 32;
constexpr int SEQ_MAXSLOTS = 32;                 // stages in flight at most (barriers); a power of two
constexpr int SEQ_SMEM     = 226 * 1024;
constexpr int SEQ_CTRL     = 1920;               // barriers, counters, three staged group descriptors, peer tables, stage offsets
constexpr int SEQ_PART     = 256;                // bytes of partial sums per slot: V * T <= 64 floats
constexpr int SEQ_MAXSEG   = 4;

struct seq_seg {            // one weight matrix of a group
    const uint8_t * W;      // packed rows
    float *         dst;    // m floats
    int             m, type, row_bytes;
    int             stage0; // first stage of this matrix in the group's stage list
};
struct seq_group {
    const float * x;        // k floats
    int           k, nseg;
    int           dep;      // 1: starts only after ALL earlier groups of the sequence are complete on every CTA
    int           interleave;   // activation layout class: 1 = IQ4_XS (pair-interleaved q8), 0 = everything else
    int           T, nteams, V, nstages;
    seq_seg       seg[SEQ_MAXSEG];
    int           nseg_all, pad;
    // split mode, per matrix of the group (also those this rank owns no rows of): this rank computes rows
    // [row_low, row_low + m_own) of m_full; dst_full = where row 0 goes
    int           row_low[SEQ_MAXSEG], m_full[SEQ_MAXSEG], m_own[SEQ_MAXSEG];
    float *       dst_full[SEQ_MAXSEG];
};                          // sizeof = 256: copied to shared memory as 16 x 16 bytes
static_assert(sizeof(seq_group) == 256, "seq_group layout");
constexpr int SEQ_GROUP_U4 = sizeof(seq_group) / 16;
struct seq_header {         // first bytes of the plan (device memory)
    int ngroups, exitc, error, act_bytes;   // act_bytes: shared memory reserved for the quantized activations
    int ring_bytes, pad[3];                  // bytes of the weight ring (the rest of shared memory)
    unsigned long long * trace;              // profiling aid (tools/seq_trace.py): 8 globaltimer stamps per (group, CTA)
    unsigned long long   pad2;
    // ---- split-mode row over `world` GPUs (one process and one plan per GPU; world == 1: everything below is unused)
    // Every rank's result buffers live at the same offsets of a "symmetric" allocation; a row this rank computes is stored
    // at that offset in EVERY rank's allocation (peer_delta[r] = rank r's base - this rank's base, NVLink-mapped), so each
    // rank ends up with the whole dst -- the next group's activations.  done[] lives in a sync block that every rank maps:
    // a CTA publishes a group with one system-scope release add per rank, consumers wait for world * gridDim.x arrivals.
    int       world, rank;
    int       epoch, pad3;                   // launches of this plan so far: the counters are never reset, targets grow
    long long peer_delta[16];
    int *     peer_done[16];                 // every rank's counter array (2 * ngroups ints); [rank] = the local one
    // "flag in the data" exchange (peer_ll[0] != null): a row computed here is stored in the peers as 8 bytes {value, tag}
    // into their exchange buffers (2 x span bytes, slot of the float at byte offset o of the span = 2 * o); every rank
    // polls the tags of the rows it expects, copies the values into its dst and publishes the group on its LOCAL counter.
    // No system-scope fence and no remote atomics on the critical path: one NVLink store latency instead of a round trip.
    char *    peer_ll[16];
    char *    base_local;                    // peer_base[rank]
    unsigned long long pad4;
};
// plan = [seq_header][done: ngroups ints, padded to 16][seq_group x ngroups]
__host__ __device__ static inline size_t seq_done_bytes(int ng) { return ((size_t) ng * 4 + 15) & ~(size_t) 15; }

#ifdef __CUDACC__
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ int ld_acquire_gpu(const int * p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void seq_stamp(unsigned long long * tr, int g, int ncta, int cta, int i) {
    if (tr) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        tr[((size_t) g * ncta + cta) * 8 + i] = t;
    }
}
__device__ __forceinline__ int ld_acquire_sys(const int * p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_relaxed_sys_add(int * p, int v) {
    asm volatile("red.relaxed.sys.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(int * p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// stage `st` of group G -> (matrix, first row, rows)
__device__ __forceinline__ int seq_locate(const seq_group * __restrict__ G, int nseg, int V, int st, int & r0, int & nr) {
    int s = 0;
#pragma unroll
    for (int i = 1; i < SEQ_MAXSEG; i++)
        if (i < nseg && st >= G->seg[i].stage0) s = i;
    r0 = (st - G->seg[s].stage0) * V;
    nr = min(V, G->seg[s].m - r0);
    return s;
}
#endif

}  // namespace b200

namespace b200 {

// One stage = V rows of one matrix in shared memory at `base` (+ c0: the bulk copy began at the 16-byte boundary below
// the first row).  Every lane dots its unit with the V rows, one transposing butterfly leaves lane group r with row r's
// warp sum, written to the team's partial slots p[row * T + wt].
template <int TYPE, int V>
__device__ __forceinline__ void seq_stage(uint32_t base, uint32_t row_bytes, uint32_t c0, int nr, int ugc, bool have, int k32,
                                          const q8_unit & u, float * p, int T, int wt, int lane) {
    if constexpr (V == 1) {
        float out[1] = {0.f};
        unit_dot<TYPE, 1>(base + c0, ugc, k32, &u, out);
        const float v = warp_sum(have ? out[0] : 0.f);
        if (lane == 0) p[wt] = v;
    } else {
        float acc[V];
        if (nr == V) {                                          // full stage: one basic block, the V dot products interleave
#pragma unroll
            for (int v = 0; v < V; v++) {
                float out[1];
                unit_dot<TYPE, 1>(base + v * row_bytes + c0, ugc, k32, &u, out);
                acc[v] = have ? out[0] : 0.f;
            }
        } else {
#pragma unroll
            for (int v = 0; v < V; v++) {
                float out[1] = {0.f};
                if (v < nr) unit_dot<TYPE, 1>(base + v * row_bytes + c0, ugc, k32, &u, out);   // warp-uniform branch
                acc[v] = have ? out[0] : 0.f;
            }
        }
        butterfly_rows<V>(acc, lane);
        if ((lane & (32 / V - 1)) == 0) p[butterfly_row<V>(lane) * T + wt] = acc[0];
    }
}
template <int TYPE>
__device__ __forceinline__ void seq_stage_v(int V, uint32_t base, uint32_t row_bytes, uint32_t c0, int nr, int ugc, bool have,
                                            int k32, const q8_unit & u, float * p, int T, int wt, int lane) {
    if (V == 8) seq_stage<TYPE, 8>(base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane);
    else if (V == 4) seq_stage<TYPE, 4>(base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane);
    else if (V == 2) seq_stage<TYPE, 2>(base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane);
    else seq_stage<TYPE, 1>(base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane);
}

// a wait of this kernel gives up after ~1 s (2^31 clocks) and marks the plan; once the plan is marked every later wait
// gives up at once, so a broken exchange costs one time-out per launch, not one per group
__device__ __forceinline__ bool seq_give_up(const long long t0, const int * error_word, int spins) {
    if ((spins & 1023) == 1023 && *reinterpret_cast<const volatile int *>(error_word) != 0) return true;
    return clock64() - t0 > (1ll << 31);
}

__device__ __forceinline__ void st_release_cta_shared(uint32_t addr, int v) {
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_cta_shared(uint32_t addr) {
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

__global__ void __launch_bounds__(SEQ_THREADS, 1) k_gemv_seq(uint8_t * __restrict__ plan, int flags) {
    // flags: bit 0 = evict-first L2 policy for the weights; bits 8..15 = how a group is published (0: fence.sc + release,
    // 1: release only, 3: relaxed -- measurement only); bits 16..23 = how many stages of a group the producer may issue
    // before the compute warps have that group's activations (0: no limit)
    const int evict_first = flags & 1, pub_mode = (flags >> 8) & 255, lookahead = (flags >> 16) & 255;
    const int dbg = (flags >> 24) & 255;   // experiments (wrong results): 1 = quantize the activations of the first group only
    extern __shared__ __align__(128) uint8_t smem[];
    seq_header *      H      = reinterpret_cast<seq_header *>(plan);
    const int         ng     = H->ngroups, world = H->world, rank = H->rank;
    int *             done   = world > 1 ? H->peer_done[rank] : reinterpret_cast<int *>(plan + sizeof(seq_header));
    const seq_group * groups = reinterpret_cast<const seq_group *>(plan + sizeof(seq_header) + seq_done_bytes(ng));
    const int         act_bytes = H->act_bytes;
    const uint32_t    R = (uint32_t) H->ring_bytes;
    constexpr int     NB = SEQ_MAXSLOTS;
    unsigned long long * const trace = H->trace;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = gridDim.x, cta = blockIdx.x;

    // shared memory: [control][partial sums: 256 B per barrier][q8 activations: act_bytes][ring: R bytes]
    //   control: full[32], empty[32] mbarriers; arrived[32] ints; issued (stages the producer has issued so far);
    //   st_off[32]: where in the ring a stage's bytes start
    // Stage number q (this CTA's stages of all groups, in order) uses barrier q % 32, phase q / 32, and a piece of the ring
    // allocated circularly by the producer: stages of different groups have different sizes (rows per stage x row bytes),
    // so every group fills the ring with a stage size that suits its row length.
    const uint32_t ctrl      = smem_u32(smem);
    const uint32_t bar_full  = ctrl, bar_empty = ctrl + 8 * SEQ_MAXSLOTS;
    int * const    arrived   = reinterpret_cast<int *>(smem + 16 * SEQ_MAXSLOTS);
    const uint32_t issued_a  = ctrl + 16 * SEQ_MAXSLOTS + 4 * SEQ_MAXSLOTS;
    const uint32_t cons_a    = issued_a + 4;                    // groups whose activations the compute warps have quantized
    // With 226 KB of shared memory carved out, L1 is a few KB: a descriptor field read from global memory is an L2 round
    // trip.  The group in work is therefore staged in shared memory, one copy for the compute warps (written at the group
    // boundary) and one for the producer (which runs ahead).
    // split mode: where a result row goes in every peer, and every peer's counters (staged: read per stored row / per group)
    seq_group * const Gc  = reinterpret_cast<seq_group *>(smem + 704);
    seq_group * const Gq2 = reinterpret_cast<seq_group *>(smem + 704 + sizeof(seq_group));   // [2]: this group / the next one
    static_assert(704 + 3 * sizeof(seq_group) <= 1472, "control block layout");
    long long * const pdelta = reinterpret_cast<long long *>(smem + 1472);     // split: peer_delta, or peer_ll in LL mode
    int ** const      pdone  = reinterpret_cast<int **>(smem + 1600);
    volatile uint32_t * const st_off = reinterpret_cast<volatile uint32_t *>(smem + 1728);
    static_assert(1728 + 4 * SEQ_MAXSLOTS <= SEQ_CTRL, "control block layout");
    float * const  partials  = reinterpret_cast<float *>(smem + SEQ_CTRL);
    uint8_t * const act      = smem + SEQ_CTRL + SEQ_PART * SEQ_MAXSLOTS;
    const uint32_t  ring     = ctrl + SEQ_CTRL + SEQ_PART * SEQ_MAXSLOTS + act_bytes;

    if (threadIdx.x == 0) {
        for (int i = 0; i < SEQ_MAXSLOTS; i++) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 1);
            arrived[i] = 0;
        }
        *reinterpret_cast<volatile int *>(smem + 16 * SEQ_MAXSLOTS + 4 * SEQ_MAXSLOTS) = 0;
        *reinterpret_cast<volatile int *>(smem + 16 * SEQ_MAXSLOTS + 4 * SEQ_MAXSLOTS + 4) = 0;
        mbar_fence_init();
    }
    const bool   ll         = world > 1 && H->peer_ll[0] != nullptr;
    char * const base_local = H->base_local;
    if (threadIdx.x < 16) {
        pdelta[threadIdx.x] = ll ? (long long) H->peer_ll[threadIdx.x] : H->peer_delta[threadIdx.x];
        pdone[threadIdx.x]  = H->peer_done[threadIdx.x];
    }
    __syncthreads();
    pdl_launch_dependents();

    auto nloc_of = [&](int g, int & s0) {
        const int nstages = groups[g].nstages;
        s0 = (int) ((int64_t) cta * nstages / G);
        return (int) ((int64_t) (cta + 1) * nstages / G) - s0;
    };

    if (warp == SEQ_NW) {
        // ---------------------------------------------------------------------------------------- TMA issue warp (lane 0)
        if (lane == 0) {
            const uint64_t policy = evict_first ? l2_policy_evict_first() : 0;
            int      q = 0, freed = 0;                          // stages issued / stages known to be consumed (in order)
            uint32_t head = 0, tail = 0;                        // ring: in-flight bytes lie from tail to head, circularly
            // descriptors come through cp.async, one group ahead, so that a group switch costs no L2 round trip
            auto fetch = [&](int g) {
                const uint32_t d = smem_u32(&Gq2[g & 1]);
                const uint8_t * src = reinterpret_cast<const uint8_t *>(&groups[g]);
#pragma unroll
                for (int i = 0; i < SEQ_GROUP_U4; i++)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + 16 * i), "l"(src + 16 * i) : "memory");
                asm volatile("cp.async.commit_group;" ::: "memory");
            };
            fetch(0);
            for (int g = 0; g < ng; g++) {
                if (g + 1 < ng) {
                    fetch(g + 1);
                    asm volatile("cp.async.wait_group 1;" ::: "memory");
                } else {
                    asm volatile("cp.async.wait_group 0;" ::: "memory");
                }
                const seq_group * Gq = &Gq2[g & 1];
                int       s0;
                const int nstages = Gq->nstages;
                s0 = (int) ((int64_t) cta * nstages / G);
                const int nloc = (int) ((int64_t) (cta + 1) * nstages / G) - s0;
                if (nloc == 0) continue;
                const int V = Gq->V, nseg = Gq->nseg, s1 = s0 + nloc;
                int st = s0;
                while (st < s1) {
                    // the matrix that holds stage st, and how many of this CTA's stages lie in it
                    int seg = 0;
                    for (int i = 1; i < nseg; i++)
                        if (st >= Gq->seg[i].stage0) seg = i;
                    const int m = Gq->seg[seg].m, row_bytes = Gq->seg[seg].row_bytes;
                    const int seg_end = seg + 1 < nseg ? Gq->seg[seg + 1].stage0 : nstages;
                    const int end = seg_end < s1 ? seg_end : s1;
                    int       r0 = (st - Gq->seg[seg].stage0) * V;
                    uint64_t  ga = (uint64_t) (Gq->seg[seg].W + (int64_t) r0 * row_bytes);
                    const uint32_t step = (uint32_t) V * (uint32_t) row_bytes;
                    for (; st < end; st++, q++) {
                        if (lookahead && st - s0 >= lookahead)
                            while (ld_acquire_cta_shared(cons_a) <= g) {}
                        const int      nr   = V < m - r0 ? V : m - r0;
                        const uint64_t g0   = ga & ~(uint64_t) 15;
                        const uint32_t span = (uint32_t) (((ga + (uint64_t) nr * row_bytes + 15) & ~(uint64_t) 15) - g0);
                        // (+16: the 2-byte-aligned formats read whole words, up to 4 bytes past a row's end -- keep that
                        //  inside the stage's own piece of the ring, the last piece ends where shared memory ends)
                        const uint32_t sz   = (span + 16 + 127) & ~127u;
                        // room for sz bytes: at head, or at 0 when the end of the ring is too short; otherwise wait for the
                        // oldest stage in flight (this thread is the only waiter on the empty barriers and takes them in order)
                        for (;;) {
                            if (freed == q) { head = tail = 0; break; }
                            if (q - freed < NB) {
                                if (head > tail) {
                                    if (head + sz <= R) break;
                                    if (sz <= tail) { head = 0; break; }
                                } else if (head + sz <= tail) break;
                            }
                            mbar_wait(bar_empty + 8 * (freed & (NB - 1)), (uint32_t) ((freed / NB) & 1));
                            freed++;
                            tail = st_off[freed & (NB - 1)];   // (meaningless when nothing is in flight: handled above)
                        }
                        st_off[q & (NB - 1)] = head;
                        const uint32_t bar  = bar_full + 8 * (q & (NB - 1));
                        mbar_expect_tx(bar, span);
                        if (evict_first) tma_load_1d_hint(ring + head, (const void *) g0, span, bar, policy);
                        else tma_load_1d(ring + head, (const void *) g0, span, bar);
                        head += sz;
                        ga += step, r0 += V;
                    }
                }
            }
        }
    } else {
        // -------------------------------------------------------------------------------------------- compute warps
        pdl_wait();                                             // x and dst may still be in use by the previous kernel
        // arrivals a group's counter must have reached in THIS launch: the counters only ever grow (no reset to race with)
        const int epoch  = H->epoch;
        const int target = (epoch + 1) * (ll ? 1 : world) * G;  // (LL mode: a rank's counters only count its own CTAs)
        int qbase = 0;                                          // this CTA's stages in earlier groups
        for (int g = 0; g <= ng; g++) {
            // ---- group boundary: every compute warp is done with the previous group's stages, activations and descriptor
            named_bar_sync(1, SEQ_NW * 32);
            if (threadIdx.x == 0 && g > 0) seq_stamp(trace, g - 1, G, cta, 6);   // every warp of this CTA is done with group g - 1
            if (ll && g > 0) {
                // LL exchange: collect this CTA's share of the rows of group g - 1 that OTHER ranks computed.  Gc still
                // describes group g - 1 (every CTA stages every group).  Row r of matrix i is remote when it lies outside
                // [row_low, row_low + m); its 8-byte slot in the local exchange buffer carries the value and the tag of
                // (launch, group) -- poll until the tag is there, then the value goes into dst.
                const uint32_t tag = (uint32_t) epoch * (uint32_t) ng + (uint32_t) g;    // = tag of group g - 1 (tags start at 1)
                const int nsegp = Gc->nseg_all;
                int total = 0;
                for (int i = 0; i < nsegp; i++) total += Gc->m_full[i] - Gc->m_own[i];
                for (int v = cta * (SEQ_NW * 32) + (int) threadIdx.x; v < total; v += G * SEQ_NW * 32) {
                    int i = 0, w = v;
                    while (i + 1 < nsegp && w >= Gc->m_full[i] - Gc->m_own[i]) w -= Gc->m_full[i] - Gc->m_own[i], i++;
                    const int row = w < Gc->row_low[i] ? w : w + Gc->m_own[i];
                    float * d = Gc->dst_full[i] + row;
                    const char * slot = reinterpret_cast<const char *>(pdelta[rank]) + 2 * (reinterpret_cast<char *>(d) - base_local);
                    uint32_t val, tg;
                    const long long t0 = clock64();
                    for (int spins = 0;; spins++) {
                        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(val), "=r"(tg) : "l"(slot) : "memory");
                        if (tg == tag) break;
                        if (seq_give_up(t0, &H->error, spins)) { atomicExch(&H->error, g); break; }
                    }
                    *d = __uint_as_float(val);
                }
                named_bar_sync(1, SEQ_NW * 32);                 // this CTA's share of the remote rows is in dst
            }
            if (threadIdx.x == 0 && g > 0) {
                // the barrier ordered every warp's dst stores of group g - 1 before this thread: one release publishes this
                // CTA's rows of it (system scope and one add per rank when the rows were mirrored into the peers).  Every CTA
                // arrives for every group, in order (with or without rows in it), so "done[g - 1] has all its arrivals" means
                // that ALL groups <= g - 1 are complete everywhere.
                if (ll) {
                    red_release_gpu_add(&done[g - 1], 1);       // own rows + collected rows: a LOCAL release is all it takes
                } else if (world > 1) {
                    asm volatile("fence.acq_rel.sys;" ::: "memory");   // (not __threadfence_system: that is fence.sc, far dearer)
                    for (int r = 0; r < world; r++) red_relaxed_sys_add(pdone[r] + (g - 1), 1);   // (the fence above releases)
                } else if (pub_mode == 0) {
                    __threadfence();
                    red_release_gpu_add(&done[g - 1], 1);
                } else if (pub_mode == 1) {
                    red_release_gpu_add(&done[g - 1], 1);
                } else {
                    asm volatile("red.relaxed.gpu.global.add.s32 [%0], %1;" ::"l"(&done[g - 1]), "r"(1) : "memory");
                }
                seq_stamp(trace, g - 1, G, cta, 7);             // published
            }
            if (g == ng) {
                // split mode: leave only when every rank's rows of the last group have landed here -- whatever runs next on
                // this stream reads the whole dst
                if (world > 1 && threadIdx.x == 0) {
                    const long long t0 = clock64();
                    for (int spins = 0; ld_acquire_sys(&done[ng - 1]) - target < 0; spins++)
                        if (seq_give_up(t0, &H->error, spins)) { atomicExch(&H->error, ng); break; }
                }
                break;
            }
            int       s0;
            const int nloc = nloc_of(g, s0);
            if (nloc == 0 && !ll) continue;
            if (threadIdx.x == 0) seq_stamp(trace, g, G, cta, 0);   // boundary reached: previous group done on this CTA
            if (warp == 0) {
                const int dep = lane == 0 ? __ldcg(&groups[g].dep) : 0;
                if (lane < SEQ_GROUP_U4) reinterpret_cast<uint4 *>(Gc)[lane] = __ldcg(reinterpret_cast<const uint4 *>(&groups[g]) + lane);
                if (lane == 0 && dep && g > 0 && nloc > 0) {    // all CTAs (of all ranks) have published every earlier group
                    const long long t0 = clock64();
                    for (int spins = 0; (world > 1 ? ld_acquire_sys(&done[g - 1]) : ld_acquire_gpu(&done[g - 1])) - target < 0; spins++)
                        if (seq_give_up(t0, &H->error, spins)) {   // ~1 s: leave a mark instead of hanging the GPU
                            atomicExch(&H->error, g);
                            break;
                        }
                }
            }
            named_bar_sync(1, SEQ_NW * 32);
            if (nloc == 0) continue;                            // (LL mode: staged the descriptor for the collection above)
            if (threadIdx.x == 0) seq_stamp(trace, g, G, cta, 1);   // earlier groups complete everywhere
            const int k = Gc->k, T = Gc->T, nteams = Gc->nteams, V = Gc->V, nseg = Gc->nseg;
            // ---- quantize x into shared memory (every CTA for itself; L2-coherent loads: x may be an earlier group's dst)
            const q8_layout L = q8_soa_layout(k);
            const float *   x = Gc->x;
            const bool      vec = ((uintptr_t) x & 15) == 0;
            // (8 loads per lane in flight: k <= 16384 is ONE L2 round trip per CTA -- under the weight prefetch a round trip
            //  costs 1-2 us, profiles/r02/seq_trace_*.log)
            if ((dbg & 1) && g > 0) {
            } else if (Gc->interleave) quantize_row_q8_soa<true, 8, true, true>(x, k, act, L, warp, SEQ_NW, lane, vec);
            else                quantize_row_q8_soa<false, 8, true, true>(x, k, act, L, warp, SEQ_NW, lane, vec);
            named_bar_sync(1, SEQ_NW * 32);
            if (threadIdx.x == 0) {
                if (lookahead) st_release_cta_shared(cons_a, g + 1);
                seq_stamp(trace, g, G, cta, 2);                 // activations quantized
            }

            const int team = warp / T, wt = warp - team * T;
            if (team < nteams) {                                // (spare warps of this group's team layout skip the stages)
                const int  k32 = k / 32, upr = (k + 63) / 64;
                const int  ug = wt * 32 + lane;
                const bool have = ug < upr;
                const int  ugc = have ? ug : upr - 1;
                const int  st1 = nseg > 1 ? Gc->seg[1].stage0 : INT32_MAX, st2 = nseg > 2 ? Gc->seg[2].stage0 : INT32_MAX,
                           st3 = nseg > 3 ? Gc->seg[3].stage0 : INT32_MAX;
                q8_unit u;
                int     cur_q6 = -1;                            // register layout of u: 1 = Q6_K grouping, 0 = two 32-blocks
                int     q = qbase + team;
                for (int li = team; li < nloc; li += nteams) {
                    const int st = s0 + li;
                    const int s  = st >= st3 ? 3 : st >= st2 ? 2 : st >= st1 ? 1 : 0;
                    const seq_seg * sg = &Gc->seg[s];
                    const int type = sg->type, row_bytes = sg->row_bytes;
                    const int r0 = (st - sg->stage0) * V, nr = min(V, sg->m - r0);
                    const int q6 = type == B200_TYPE_Q6_K;
                    if (q6 != cur_q6) {
                        if (q6) make_unit_q8<B200_TYPE_Q6_K, true>(u, act, L, k, ugc);
                        else    make_unit_q8<B200_TYPE_Q4_K, true>(u, act, L, k, ugc);
                        cur_q6 = q6;
                    }
                    const int slot = q & (NB - 1);
                    if (threadIdx.x == 0 && li == team) seq_stamp(trace, g, G, cta, 5);   // activations in registers
                    // parity wait on a barrier shared by stages q, q - 32, ...: safe because stage q - 32 was consumed by this
                    // same team (nteams divides 32) or belongs to an earlier group (complete before the boundary barrier),
                    // so the barrier cannot still be in the phase before the one waited for.  (An "issued" counter published
                    // by the producer with a release store per stage made the producer the bottleneck: MEMBAR.CTA per stage.)
                    mbar_wait(bar_full + 8 * slot, (uint32_t) ((q / NB) & 1));
                    if (threadIdx.x == 0 && li == team) seq_stamp(trace, g, G, cta, 3);   // first weights of the group in hand
                    const uint32_t base = ring + st_off[slot];
                    const uint32_t c0   = (uint32_t) ((uint64_t) (sg->W + (int64_t) r0 * row_bytes) & 15);
                    float * p = partials + slot * (SEQ_PART / 4);
                    switch (type) {
                        case B200_TYPE_Q4_K:   seq_stage_v<B200_TYPE_Q4_K>(V, base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane); break;
                        case B200_TYPE_Q6_K:   seq_stage_v<B200_TYPE_Q6_K>(V, base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane); break;
                        case B200_TYPE_Q4_0:   seq_stage_v<B200_TYPE_Q4_0>(V, base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane); break;
                        case B200_TYPE_Q8_0:   seq_stage_v<B200_TYPE_Q8_0>(V, base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane); break;
                        default:               seq_stage_v<B200_TYPE_IQ4_XS>(V, base, row_bytes, c0, nr, ugc, have, k32, u, p, T, wt, lane); break;
                    }
                    // the LAST warp of the team to get here sums the T partials of every row in a fixed order
                    // (deterministic), stores them and hands the slot back to the producer
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
                        if (lane < nr) {
                            const volatile float * pq = p + lane * T;
                            float v = 0.f;
                            for (int w = 0; w < T; w++) v += pq[w];
                            float * d = sg->dst + r0 + lane;
                            *d = v;
                            if (ll) {
                                const long long o2 = 2 * (reinterpret_cast<char *>(d) - base_local);
                                const uint32_t  tg = (uint32_t) epoch * (uint32_t) ng + (uint32_t) g + 1u;
                                for (int r = 0; r < world; r++)
                                    if (r != rank)
                                        asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"((char *) pdelta[r] + o2),
                                                     "r"(__float_as_uint(v)), "r"(tg) : "memory");
                            } else {
                                for (int r = 0; r < world; r++)     // (world == 1: no iterations; peer_delta[rank] == 0 is skipped)
                                    if (r != rank && world > 1)
                                        *reinterpret_cast<float *>(reinterpret_cast<char *>(d) + pdelta[r]) = v;
                            }
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_empty + 8 * slot);
                    }
                    q += nteams;
                }
                if (threadIdx.x == 0) seq_stamp(trace, g, G, cta, 4);   // team 0's share of the group done
            }
            qbase += nloc;
        }
    }
    // ---- exit: the last CTA to leave advances the plan's epoch for the next launch
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&H->exitc, 1) == G - 1) {
            H->epoch = H->epoch + 1;                            // every CTA has read it; the next launch's targets move on
            __threadfence();
            atomicExch(&H->exitc, 0);
        }
    }
}

}  // namespace b200

// ------------------------------------------------------------------------------------------------ host side
using namespace b200;

static int seq_block_elems(int type) { return type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 ? 32 : 256; }
static int seq_align(int type) { return type == B200_TYPE_Q4_K ? 16 : type == B200_TYPE_IQ4_XS ? 8 : 2; }

// per-group geometry: the team layout and the rows per stage V.  V is the largest of 8 / 4 / 2 / 1 (4 for rows of four or
// more warps: shorter stages, more of them in flight -- the measured optimum of the per-call planner, mmvq.cu make_plan)
// such that the ring holds a stage for every team plus one (two at least), the team's partial sums fit and every team of
// the CTA gets a stage; false if two stages of one row do not fit or the row is too long for a CTA
static bool seq_geometry(seq_group & g, int ring_bytes, int ncta) {
    const int upr = (g.k + 63) / 64;
    g.T = (upr + 31) / 32;
    if (g.T > SEQ_NW) return false;
    g.nteams = 1;
    while (g.nteams * 2 * g.T <= SEQ_NW) g.nteams *= 2;       // a power of two: the kernel's barrier-parity argument needs nteams | 32
    int     rowmax = 0;
    int64_t rows = 0;
    for (int i = 0; i < g.nseg; i++) rowmax = g.seg[i].row_bytes > rowmax ? g.seg[i].row_bytes : rowmax, rows += g.seg[i].m;
    const int64_t rows_per_cta = (rows + ncta - 1) / ncta;
    auto stage_bytes = [&](int v) { return (v * rowmax + 15 + 16 + 16 + 127) & ~127; };   // = the kernel's worst-case `sz`
    const int want = g.nteams + 1 > 2 ? g.nteams + 1 : 2;
    int V = g.T >= 4 ? env_knob("B200_SEQ_V_LONG", 4) : 8;
    while (V > 1 && (ring_bytes / stage_bytes(V) < want || V * g.T > SEQ_PART / 4 || rows_per_cta < (int64_t) g.nteams * V)) V >>= 1;
    if (ring_bytes / stage_bytes(V) < 2) return false;
    g.V = V;
    if (g.nseg == 0) g.nteams = 1;
    int st = 0;
    for (int i = 0; i < g.nseg; i++) g.seg[i].stage0 = st, st += (g.seg[i].m + V - 1) / V;
    g.nstages = st;
    return true;
}

extern "C" {

size_t b200_sequence_plan_size(int ngroups) {
    if (ngroups <= 0) return 0;
    const int ng = 2 * ngroups;   // a group that mixes IQ4_XS with other types is split in two (different q8 layouts)
    return sizeof(seq_header) + seq_done_bytes(ng) + (size_t) ng * sizeof(seq_group);
}

struct seq_split_info {     // row split: which rows of every matrix this rank computes, and the exchange set-up
    int world, rank;
    void * const *    peer_base;
    size_t            span_bytes;
    int32_t * const * peer_sync;
    void * const *    peer_ll;   // null: plain mirrored stores + system-scope counters
};

static int seq_build(const b200_seq_group * groups, int ngroups, void * plan_device, size_t plan_bytes, const seq_split_info * sp,
                     cudaStream_t stream) {
    if (!groups || ngroups <= 0) return B200_ERR_SHAPE;
    if (!plan_device || plan_bytes < b200_sequence_plan_size(ngroups)) return B200_ERR_WORKSPACE;
    if ((uintptr_t) plan_device % 16 != 0) return B200_ERR_ALIGNMENT;
    const int ncta = sm_count_current_device();
    if (!ncta) return B200_ERR_CUDA;
    std::vector<seq_group> out;
    int64_t kmax = 0;
    for (int gi = 0; gi < ngroups; gi++) {
        const b200_seq_group & in = groups[gi];
        if (in.nmat < 1 || in.nmat > SEQ_MAXSEG || in.k <= 0 || in.k > 32768 || !in.x_device) return B200_ERR_SHAPE;
        if ((uintptr_t) in.x_device % 4 != 0) return B200_ERR_ALIGNMENT;
        kmax = in.k > kmax ? in.k : kmax;
        for (int cls = 0; cls < 2; cls++) {                     // 0: plain q8 layout, 1: IQ4_XS (pair-interleaved)
            seq_group g = {};
            g.x = in.x_device, g.k = (int) in.k, g.interleave = cls;
            g.dep = (in.dependent || sp) ? 1 : 0;               // row split: every group waits for the exchange of the one before
            bool has_class = false;   // a group is kept even when this rank owns no rows of it: in a row split every rank
                                      // publishes every group, so the group list is the same on all ranks
            for (int i = 0; i < in.nmat; i++) {
                const int type = in.type[i];
                if (!b200_type_supported(type)) return B200_ERR_TYPE;
                if ((type == B200_TYPE_IQ4_XS) != (cls == 1)) continue;
                has_class = true;
                if (in.k % seq_block_elems(type) != 0 || in.m[i] < 0 || in.m[i] > INT32_MAX) return B200_ERR_SHAPE;
                int64_t lo = 0, hi = in.m[i];
                if (sp) {                                       // in.m = the whole matrix; W = this rank's rows, dst = row 0
                    int st = b200_row_split(in.m[i], sp->world, sp->rank, nullptr, &lo, &hi);
                    if (st != B200_OK) return st;
                    const char * d = (const char *) in.dst_device[i];
                    if (in.m[i] > 0 && (!d || d < (const char *) sp->peer_base[sp->rank] ||
                                        d + in.m[i] * 4 > (const char *) sp->peer_base[sp->rank] + sp->span_bytes))
                        return B200_ERR_SHAPE;                  // a result outside the symmetric allocation cannot be exchanged
                    const int j = g.nseg_all++;
                    g.row_low[j] = (int) lo, g.m_full[j] = (int) in.m[i], g.m_own[j] = (int) (hi - lo), g.dst_full[j] = in.dst_device[i];
                }
                if (hi == lo) continue;
                if (!in.W_device[i] || !in.dst_device[i]) return B200_ERR_SHAPE;
                if ((uintptr_t) in.W_device[i] % seq_align(type) != 0 || (uintptr_t) in.dst_device[i] % 4 != 0) return B200_ERR_ALIGNMENT;
                seq_seg & s = g.seg[g.nseg++];
                s.W = (const uint8_t *) in.W_device[i], s.dst = in.dst_device[i] + lo, s.m = (int) (hi - lo), s.type = type;
                s.row_bytes = (int) b200_row_size(type, in.k);
            }
            if (has_class) out.push_back(g);
        }
    }
    if (out.empty()) return B200_ERR_SHAPE;
    if (sp && sp->peer_ll)                                      // an exchange slot must not be rewritten before it was collected:
        for (size_t a = 0; a + 1 < out.size(); a++)             // consecutive groups may not share result rows
            for (int i = 0; i < out[a].nseg_all; i++)
                for (int j = 0; j < out[a + 1].nseg_all; j++) {
                    const float *p0 = out[a].dst_full[i], *p1 = p0 + out[a].m_full[i];
                    const float *q0 = out[a + 1].dst_full[j], *q1 = q0 + out[a + 1].m_full[j];
                    if (p0 < q1 && q0 < p1) return B200_ERR_SHAPE;
                }
    seq_header H = {};
    H.world      = 1;
    H.ngroups    = (int) out.size();
    H.act_bytes  = (int) ((q8_soa_layout(kmax).bytes + 127) & ~127u);
    H.ring_bytes = (SEQ_SMEM - SEQ_CTRL - SEQ_PART * SEQ_MAXSLOTS - H.act_bytes) & ~127;
    for (auto & g : out)
        if (!seq_geometry(g, H.ring_bytes, ncta)) return B200_ERR_UNSUPPORTED;
    if (sp) {
        H.world = sp->world, H.rank = sp->rank, H.epoch = 0;
        H.base_local = (char *) sp->peer_base[sp->rank];
        for (int r = 0; r < sp->world; r++) {
            H.peer_delta[r] = (long long) ((const char *) sp->peer_base[r] - (const char *) sp->peer_base[sp->rank]);
            H.peer_done[r]  = sp->peer_sync[r];
            H.peer_ll[r]    = sp->peer_ll ? (char *) sp->peer_ll[r] : nullptr;
        }
    }
    const size_t done_b = seq_done_bytes(H.ngroups);
    std::vector<uint8_t> host(sizeof(H) + done_b + out.size() * sizeof(seq_group), 0);
    memcpy(host.data(), &H, sizeof(H));
    memcpy(host.data() + sizeof(H) + done_b, out.data(), out.size() * sizeof(seq_group));
    // pageable source: cudaMemcpyAsync returns once the bytes are staged, so `host` may go out of scope
    B200_CUDA_TRY(cudaMemcpyAsync(plan_device, host.data(), host.size(), cudaMemcpyHostToDevice, stream));
    return B200_OK;
}

int b200_sequence_plan_build(const b200_seq_group * groups, int ngroups, void * plan_device, size_t plan_bytes,
                             b200_stream_t stream_) {
    return seq_build(groups, ngroups, plan_device, plan_bytes, nullptr, (cudaStream_t) stream_);
}

// The plan of ONE RANK of a row split over `world` GPUs (one process per GPU).  `groups` describe the WHOLE matrices (m = all
// rows, dst = where row 0 goes) except that W points at THIS RANK's rows, which are rows [lo, hi) = b200_row_split(m, world,
// rank).  Every dst lies inside this rank's symmetric allocation [peer_base[rank], + span_bytes); peer_base[r], peer_sync[r]
// (>= 2 * ngroups ints, zeroed once) and peer_ll[r] (2 * span_bytes, zeroed, or peer_ll = NULL) are rank r's buffers as
// mapped into THIS process (b200_ipc_open_handle).  Every rank runs its plan the same number of times;
// b200_sequence_run returns only after all ranks' rows have landed in this rank's dst.
int b200_sequence_plan_build_rowsplit(const b200_seq_group * groups, int ngroups, void * plan_device, size_t plan_bytes, int world,
                                      int rank, void * const * peer_base, size_t span_bytes, int32_t * const * peer_sync,
                                      void * const * peer_ll, b200_stream_t stream_) {
    if (world < 1 || world > 16 || rank < 0 || rank >= world || !peer_base || !peer_sync) return B200_ERR_SHAPE;
    for (int r = 0; r < world; r++) {
        if (!peer_base[r] || !peer_sync[r] || (uintptr_t) peer_base[r] % 8 != 0) return B200_ERR_ALIGNMENT;
        if (peer_ll && (!peer_ll[r] || (uintptr_t) peer_ll[r] % 8 != 0)) return B200_ERR_ALIGNMENT;
    }
    seq_split_info sp = {world, rank, peer_base, span_bytes, peer_sync, world > 1 ? peer_ll : nullptr};
    if (sp.peer_ll) {
        // a new plan starts its tags at 1 again: slots of THIS rank's exchange buffer that an earlier plan tagged must not
        // look valid to it.  (The peers write into this buffer as soon as they RUN their plans: the ranks need a barrier
        // between building and the first run, which they need anyway to know that every plan exists.)
        B200_CUDA_TRY(cudaMemsetAsync(peer_ll[rank], 0, 2 * span_bytes, (cudaStream_t) stream_));
    }
    return seq_build(groups, ngroups, plan_device, plan_bytes, &sp, (cudaStream_t) stream_);
}

int b200_sequence_run(void * plan_device, b200_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t) stream_;
    if (!plan_device) return B200_ERR_WORKSPACE;
    const int ncta = sm_count_current_device();
    if (!ncta) return B200_ERR_CUDA;
    B200_SMEM_ATTR_ONCE(k_gemv_seq, SEQ_SMEM);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) ncta, 1, 1);
    cfg.blockDim         = dim3(SEQ_THREADS, 1, 1);
    cfg.dynamicSmemBytes = SEQ_SMEM;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = env_knob("B200_PDL", 1);
    cfg.attrs = attr, cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_gemv_seq, (uint8_t *) plan_device, (env_knob("B200_GEMV_EVICT_FIRST", 1) & 1) | (env_knob("B200_SEQ_PUBLISH", 1) & 255) << 8 |
                                                                   (env_knob("B200_SEQ_LOOKAHEAD", 0) & 255) << 16 | (env_knob("B200_SEQ_DEBUG", 0) & 255) << 24));
    note_launch();
    return B200_OK;
}

// profiling aid: 8 globaltimer stamps per (group, CTA) are written to `trace_device` (ngroups * sm_count * 8 u64) by every
// later run of the plan; NULL switches it off
int b200_debug_sequence_trace(void * plan_device, void * trace_device, b200_stream_t stream_) {
    if (!plan_device) return B200_ERR_SHAPE;
    B200_CUDA_TRY(cudaMemcpyAsync((uint8_t *) plan_device + offsetof(seq_header, trace), &trace_device, sizeof(void *),
                                  cudaMemcpyHostToDevice, (cudaStream_t) stream_));
    B200_CUDA_TRY(cudaStreamSynchronize((cudaStream_t) stream_));
    return B200_OK;
}

// 0 = the last runs of this plan completed normally; g > 0 = a dependent group g gave up waiting for earlier groups
// (copies 4 bytes device -> host and synchronizes the stream)
int b200_sequence_status(void * plan_device, int * error_group, b200_stream_t stream_) {
    if (!plan_device || !error_group) return B200_ERR_SHAPE;
    seq_header H;
    B200_CUDA_TRY(cudaMemcpyAsync(&H, plan_device, sizeof(H), cudaMemcpyDeviceToHost, (cudaStream_t) stream_));
    B200_CUDA_TRY(cudaStreamSynchronize((cudaStream_t) stream_));
    *error_group = H.error;
    if (H.error) {   // reported once: clear the mark so that later launches wait normally again
        const int zero = 0;
        B200_CUDA_TRY(cudaMemcpyAsync((uint8_t *) plan_device + offsetof(seq_header, error), &zero, sizeof(int), cudaMemcpyHostToDevice,
                                      (cudaStream_t) stream_));
        B200_CUDA_TRY(cudaStreamSynchronize((cudaStream_t) stream_));
    }
    return B200_OK;
}

}  // extern "C"
