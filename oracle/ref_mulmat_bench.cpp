// oracle/ref_mulmat_bench.cpp -- TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Times the UNMODIFIED reference CPU implementation of the hot path (ggml-cpu: ggml_compute_forward_mul_mat,
// ggml/src/ggml-cpu/ggml-cpu.c:1266, reached through ggml_backend_graph_compute, ggml/src/ggml-backend.cpp:327-335)
// on this host's cores.  It only uses the reference's PUBLIC C API (ggml.h, ggml-backend.h) and is linked against
// the reference libraries built by oracle/Makefile.ref into oracle/_ref/.  bench.py runs it for the `cpu_baseline`
// object and for `--impl reference`.
//
//   ref-mulmat-bench --threads T --steps K --warmup W --n N --copies C  type:m:k:count [type:m:k:count ...]
//
// One "step" = one graph holding, for every spec, `count` MUL_MAT nodes of W[type, m x k] (cycling over `copies`
// distinct weight matrices per spec so that the weights do not stay cache-resident) against f32 activations [k x N].
// Weights are random VALID blocks written directly in the quantized format (random payload bytes, sane fp16 scales),
// the same generator as tests/qref.py:random_blocks.  Prints one JSON line.
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

struct spec {
    ggml_type type;
    int64_t   m, k;
    int       count;
};

static ggml_type type_by_name(const std::string & s) {
    if (s == "q4_0") return GGML_TYPE_Q4_0;
    if (s == "q8_0") return GGML_TYPE_Q8_0;
    if (s == "q4_K") return GGML_TYPE_Q4_K;
    if (s == "q6_K") return GGML_TYPE_Q6_K;
    if (s == "iq4_xs") return GGML_TYPE_IQ4_XS;
    fprintf(stderr, "unknown type %s\n", s.c_str());
    exit(2);
}

// byte offsets of the fp16 scale fields inside a block (ggml/src/ggml-common.h:167-172,209-214,285-296,320-326,412-418)
static void fix_scales(ggml_type t, uint8_t * blk, std::mt19937 & rng) {
    std::uniform_real_distribution<float> u(0.1f, 1.0f);
    auto put = [&](int off, float v) {
        const ggml_fp16_t h = ggml_fp32_to_fp16(v);
        memcpy(blk + off, &h, 2);
    };
    const float sgn = (rng() & 1) ? 1.f : -1.f;
    switch (t) {
        case GGML_TYPE_Q4_0: put(0, sgn * u(rng) * 0.02f / 4); break;
        case GGML_TYPE_Q8_0: put(0, sgn * u(rng) * 0.02f / 64); break;
        case GGML_TYPE_Q4_K: put(0, u(rng) * 0.02f / 200); put(2, u(rng) * 0.02f / 60); break;
        case GGML_TYPE_Q6_K: put(208, sgn * u(rng) * 0.02f / 2000); break;
        case GGML_TYPE_IQ4_XS: put(0, sgn * u(rng) * 0.02f / 2000); break;
        default: break;
    }
}

int main(int argc, char ** argv) {
    int threads = (int) std::thread::hardware_concurrency(), steps = 3, warmup = 1, n = 1, copies = 2;
    std::vector<spec> specs;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() { return i + 1 < argc ? atoi(argv[++i]) : 0; };
        if (a == "--threads") threads = next();
        else if (a == "--steps") steps = next();
        else if (a == "--warmup") warmup = next();
        else if (a == "--n") n = next();
        else if (a == "--copies") copies = next();
        else {
            spec s;
            char name[32];
            long long m, k;
            int c;
            if (sscanf(a.c_str(), "%31[^:]:%lld:%lld:%d", name, &m, &k, &c) != 4) {
                fprintf(stderr, "bad spec %s\n", a.c_str());
                return 2;
            }
            s.type = type_by_name(name), s.m = m, s.k = k, s.count = c;
            specs.push_back(s);
        }
    }
    if (specs.empty() || threads < 1) return 2;

    ggml_backend_load_all();  // dlopens libggml-cpu-*.so beside this executable, best-scoring variant wins
    ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    if (!cpu) {
        fprintf(stderr, "no CPU backend found\n");
        return 3;
    }
    ggml_backend_reg_t reg = ggml_backend_dev_backend_reg(ggml_backend_get_device(cpu));
    auto set_threads = (ggml_backend_set_n_threads_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_set_n_threads");
    if (set_threads) set_threads(cpu, threads);

    size_t n_nodes = 0, n_tensors = 0;
    for (auto & s : specs) n_nodes += s.count, n_tensors += (size_t) std::min(copies, s.count) + 1 + s.count;
    ggml_init_params ip = {ggml_tensor_overhead() * (n_tensors + 16) + ggml_graph_overhead_custom(n_nodes + 16, false),
                           nullptr, true};
    ggml_context * ctx = ggml_init(ip);
    ggml_cgraph *  gf  = ggml_new_graph_custom(ctx, n_nodes + 16, false);

    struct built {
        std::vector<ggml_tensor *> w;
        ggml_tensor *              x;
    };
    std::vector<built> bs;
    double weight_bytes = 0, flops = 0;
    for (auto & s : specs) {
        built b;
        const int nc = std::min(copies, s.count);
        for (int c = 0; c < nc; c++) b.w.push_back(ggml_new_tensor_2d(ctx, s.type, s.k, s.m));
        b.x = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, s.k, n);
        for (int i = 0; i < s.count; i++) {
            ggml_tensor * out = ggml_mul_mat(ctx, b.w[i % nc], b.x);
            ggml_build_forward_expand(gf, out);
            weight_bytes += (double) ggml_nbytes(b.w[i % nc]);
            flops += 2.0 * s.m * s.k * n;
        }
        bs.push_back(b);
    }
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, cpu);
    if (!buf) {
        fprintf(stderr, "allocation failed\n");
        return 4;
    }
    std::mt19937 rng(1234);
    for (size_t si = 0; si < specs.size(); si++) {
        const size_t ts = ggml_type_size(specs[si].type);
        for (ggml_tensor * w : bs[si].w) {
            uint8_t *    p  = (uint8_t *) w->data;
            const size_t nb = ggml_nbytes(w);
            uint64_t *   q  = (uint64_t *) p;
            uint64_t     st = rng() | 1;
            for (size_t i = 0; i < nb / 8; i++) {  // xorshift payload
                st ^= st << 13, st ^= st >> 7, st ^= st << 17;
                q[i] = st;
            }
            for (size_t i = nb / 8 * 8; i < nb; i++) p[i] = (uint8_t) rng();
            for (size_t o = 0; o + ts <= nb; o += ts) fix_scales(specs[si].type, p + o, rng);
        }
        float * x = (float *) bs[si].x->data;
        std::uniform_real_distribution<float> u(-1.f, 1.f);
        for (int64_t i = 0; i < ggml_nelements(bs[si].x); i++) x[i] = u(rng);
    }

    for (int i = 0; i < warmup; i++) ggml_backend_graph_compute(cpu, gf);
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) ggml_backend_graph_compute(cpu, gf);
    const auto   t1 = std::chrono::steady_clock::now();
    const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count() / steps;
    // checksum so that the work cannot be optimised away and runs can be compared
    double cks = 0;
    for (int i = 0; i < ggml_graph_n_nodes(gf); i++) {
        ggml_tensor * t = ggml_graph_node(gf, i);
        cks += ((float *) t->data)[0];
    }
    printf("{\"ms_per_step\": %.4f, \"steps\": %d, \"warmup\": %d, \"threads\": %d, \"n\": %d, \"nodes\": %zu, "
           "\"weight_bytes_per_step\": %.0f, \"flops_per_step\": %.0f, \"weight_gbs\": %.2f, \"checksum\": %.6g}\n",
           ms, steps, warmup, threads, n, n_nodes, weight_bytes, flops, weight_bytes / ms / 1e6, cks);
    ggml_backend_buffer_free(buf);
    ggml_free(ctx);
    ggml_backend_free(cpu);
    return 0;
}
