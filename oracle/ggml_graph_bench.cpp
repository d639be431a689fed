// oracle/ggml_graph_bench.cpp -- MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The decode token of a Llama-shaped model THROUGH THE REFERENCE'S PUBLIC ggml API, on any registered backend:
//   host activations -> ggml_backend_tensor_set -> ggml_backend_graph_compute -> ggml_backend_tensor_get -> host results
// per group of MUL_MAT nodes that read the same activations (q/k/v | attn_output | gate/up | down | output) -- the splits
// ggml_backend_sched hands a MUL_MAT-only backend (ggml/src/ggml-backend.cpp:1355-1400: copy inputs, graph_compute).
// With GGML_BACKEND_PATH=.../libggml-b200.so and --backend B200_0 this is what a ggml caller gets from the plug-in
// (bench.py: "e2e_ggml"); with --backend CPU it is the unmodified ggml-cpu through the same calls.
// Uses only ggml.h / ggml-backend.h and links the reference libraries built by oracle/Makefile.ref.
//
// Default: the ASYNCHRONOUS forms of the three calls with one ggml_backend_synchronize per token, which is how
// ggml_backend_sched drives a backend that offers them (ggml-backend.cpp:1363-1400); --sync uses the blocking forms.
//   ggml-graph-bench --backend NAME [--threads T] --steps K --warmup W --layers L --embd E --ff F --kv KV --vocab V [--q4_0] [--sync]
// Prints one JSON line.
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <random>
#include <string>
#include <thread>
#include <vector>

static void fix_scales(ggml_type t, uint8_t * blk, std::mt19937 & rng) {
    std::uniform_real_distribution<float> u(0.1f, 1.0f);
    auto put = [&](int off, float v) {
        const ggml_fp16_t h = ggml_fp32_to_fp16(v);
        memcpy(blk + off, &h, 2);
    };
    const float sgn = (rng() & 1) ? 1.f : -1.f;
    switch (t) {
        case GGML_TYPE_Q4_0: put(0, sgn * u(rng) * 0.02f / 4); break;
        case GGML_TYPE_Q4_K: put(0, u(rng) * 0.02f / 200); put(2, u(rng) * 0.02f / 60); break;
        case GGML_TYPE_Q6_K: put(208, sgn * u(rng) * 0.02f / 2000); break;
        default: break;
    }
}
// src/llama-quant.cpp:129-131 of the reference
static bool use_more_bits(int i, int n) { return i < n / 8 || i >= 7 * n / 8 || (i - n / 8) % 3 == 2; }

struct group {
    ggml_cgraph *              gf;
    ggml_tensor *              x;
    std::vector<ggml_tensor *> dst;
};

int main(int argc, char ** argv) {
    std::string backend_name = "CPU";
    int  threads = (int) std::thread::hardware_concurrency(), steps = 8, warmup = 2;
    int  L = 32, E = 4096, F = 14336, KV = 1024, V = 128256;
    bool q4_0 = false, sync_calls = false;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--backend") backend_name = next();
        else if (a == "--threads") threads = atoi(next());
        else if (a == "--steps") steps = atoi(next());
        else if (a == "--warmup") warmup = atoi(next());
        else if (a == "--layers") L = atoi(next());
        else if (a == "--embd") E = atoi(next());
        else if (a == "--ff") F = atoi(next());
        else if (a == "--kv") KV = atoi(next());
        else if (a == "--vocab") V = atoi(next());
        else if (a == "--q4_0") q4_0 = true;
        else if (a == "--sync") sync_calls = true;
    }
    ggml_backend_load_all();   // CPU variants beside the executable + whatever GGML_BACKEND_PATH names
    ggml_backend_dev_t dev = ggml_backend_dev_by_name(backend_name.c_str());
    if (!dev) {
        fprintf(stderr, "backend device %s not registered; have:", backend_name.c_str());
        for (size_t i = 0; i < ggml_backend_dev_count(); i++) fprintf(stderr, " %s", ggml_backend_dev_name(ggml_backend_dev_get(i)));
        fprintf(stderr, "\n");
        return 3;
    }
    ggml_backend_t be = ggml_backend_dev_init(dev, nullptr);
    if (!be) return 3;
    if (ggml_backend_dev_type(dev) == GGML_BACKEND_DEVICE_TYPE_CPU) {
        auto set_threads = (ggml_backend_set_n_threads_t) ggml_backend_reg_get_proc_address(ggml_backend_dev_backend_reg(dev),
                                                                                          "ggml_backend_set_n_threads");
        if (set_threads) set_threads(be, threads);
    }
    const size_t n_mat = (size_t) L * 7 + 1, n_groups = (size_t) L * 4 + 1;
    ggml_init_params ip = {ggml_tensor_overhead() * (2 * n_mat + n_groups + 64) + n_groups * ggml_graph_overhead_custom(8, false),
                           nullptr, true};
    ggml_context * ctx = ggml_init(ip);
    std::vector<group> groups;
    std::vector<ggml_tensor *> weights;
    auto add_group = [&](int64_t k, std::initializer_list<std::pair<ggml_type, int64_t>> mats) {
        group g;
        g.gf = ggml_new_graph_custom(ctx, 8, false);
        g.x  = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, k, 1);
        ggml_set_input(g.x);
        for (auto & tm : mats) {
            ggml_tensor * w = ggml_new_tensor_2d(ctx, tm.first, k, tm.second);
            weights.push_back(w);
            ggml_tensor * d = ggml_mul_mat(ctx, w, g.x);
            ggml_set_output(d);
            ggml_build_forward_expand(g.gf, d);
            g.dst.push_back(d);
        }
        groups.push_back(g);
    };
    for (int l = 0; l < L; l++) {
        const ggml_type base = q4_0 ? GGML_TYPE_Q4_0 : GGML_TYPE_Q4_K;
        const ggml_type hi   = q4_0 ? GGML_TYPE_Q4_0 : (use_more_bits(l, L) ? GGML_TYPE_Q6_K : GGML_TYPE_Q4_K);
        add_group(E, {{base, E}, {base, KV}, {hi, KV}});
        add_group(E, {{base, E}});
        add_group(E, {{base, F}, {base, F}});
        add_group(F, {{hi, E}});
    }
    add_group(E, {{GGML_TYPE_Q6_K, V}});
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, be);
    if (!buf) {
        fprintf(stderr, "allocation failed\n");
        return 4;
    }
    ggml_backend_buffer_set_usage(buf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
    // weights: one host image per distinct (type, k, m), uploaded to every matrix of that shape
    std::mt19937 rng(1234);
    std::map<std::tuple<int, int64_t, int64_t>, std::vector<uint8_t>> images;
    double weight_bytes = 0;
    for (ggml_tensor * w : weights) {
        auto key = std::make_tuple((int) w->type, w->ne[0], w->ne[1]);
        auto & img = images[key];
        if (img.empty()) {
            img.resize(ggml_nbytes(w));
            uint64_t * q = (uint64_t *) img.data();
            uint64_t   st = rng() | 1;
            for (size_t i = 0; i < img.size() / 8; i++) st ^= st << 13, st ^= st >> 7, st ^= st << 17, q[i] = st;
            const size_t ts = ggml_type_size(w->type);
            for (size_t o = 0; o + ts <= img.size(); o += ts) fix_scales(w->type, img.data() + o, rng);
        }
        ggml_backend_tensor_set(w, img.data(), 0, img.size());
        weight_bytes += (double) img.size();
    }
    // host side: the device's pinned host buffer type when it offers one (what llama.cpp uses for inputs / outputs)
    size_t host_bytes = 0, max_k = 0;
    for (auto & g : groups) {
        max_k = std::max(max_k, (size_t) g.x->ne[0]);
        for (auto d : g.dst) host_bytes += ggml_nbytes(d);
    }
    ggml_backend_buffer_type_t hbt = ggml_backend_dev_host_buffer_type(dev);
    ggml_backend_buffer_t      hbuf = hbt ? ggml_backend_buft_alloc_buffer(hbt, host_bytes + max_k * 4 + 256) : nullptr;
    std::vector<uint8_t>       pageable;
    uint8_t * host = nullptr;
    if (hbuf) host = (uint8_t *) ggml_backend_buffer_get_base(hbuf);
    else pageable.resize(host_bytes + max_k * 4 + 256), host = pageable.data();
    float * hx = (float *) host;
    std::uniform_real_distribution<float> u(-1.f, 1.f);
    for (size_t i = 0; i < max_k; i++) hx[i] = u(rng);
    uint8_t * hout = host + ((max_k * 4 + 255) & ~(size_t) 255);

    double h2d = 0, d2h = 0;
    auto token = [&]() {
        size_t off = 0;
        h2d = d2h = 0;
        for (auto & g : groups) {
            if (sync_calls) ggml_backend_tensor_set(g.x, hx, 0, ggml_nbytes(g.x));
            else ggml_backend_tensor_set_async(be, g.x, hx, 0, ggml_nbytes(g.x));
            h2d += (double) ggml_nbytes(g.x);
            if ((sync_calls ? ggml_backend_graph_compute(be, g.gf) : ggml_backend_graph_compute_async(be, g.gf)) != GGML_STATUS_SUCCESS) exit(5);
            for (auto d : g.dst) {
                if (sync_calls) ggml_backend_tensor_get(d, hout + off, 0, ggml_nbytes(d));
                else ggml_backend_tensor_get_async(be, d, hout + off, 0, ggml_nbytes(d));
                off += ggml_nbytes(d), d2h += (double) ggml_nbytes(d);
            }
        }
        ggml_backend_synchronize(be);   // the caller reads this token's results on the host
    };
    for (int i = 0; i < warmup; i++) token();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) token();
    const auto   t1 = std::chrono::steady_clock::now();
    const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count() / steps;
    double cks = 0;
    for (size_t i = 0; i < host_bytes / 4; i += 997) cks += ((float *) hout)[i];
    printf("{\"backend\": \"%s\", \"ms_per_step\": %.4f, \"tok_s\": %.2f, \"steps\": %d, \"warmup\": %d, \"threads\": %d, "
           "\"nodes\": %zu, \"graph_compute_calls\": %zu, \"weight_bytes_per_step\": %.0f, \"weight_gbs\": %.2f, "
           "\"h2d_bytes_per_step\": %.0f, \"d2h_bytes_per_step\": %.0f, \"host_buffer\": \"%s\", \"checksum\": %.6g}\n",
           ggml_backend_dev_name(dev), ms, 1e3 / ms, steps, warmup, threads, weights.size(), groups.size(), weight_bytes,
           weight_bytes / ms / 1e6, h2d, d2h, hbuf ? ggml_backend_buft_name(hbt) : "pageable", sync_calls ? "blocking" : "async", cks);
    if (hbuf) ggml_backend_buffer_free(hbuf);
    ggml_backend_buffer_free(buf);
    ggml_free(ctx);
    ggml_backend_free(be);
    return 0;
}
