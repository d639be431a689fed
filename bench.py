#!/usr/bin/env python
"""bench.py -- the measurement contract for the quantized MUL_MAT hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one decode token's pass over the hot path: every quantized MUL_MAT of a Llama-shaped transformer
(7 per layer + the output head) at batch 1, weights random-init in the GGUF block formats and resident in HBM.
ONE workload for every N, so that the driver's 1 -> 8 GPU scaling compares like with like: BASELINE.json configs[4],
Llama-3-70B-shaped Q4_K_M (561 matmuls, 41.9 GB of weights per token; it fits one B200).
  * N = 1  : all rows on one GPU.  The line also carries BASELINE.json configs[1] -- Llama-3-8B-shaped Q4_K_M: tg128,
             pp512 (tensor roofline), e2e, MUL_MAT_ID -- as the sub-object "llama3_8b_q4_k_m", measured by a sub-run of
             this script (--workload llama3-8b-q4_k_m) on the same GPU.
  * N > 1  : split-mode row: every weight's output rows are sharded over the ranks (get_row_split rule) and every rank
             ends up with the whole result (fused peer stores from the GEMV epilogue; NCCL all-gather as the baseline).
             Strong scaling: the model is fixed, per-GPU work shrinks with N.
value   = tokens/s with inputs resident in HBM (device-timed with CUDA events, max over ranks).
e2e     = the same step through the host-buffer C-ABI call b200_mul_mat_host: activations start in pinned host
          memory and results end there, per matmul (what the ggml scheduler does around a MUL_MAT-only backend).
roofline= algorithmic bytes of the step's GEMV launches / their device time, against MEASURED_PEAKS.json hbm_gbs.
cpu_baseline / --impl reference = the UNMODIFIED reference ggml-cpu backend (oracle/_ref, built by
          oracle/Makefile.ref from /root/reference) on this host's cores, timed by oracle/_ref/ref-mulmat-bench.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import zlib

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

Q4_0, Q8_0, Q4_K, Q6_K, IQ4_XS = 2, 8, 12, 14, 23
TNAME = {Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q6_K: "q6_K", IQ4_XS: "iq4_xs"}
BLOCK = {Q4_0: 32, Q8_0: 32, Q4_K: 256, Q6_K: 256, IQ4_XS: 256}
TSIZE = {Q4_0: 18, Q8_0: 34, Q4_K: 144, Q6_K: 210, IQ4_XS: 136}

MODELS = {
    # name: n_layer, n_embd, n_ff, n_kv (= n_head_kv * head_dim), vocab, recipe
    "llama3-8b-q4_k_m": dict(n_layer=32, n_embd=4096, n_ff=14336, n_kv=1024, vocab=128256, recipe="q4_k_m"),
    "llama3-70b-q4_k_m": dict(n_layer=80, n_embd=8192, n_ff=28672, n_kv=1024, vocab=128256, recipe="q4_k_m"),
    "llama2-7b-q4_0": dict(n_layer=32, n_embd=4096, n_ff=11008, n_kv=4096, vocab=32000, recipe="q4_0"),
}


def use_more_bits(i, n):
    # src/llama-quant.cpp:129-131 of the reference
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def matmul_list(model):
    """[(name, ggml_type, m, k)] in graph order for one token (llm_build_llama, src/llama-model.cpp:4408).
    Q4_K_M recipe (src/llama-quant.cpp:166-168,235-236,291-297): attn_v and ffn_down are Q6_K in use_more_bits layers,
    output.weight is Q6_K, everything else Q4_K (the 70B recipe's Q5_K attn_v is replaced by Q4_K: out of scope)."""
    c = MODELS[model]
    L, e, ff, kv, V = c["n_layer"], c["n_embd"], c["n_ff"], c["n_kv"], c["vocab"]
    out = []
    for i in range(L):
        if c["recipe"] == "q4_k_m":
            hi = Q6_K if use_more_bits(i, L) else Q4_K
            base = Q4_K
        else:
            hi = base = Q4_0
        out += [("attn_q.%d" % i, base, e, e), ("attn_k.%d" % i, base, kv, e), ("attn_v.%d" % i, hi, kv, e),
                ("attn_output.%d" % i, base, e, e), ("ffn_gate.%d" % i, base, ff, e), ("ffn_up.%d" % i, base, ff, e),
                ("ffn_down.%d" % i, hi, e, ff)]
    out.append(("output", Q6_K, V, e))   # output.weight is Q6_K under both recipes (src/llama-quant.cpp:166-168)
    return out


def row_bytes(t, k):
    return k // BLOCK[t] * TSIZE[t]


def gemv_bytes(t, m, k, n):
    """algorithmic bytes of one matmul: weights + f32 activations + f32 result"""
    return m * row_bytes(t, k) + n * k * 4 + n * m * 4


# ------------------------------------------------------------------------------------------------ reference (CPU) arm
def host_threads():
    try:
        import psutil
        phys = psutil.cpu_count(logical=False) or 0
    except Exception:  # noqa: BLE001
        phys = 0
    avail = len(os.sched_getaffinity(0))
    n = min(phys, avail) if phys else avail
    return max(1, min(n, 512))


def ref_specs(model):
    agg = {}
    for _, t, m, k in matmul_list(model):
        agg[(t, m, k)] = agg.get((t, m, k), 0) + 1
    return ["%s:%d:%d:%d" % (TNAME[t], m, k, c) for (t, m, k), c in agg.items()]


def run_reference(model, steps, warmup, threads, copies=3):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref-mulmat-bench")
    if not os.path.exists(exe):
        raise RuntimeError("oracle/_ref/ref-mulmat-bench missing: run __graft_entry__.build() where /root/reference exists")
    cmd = [exe, "--threads", str(threads), "--steps", str(steps), "--warmup", str(warmup), "--n", "1", "--copies",
           str(copies)] + ref_specs(model)
    out = subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    line = [ln for ln in out.splitlines() if ln.startswith("{")][-1]
    return json.loads(line)


def metric_name(model):
    return "tg128 tok/s (%s-shaped, every quantized MUL_MAT of one decode token, batch 1)" % model


def reference_arm(args, model):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    try:
        r = run_reference(model, args.steps, args.warmup, threads)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"impl": "reference", "unavailable": str(e)[:200]}))
        return
    toks = 1e3 / r["ms_per_step"]
    sample = "%d warm-up + %d timed tokens, each all %d MUL_MATs of %s, 3 weight copies per shape cycled" % (
        args.warmup, args.steps, r["nodes"], model)
    print(json.dumps({
        "impl": "reference", "metric": metric_name(model), "value": toks, "unit": "tok/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
        "config": {"workload": model + " decode (ggml-cpu, %d threads)" % threads},
        "cpu_baseline": {"value": toks, "unit": "tok/s", "cores": threads, "kind": "reference", "sample": sample,
                         "weight_gbs": r["weight_gbs"]},
        "e2e": {"value": toks, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------------------------------ clocks sampler
class ClockSampler:
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, index):
        self.samples, self.reasons, self.stop = [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop and self.nv is not None:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                bits = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.02)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=1)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------ device weights
def gen_weights(torch, t, rows, k, gen, device):
    """random VALID blocks written directly in the quantized format on the device (same rule as
    tests/qref.py:random_blocks): random payload bytes, sane finite fp16 scales"""
    nb = rows * (k // BLOCK[t])
    raw = torch.randint(0, 256, (nb, TSIZE[t]), dtype=torch.uint8, device=device, generator=gen)

    def put(col, scale, signed):
        v = (torch.rand(nb, device=device, generator=gen) * 0.9 + 0.1) * scale
        if signed:
            v = v * (torch.randint(0, 2, (nb,), device=device, generator=gen) * 2 - 1)
        raw[:, col:col + 2] = v.to(torch.float16).view(torch.uint8).view(nb, 2)

    if t == Q4_0:
        put(0, 0.02 / 4, True)
    elif t == Q8_0:
        put(0, 0.02 / 64, True)
    elif t == Q4_K:
        put(0, 0.02 / 200, False)
        put(2, 0.02 / 60, False)
    elif t == Q6_K:
        put(208, 0.02 / 2000, True)
    elif t == IQ4_XS:
        put(0, 0.02 / 2000, True)
    return raw.view(rows, -1)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def config3_sweep(args):
    """BASELINE config 3: MUL_MAT m=4096 k=14336, every weight type, n in {1,2,4,8,512} -- B200 (b200_mul_mat through the
    C ABI, weights cycled over copies that exceed L2) beside the unmodified ggml-cpu (oracle/_ref/ref-mulmat-bench, all host
    threads).  Prints one JSON line with the table; tools read it into BASELINE.md.  Not the bench contract line."""
    import torch
    import b200pkg
    qmm = b200pkg.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    gen = torch.Generator(device=dev)
    gen.manual_seed(3)
    m, k, copies = 4096, 14336, 8
    threads = host_threads()
    rows = []
    for t in (Q4_0, Q8_0, Q4_K, Q6_K, IQ4_XS):
        ws = [gen_weights(torch, t, m, k, gen, dev) for _ in range(copies)]       # 8 x 24..62 MB > 126 MB of L2
        wbytes = m * (k // BLOCK[t]) * TSIZE[t]
        planes = None
        for n in (1, 2, 4, 8, 512):
            x = torch.rand((n, k), device=dev, generator=gen) * 2 - 1
            out = torch.empty((n, m), device=dev)
            if n >= 512 and planes is None:
                planes = [qmm.build_plane(t, w, m, k) for w in ws]      # what the plug-in caches beside the weights

            def calls():
                for i in range(copies):
                    if n >= 512:
                        qmm.mul_mat_mmq(t, planes[i], x, m, k, out=out)
                    else:
                        qmm.mul_mat(t, ws[i], x, m, k, out=out)
            st = torch.cuda.Stream(device=dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(st):
                calls()
                st.synchronize()
                with torch.cuda.graph(g, stream=st):
                    calls()
                for _ in range(2):
                    g.replay()
                st.synchronize()
                reps = 10 if n < 512 else 5
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(reps):
                    g.replay()
                e1.record(st)
                st.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
            del g
            cpu_us = None
            try:
                exe = os.path.join(ROOT, "oracle", "_ref", "ref-mulmat-bench")
                cmd = [exe, "--threads", str(threads), "--steps", "4" if n < 512 else "2", "--warmup", "1", "--n", str(n),
                       "--copies", "4", "%s:%d:%d:4" % (TNAME[t], m, k)]
                r = json.loads([ln for ln in subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                                            text=True, timeout=300).stdout.splitlines() if ln.startswith("{")][-1])
                cpu_us = r["ms_per_step"] * 1e3 / 4
            except Exception:  # noqa: BLE001
                pass
            rows.append({"type": TNAME[t], "n": n, "b200_us": us, "b200_gbs": (wbytes + n * k * 4 + n * m * 4) / us / 1e3,
                         "b200_tops": 2.0 * m * k * n / us / 1e6, "ggml_cpu_us": cpu_us, "cpu_threads": threads,
                         "path": "tensor cores (cached int8 plane)" if n >= 512 else "gemv"})
        del ws, planes
    print(json.dumps({"config3_sweep": rows, "m": m, "k": k, "note": "CUDA-graph replay of b200_mul_mat (n <= 8) / b200_mul_mat_mmq "
                      "with cached planes (n = 512) over %d weight copies (> L2); activation quantizer included" % copies}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=list(MODELS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-pp", action="store_true", help="skip the pp512 leg")
    ap.add_argument("--sweep", action="store_true", help="BASELINE config 3 table (MUL_MAT 4096x14336, all types, n=1..512) instead of the bench line")
    ap.add_argument("--collective", default="fused", choices=["fused", "nccl"],
                    help="N>1: how the row slices come together (fused = peer stores from the GEMV epilogue)")
    ap.add_argument("--per-node", action="store_true",
                    help="N=1: one b200_mul_mat call per MUL_MAT node instead of b200_mul_mat_multi per src1 group")
    ap.add_argument("--no-single", action="store_true", help="N>1: skip the same-workload single-GPU leg on rank 0")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA-graph replay")
    ap.add_argument("--no-sub", action="store_true", help="N=1: skip the llama3-8b-q4_k_m sub-run (configs[1])")
    ap.add_argument("--sub", action="store_true", help="this IS the sub-run: full single-GPU legs, no sub-run of its own")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 1)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    model = args.workload or "llama3-70b-q4_k_m"     # the same workload at every N (see the module docstring)
    # the headline run at N = 1 leaves pp512 / MUL_MAT_ID (quoted on the 8B config) to the sub-run
    with_sub = world == 1 and args.gpus == 1 and not args.sub and not args.no_sub and model != "llama3-8b-q4_k_m" \
        and args.impl == "b200"
    if with_sub:
        args.no_pp = True

    if args.sweep:
        config3_sweep(args)
        return
    if args.impl == "reference":
        return reference_arm(args, model)

    import torch
    import b200pkg
    qmm = b200pkg.load()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    assert qmm.lib().b200_device_ok(local_rank) == 1, "not an sm_100 device"
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    mms = matmul_list(model)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)          # same weights on every rank before slicing
    n = 1
    # ---- weights: this rank's row slice of every matrix, resident in HBM
    weights, slices = [], []
    for name, t, m, k in mms:
        lo, hi = qmm.row_split(m, world, rank)
        full_rows = m
        # generate only this rank's slice (seeded per matrix so that slices are reproducible)
        g = torch.Generator(device=dev)
        g.manual_seed(zlib.crc32(("%s/%d" % (name, rank)).encode()))
        weights.append(gen_weights(torch, t, hi - lo, k, g, dev) if hi > lo else None)
        slices.append((lo, hi, full_rows))
    ks = sorted({k for _, _, _, k in mms})
    xs = {k: torch.rand((n, k), device=dev, generator=gen) * 2 - 1 for k in ks}
    # results: one buffer per distinct m; in row-split mode each rank writes its slice and the slices are gathered
    maxrows = {m: max(qmm.row_split(m, world, r)[1] - qmm.row_split(m, world, r)[0] for r in range(world))
               for _, _, m, _ in mms}
    out_local = {m: torch.zeros((n, maxrows[m]), device=dev) for m in maxrows}
    out_full = {m: torch.zeros((world, n, maxrows[m]), device=dev) for m in maxrows} if world > 1 else None

    def token_step_per_node():
        for (name, t, m, k), w, (lo, hi, _) in zip(mms, weights, slices):
            if w is not None:
                qmm.mul_mat(t, w, xs[k], hi - lo, k, out=out_local[m][:, :hi - lo])
            if world > 1:
                dist.all_gather_into_tensor(out_full[m], out_local[m])

    # N = 1: MUL_MATs that read the same activations (q/k/v of a layer; gate/up) go through ONE b200_mul_mat_multi call:
    # x quantized once, same-type matrices in one launch.  Same nodes, same results; --per-node keeps one call per node.
    token_step = token_step_per_node
    grouping = "one b200_mul_mat per MUL_MAT node"
    if world == 1 and not args.per_node:
        def group_of(name):
            base = name.split(".")[0]
            return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
        node_groups = []
        for (name, t, m, k), w in zip(mms, weights):
            key = (group_of(name), name.split(".")[-1])
            if not node_groups or node_groups[-1][0] != key:
                node_groups.append((key, k, []))
            node_groups[-1][2].append((name.split(".")[0], t, m, w))
        out_node = {}
        calls = []
        for _, k, members in node_groups:
            outs = []
            for base, t, m, w in members:
                if base not in out_node:
                    out_node[base] = torch.zeros((n, m), device=dev)
                outs.append(out_node[base])
            calls.append((qmm.MultiMatMul([t for _, t, _, _ in members], [w for _, _, _, w in members],
                                          [m for _, _, m, _ in members], k, outs, n), xs[k]))

        def token_step_grouped():
            for call, x in calls:
                call(x)
        token_step = token_step_grouped
        grouping = "b200_mul_mat_multi per group of nodes sharing src1 (%d calls for %d nodes)" % (len(calls), len(mms))

    weight_bytes_local = sum(w.numel() for w in weights if w is not None)
    step_bytes = sum(gemv_bytes(t, hi - lo, k, n) for (_, t, _, k), (lo, hi, _) in zip(mms, slices) if hi > lo)
    total_bytes_all_ranks = sum(gemv_bytes(t, m, k, n) for _, t, m, k in mms)

    # ---- warm-up (eager), count launches per step
    before = qmm.launch_count()
    token_step()
    torch.cuda.synchronize()
    launches_per_step = qmm.launch_count() - before
    for _ in range(2):
        token_step()
    torch.cuda.synchronize()

    if os.environ.get("B200_BENCH_NVTX_STEP"):
        # profiling aid (tools/gpu_scripts/final_1gpu.sh): ONE eager token inside an NVTX range, so that `ncu --nvtx --nvtx-include
        # b200_step/` sees exactly the kernels of a step (DRAM bytes for roofline.traffic, the launch list); then stop
        torch.cuda.nvtx.range_push("b200_step")
        token_step()
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        print(json.dumps({"nvtx_step": True, "launches_per_step": launches_per_step, "algorithmic_bytes_per_step": step_bytes,
                          "workload": model}))
        return

    stream = torch.cuda.Stream(device=dev)
    # one CUDA graph per token (NCCL collectives are capturable too); B200_BENCH_GRAPH_MULTI=0 keeps N>1 eager
    use_graph = not args.no_graph and (world == 1 or os.environ.get("B200_BENCH_GRAPH_MULTI", "1") == "1")
    if use_graph:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(stream):
            token_step()
            stream.synchronize()
            with torch.cuda.graph(graph, stream=stream):
                token_step()
        run = graph.replay
    else:
        run = token_step

    def timed(fn, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                fn()
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
                torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(steps):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
                torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            tms = torch.tensor([ms], device=dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms.item())
        return ms / steps

    with ClockSampler(local_rank) as clk:
        ms_step = timed(run, args.steps, args.warmup)
        # make sure the sampler saw the loaded state even for very short runs
        t_end = time.time() + 0.3
        while len(clk.samples) < 8 and time.time() < t_end:
            run()
        torch.cuda.synchronize()
    clocks = clk.summary()
    ms_eager = timed(token_step, max(4, args.steps // 8), 3) if use_graph else ms_step
    per_node = None
    if world == 1 and token_step is not token_step_per_node:
        # the same token with one call (one quantizer + one GEMV launch) per MUL_MAT node, for comparison
        b0 = qmm.launch_count()
        token_step_per_node()
        torch.cuda.synchronize()
        l_node = qmm.launch_count() - b0
        run_n = token_step_per_node
        if use_graph:
            gn = torch.cuda.CUDAGraph()
            with torch.cuda.stream(stream):
                token_step_per_node()
                stream.synchronize()
                with torch.cuda.graph(gn, stream=stream):
                    token_step_per_node()
            run_n = gn.replay
        ms_node = timed(run_n, max(8, args.steps // 4), 3)
        per_node = {"value": 1e3 / ms_node, "unit": "tok/s", "ms_per_step": ms_node, "launches_per_step": l_node,
                    "hbm_gbs": step_bytes / ms_node / 1e6}
        del run_n

    # ---- N = 1: the same token as ONE persistent launch (b200_sequence_run): every group marked dependent on the groups
    #      before it, i.e. the ordering consecutive calls on a stream have.  Checked bit for bit against the grouped calls.
    sequence = None
    if world == 1 and token_step is not token_step_per_node:
        seq_groups = []
        for (_, k, members), (call, x) in zip(node_groups, calls):
            seq_groups.append((x, k, [(t, w, m, o) for (_, t, m, w), o in zip(members, call.keep[1])], True))
        token_step()
        torch.cuda.synchronize()
        want = {b: o.clone() for b, o in out_node.items()}
        for o in out_node.values():
            o.zero_()
        seq = qmm.Sequence(seq_groups)
        seq()
        torch.cuda.synchronize()
        same = all(torch.equal(out_node[b], want[b]) for b in out_node) and seq.status() == 0
        gs = torch.cuda.CUDAGraph()
        with torch.cuda.stream(stream):
            seq()
            stream.synchronize()
            with torch.cuda.graph(gs, stream=stream):
                seq()
        ms_seq = timed(gs.replay, max(8, args.steps // 2), 3)
        ms_seq_eager = timed(seq, max(8, args.steps // 4), 3)
        sequence = {"value": 1e3 / ms_seq, "unit": "tok/s", "ms_per_step": ms_seq, "ms_per_step_eager": ms_seq_eager,
                    "launches_per_step": 1, "groups": len(seq_groups), "bit_identical_to_grouped_calls": bool(same),
                    "hbm_gbs": step_bytes / ms_seq / 1e6,
                    "api": "b200_sequence_plan_build once + b200_sequence_run per token (one persistent kernel, in-CTA "
                           "activation quantizer, release/acquire counters between dependent groups)"}
        del gs

    # ---- N = 1: e2e THROUGH THE REFERENCE'S ggml API and the plug-in: host tensors -> ggml_backend_tensor_set ->
    #      ggml_backend_graph_compute(libggml-b200.so) -> ggml_backend_tensor_get, per group of nodes sharing src1
    #      (oracle/ggml_graph_bench.cpp; its own weights, same shapes).  What a ggml caller gets.
    e2e_ggml = None
    if world == 1:
        exe = os.path.join(ROOT, "oracle", "_ref", "ggml-graph-bench")
        plug = os.path.join(ROOT, "llama.cpp.dsp_b200", "lib", "libggml-b200.so")
        if os.path.exists(exe) and os.path.exists(plug):
            c = MODELS[model]
            cmd = [exe, "--backend", "B200_%d" % local_rank, "--steps", str(max(8, args.steps // 4)), "--warmup", "3",
                   "--layers", str(c["n_layer"]), "--embd", str(c["n_embd"]), "--ff", str(c["n_ff"]), "--kv", str(c["n_kv"]),
                   "--vocab", str(c["vocab"])] + (["--q4_0"] if c["recipe"] == "q4_0" else [])
            try:
                out = subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=600,
                                     env=dict(os.environ, GGML_BACKEND_PATH=plug)).stdout
                r = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
                e2e_ggml = {"value": r["tok_s"], "unit": "tok/s", "ms_per_step": r["ms_per_step"],
                            "h2d_bytes_per_step": r["h2d_bytes_per_step"], "d2h_bytes_per_step": r["d2h_bytes_per_step"],
                            "graph_compute_calls_per_step": r["graph_compute_calls"], "host_buffer": r["host_buffer"],
                            "api": "ggml_backend_tensor_set + ggml_backend_graph_compute + ggml_backend_tensor_get through "
                                   "libggml-b200.so (GGML_BACKEND_PATH), one graph per group of MUL_MAT nodes sharing src1"}
            except Exception as e:  # noqa: BLE001
                e2e_ggml = {"value": None, "error": str(e)[:200]}

    # ---- N > 1: the fused path -- each GEMV stores its rows into every rank's dst over NVLink, a stream wait-value
    #      replaces the collective (b200_mul_mat_rowsplit_fused + b200_rowsplit_wait).  When it runs, it is the headline
    #      and the NCCL all-gather figure above is kept beside it as the baseline.
    collective, nccl_leg = "nccl_all_gather", None
    if world > 1 and args.collective == "fused":
        from llama_cpp_dsp_b200.rowsplit import PeerBuffers, RowSplit, mul_mat_row_split_fused, row_split_wait
        per_m = sorted({m for _, _, m, _ in mms})
        peers = {m: PeerBuffers(n, m, 1, world, rank, dev) for m in per_m}
        splits = {m: RowSplit(m, world, rank) for m in per_m}
        fws = torch.empty(max(qmm.lib().b200_mul_mat_rowsplit_workspace_size(t, n, k) for _, t, _, k in mms) + 256,
                          dtype=torch.uint8, device=dev)
        # dependency groups of the decode graph: {q,k,v} -> {attn_output} -> {gate,up} -> {down} -> next layer ... ->
        # output.  Matmuls of one group read the same x and are independent; the first matmul of a group needs the
        # previous group's results whole on this rank.  One sync block per group.
        def group_of(name):
            base = name.split(".")[0]
            return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
        groups = []
        for (name, t, m, k), w in zip(mms, weights):
            key = (group_of(name), name.split(".")[-1])
            if not groups or groups[-1][0] != key:
                groups.append((key, []))
            groups[-1][1].append((t, m, k, w))
        gsync = PeerBuffers(1, 1, len(groups), world, rank, dev)
        dist.barrier()

        def token_step_fused():
            prev, prev_calls = None, 0
            for gi, (_, members) in enumerate(groups):
                ps = gsync.ptr_sync[gi]
                for j, (t, m, k, w) in enumerate(members):
                    mul_mat_row_split_fused(t, w, xs[k], splits[m], k, peers[m].ptr_dst, ps, m, fws,
                                            wait_sync=prev if j == 0 else None, wait_calls=prev_calls)
                prev, prev_calls = ps, len(members)
            row_split_wait(prev, rank, world, calls=prev_calls)

        token_step_fused()
        torch.cuda.synchronize()
        dist.barrier()
        # cross-check against the all-gather path before timing anything
        token_step()
        torch.cuda.synchronize()
        problems = []
        for m in per_m:
            ref_full = torch.cat([out_full[m][r, :, :hi - lo] for r, (lo, hi) in enumerate(splits[m].ranges)], dim=1)
            if not torch.equal(ref_full, peers[m].dst):
                bad = (ref_full != peers[m].dst).nonzero()[:, 1]
                i0 = int(bad.min())
                problems.append("m=%d: %d elements differ, rows %d..%d (own rows %d..%d); got %s want %s" % (
                    m, bad.numel(), i0, int(bad.max()), splits[m].lo, splits[m].hi,
                    peers[m].dst[0, i0:i0 + 3].tolist(), ref_full[0, i0:i0 + 3].tolist()))
        everyone = [None] * world
        dist.all_gather_object(everyone, problems)
        anybad = any(everyone)
        if anybad and os.environ.get("B200_BENCH_DIAG"):
            for m in per_m:
                sp = splits[m]
                plo, phi = sp.ranges[1 - rank] if world == 2 else sp.ranges[0]
                hits, own = [], []
                for (name, t, mm_, k), w in zip(mms, weights):
                    if mm_ != m:
                        continue
                    qmm.mul_mat(t, w, xs[k], sp.rows, k, out=out_local[m][:, :sp.rows])
                    dist.all_gather_into_tensor(out_full[m], out_local[m])
                    full = torch.cat([out_full[m][r, :, :hi - lo] for r, (lo, hi) in enumerate(sp.ranges)], dim=1)
                    if torch.equal(full[:, plo:phi], peers[m].dst[:, plo:phi]):
                        hits.append(name)
                    if torch.equal(full[:, sp.lo:sp.hi], peers[m].dst[:, sp.lo:sp.hi]):
                        own.append(name)
                print("DIAG rank %d m=%d peer rows match: %s ; own rows match: %s" % (rank, m, hits[-3:], own[-3:]),
                      file=sys.stderr, flush=True)
        if anybad:
            print("bench.py: rank %d: fused row-split result differs from the all-gather result: %s"
                  % (rank, "; ".join(problems) or "(this rank is fine)"), file=sys.stderr, flush=True)
            dist.barrier()
            os._exit(3)
        run_f = token_step_fused
        fused_graph = False
        if use_graph:
            gf = torch.cuda.CUDAGraph()
            with torch.cuda.stream(stream):
                token_step_fused()
                stream.synchronize()
                with torch.cuda.graph(gf, stream=stream):
                    token_step_fused()
            run_f, fused_graph = gf.replay, True
        dist.barrier()
        with ClockSampler(local_rank) as clk2:
            ms_fused = timed(run_f, args.steps, args.warmup)
        fused_info = {"sync_groups_per_step": len(groups), "stream_waits_per_step": 1,
                      "in_kernel_waits_per_step": len(groups) - 1}
        nccl_leg = {"value": 1e3 / ms_step, "unit": "tok/s", "ms_per_step": ms_step,
                    "launch": "cuda_graph_replay" if use_graph else "eager",
                    "note": "same slices, dist.all_gather_into_tensor per matmul (the non-fused baseline)"}
        ms_step, clocks, collective = ms_fused, clk2.summary(), "fused_peer_store"
        use_graph = fused_graph

    # ---- N > 1: the whole token as ONE persistent launch per rank (b200_sequence_plan_build_rowsplit): the epilogue mirrors
    #      every row into all ranks' symmetric result buffers over NVLink, groups are published with system-scope counters --
    #      no per-matmul launches, waits or collectives at all.  Verified against an all-gather of per-rank results first.
    seq_split = None
    if world > 1 and args.collective == "fused":
        bases, off = {}, 0
        for name, t, m, k in mms:
            b_ = name.split(".")[0]
            if b_ not in bases:
                bases[b_] = (off, m)
                off += (m + 63) // 64 * 64
        sgroups, sgrp_src = [], []
        for (name, t, m, k), w, (lo, hi, _) in zip(mms, weights, slices):
            key = (group_of(name), name.split(".")[-1])
            if not sgrp_src or sgrp_src[-1][0] != key:
                sgrp_src.append((key, k, []))
            sgrp_src[-1][2].append((name.split(".")[0], t, w, lo, hi))
        sym = PeerBuffers(1, off, 2 * len(sgrp_src) + 8, world, rank, dev)
        for _, k, members in sgrp_src:
            mats = []
            for b_, t, w, lo, hi in members:
                o0 = bases[b_][0]
                mats.append((t, w if w is not None else fws, bases[b_][1], sym.dst[0, o0:o0 + bases[b_][1]]))   # whole matrix, w = own rows
            sgroups.append((xs[k], k, mats, True))
        use_ll = os.environ.get("B200_SEQ_SPLIT_LL", "1") != "0"
        sym_ll = PeerBuffers(1, 2 * off, 1, world, rank, dev) if use_ll else None
        seqs = qmm.Sequence(sgroups, split=(world, rank, sym.ptr_dst, off * 4, sym.ptr_sync[0], sym_ll.ptr_dst if use_ll else None))
        dist.barrier()
        seqs()
        torch.cuda.synchronize()
        dist.barrier()
        # check: the last layer's matrices of every base (earlier layers were overwritten) against per-rank results gathered by NCCL
        bad = []
        last = {}
        for (name, t, m, k), w, (lo, hi, _) in zip(mms, weights, slices):
            last[name.split(".")[0]] = (t, m, k, w, lo, hi)
        for b_, (t, m, k, w, lo, hi) in last.items():
            loc = torch.zeros((1, maxrows[m]), device=dev)
            if w is not None:
                qmm.mul_mat(t, w, xs[k], hi - lo, k, out=loc[:, :hi - lo])
            allv = torch.empty((world, 1, maxrows[m]), device=dev)
            dist.all_gather_into_tensor(allv, loc)
            full = torch.cat([allv[r, :, :h - l] for r, (l, h) in enumerate(splits[m].ranges)], dim=1)
            o0 = bases[b_][0]
            if not torch.equal(full[0], sym.dst[0, o0:o0 + m]):
                bad.append(b_)
        everyone = [None] * world
        dist.all_gather_object(everyone, (bad, seqs.status()))
        ok = not any(b for b, _ in everyone) and not any(e for _, e in everyone)
        if ok:
            gq = torch.cuda.CUDAGraph()
            with torch.cuda.stream(stream):
                seqs()
                stream.synchronize()
                dist.barrier()
                with torch.cuda.graph(gq, stream=stream):
                    seqs()
            dist.barrier()
            ms_sq = timed(gq.replay, args.steps, args.warmup)
            seq_split = {"value": 1e3 / ms_sq, "unit": "tok/s", "ms_per_step": ms_sq, "launches_per_step": 1,
                         "bit_identical_to_per_rank_results": True, "groups": len(sgroups),
                         "exchange": "flag in the data: 8-byte {value, tag} peer stores, local collection + local counters" if use_ll
                                     else "mirrored stores + system-scope fence + remote counters",
                         "hbm_gbs_per_gpu": step_bytes / ms_sq / 1e6,
                         "api": "b200_sequence_plan_build_rowsplit + b200_sequence_run per rank: one persistent kernel per GPU, rows "
                                "stored into every rank over NVLink from the epilogue"}
            if os.environ.get("B200_BENCH_SEQ_TRACE"):
                # in-kernel globaltimer stamps of one token (b200_debug_sequence_trace): where a group's time goes on this rank
                import numpy as np
                ngp, ncta = len(sgroups), qmm.lib().b200_sm_count(local_rank)
                trace = torch.zeros((2 * ngp, ncta, 8), dtype=torch.int64, device=dev)
                qmm.check(qmm.lib().b200_debug_sequence_trace(seqs.plan.data_ptr(), trace.data_ptr(), None))
                dist.barrier()
                seqs()
                seqs()
                torch.cuda.synchronize()
                qmm.check(qmm.lib().b200_debug_sequence_trace(seqs.plan.data_ptr(), None, None))
                tr = trace.cpu().numpy()[:ngp]
                ok = (tr[:-1, :, 0] > 0) & (tr[1:, :, 0] > 0) & (tr[:-1, :, 6] > 0)
                a, b = tr[:-1], tr[1:]
                med = lambda v: float(np.median(v[ok])) / 1e3
                seq_split["trace_us_median"] = {
                    "wait_for_all_ranks": med(a[:, :, 1] - a[:, :, 0]), "quantize": med(a[:, :, 2] - a[:, :, 1]),
                    "first_weights": med(a[:, :, 3] - a[:, :, 2]), "dot_products": med(a[:, :, 4] - a[:, :, 3]),
                    "team0_done_to_cta_done": med(a[:, :, 6] - a[:, :, 4]), "publish_fence_and_adds": med(a[:, :, 7] - a[:, :, 6]),
                    "to_next_boundary": med(b[:, :, 0] - a[:, :, 7]),
                    "group_period": float(np.median((b[:, :, 0] - a[:, :, 0])[ok])) / 1e3}
            if ms_sq < ms_step:
                fused_leg = {"value": 1e3 / ms_step, "unit": "tok/s", "ms_per_step": ms_step}
                ms_step, collective = ms_sq, "sequence_peer_store"
                seq_split["fused_per_matmul_launches"] = fused_leg
        else:
            seq_split = {"value": None, "error": "mismatch or timeout: %s" % (everyone,)}

    # ---- e2e: host buffers, per-matmul H2D -> kernel -> D2H through b200_mul_mat_host (rank-local slice)
    hx = {k: xs[k].cpu().pin_memory() for k in ks}
    hout = {m: torch.zeros((n, maxrows[m])).pin_memory() for m in maxrows}
    h2d = sum(n * k * 4 for (_, _, _, k), w in zip(mms, weights) if w is not None)
    d2h = sum(n * (hi - lo) * 4 for (lo, hi, _), w in zip(slices, weights) if w is not None)

    def token_step_host():
        for (name, t, m, k), w, (lo, hi, _) in zip(mms, weights, slices):
            if w is not None:
                dst = hout[m] if hi - lo == maxrows[m] else hout[m][:, :hi - lo].contiguous()
                qmm.mul_mat_host(t, w, hx[k], dst, hi - lo, k)
        torch.cuda.current_stream().synchronize()   # the caller reads the results of this token on the host

    e2e = None
    if world == 1:
        ms_e2e_eager = timed(token_step_host, max(8, args.steps // 4), 3)
        ms_e2e, e2e_launch = ms_e2e_eager, "eager"
        if use_graph:
            # the same calls captured once and replayed: every replay still does every H2D copy, kernel and D2H copy
            # (b200_mul_mat_host is stream-ordered, so a caller can capture it like any other stream work)
            if token_step is token_step_per_node:
                def host_calls_only():
                    for (name, t, m, k), w, (lo, hi, _) in zip(mms, weights, slices):
                        qmm.mul_mat_host(t, w, hx[k], hout[m], hi - lo, k)
            else:
                # grouped: one H2D copy of x per group, results of a group adjacent on the host -> one D2H copy per group
                hcalls, hgroup_out = [], {}
                for gkey, k, members in node_groups:
                    total = sum(m for _, _, m, _ in members)
                    sig = (gkey[0], total)
                    if sig not in hgroup_out:
                        hgroup_out[sig] = torch.zeros(n * total).pin_memory()
                    outs, off = [], 0
                    for _, _, m, _ in members:
                        outs.append(hgroup_out[sig][off:off + n * m].view(n, m))
                        off += n * m
                    hcalls.append((qmm.MultiMatMulHost([t for _, t, _, _ in members], [w for _, _, _, w in members],
                                                       [m for _, _, m, _ in members], k, outs, n), hx[k]))

                def host_calls_only():
                    for call, xh in hcalls:
                        call(xh)
                h2d = sum(n * k * 4 for _, k, _ in node_groups)
            gh = torch.cuda.CUDAGraph()
            with torch.cuda.stream(stream):
                host_calls_only()
                stream.synchronize()
                with torch.cuda.graph(gh, stream=stream):
                    host_calls_only()

            def token_step_host_graph():
                gh.replay()
                torch.cuda.current_stream().synchronize()   # the caller reads this token's results on the host
            ms_e2e, e2e_launch = timed(token_step_host_graph, max(8, args.steps // 4), 3), "cuda_graph_replay"
        e2e = {"value": 1e3 / ms_e2e, "unit": "tok/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": ms_e2e, "launch": e2e_launch, "ms_per_step_eager": ms_e2e_eager,
               "api": ("b200_mul_mat_host per MUL_MAT" if token_step is token_step_per_node or not use_graph else
                       "b200_mul_mat_multi_host per group of MUL_MATs sharing src1") +
                      " (pinned host x -> read over PCIe by the quantizer kernel, no staging copy -> GEMV -> results"
                      " stored into pinned host memory by a copy kernel), stream sync once per token",
               "h2d": "zero-copy: the activation / result bytes counted here cross the bus as kernel loads from and "
                      "kernel stores to pinned host memory (B200_HOST_ZEROCOPY=0 restores cudaMemcpyAsync both ways)"
               if os.environ.get("B200_HOST_ZEROCOPY", "1") != "0" else "cudaMemcpyAsync per call"}
    else:
        # row-split: host activations are broadcast to every rank's device, the gathered result is read back on rank 0
        def token_step_host_split():
            for k in ks:
                xs[k].copy_(hx[k], non_blocking=True)
            if collective == "sequence_peer_store":
                seqs()
                for m, (o0, _) in seq_last_of_m.items():
                    hfull[m].copy_(sym.dst[:, o0:o0 + m], non_blocking=True)
            elif collective == "fused_peer_store":
                token_step_fused()
                for m in maxrows:
                    hfull[m].copy_(peers[m].dst, non_blocking=True)
            else:
                token_step()
                for m in maxrows:
                    hout[m].copy_(out_full[m][0], non_blocking=True)
            torch.cuda.current_stream().synchronize()
        hfull = {m: torch.zeros((n, m)).pin_memory() for m in maxrows}
        if collective == "sequence_peer_store":
            seq_last_of_m = {m: (bases[b_][0], b_) for b_, (_, m, _, _, _, _) in last.items()}
        ms_e2e = timed(token_step_host_split, max(8, args.steps // 4), 3)
        fusedp = collective in ("fused_peer_store", "sequence_peer_store")
        e2e = {"value": 1e3 / ms_e2e, "unit": "tok/s", "h2d_bytes_per_step": sum(n * k * 4 for k in ks),
               "d2h_bytes_per_step": sum(n * (m if fusedp else maxrows[m]) * 4 for m in maxrows), "ms_per_step": ms_e2e,
               "api": "pinned host activations -> H2D once per token -> row-split matmuls (%s) -> D2H of the last "
                      "result per distinct m" % collective}

    # ---- pp512 leg (prefill): the same matmuls at n = 512 (output head at n = 1, as llama does for the last token)
    pp = None
    if not args.no_pp and world == 1:
        npp = 512
        xpp = {k: torch.rand((npp, k), device=dev, generator=gen) * 2 - 1 for k in ks}
        opp = {m: torch.empty((npp, m), device=dev) for _, _, m, _ in mms[:-1]}

        # int8 "prefill planes" of the weights: built once per tensor (the ggml plug-in builds them lazily at first
        # prefill and keeps them with the weight buffer); +1 byte of HBM per weight
        t_pl0 = time.time()
        planes = [qmm.build_plane(t, w, m, k) for (name, t, m, k), w in zip(mms[:-1], weights[:-1])]
        torch.cuda.synchronize()
        plane_build_s = time.time() - t_pl0
        plane_bytes = sum(p_.numel() for p_ in planes)

        def prefill_step_per_node():
            for (name, t, m, k), pl in zip(mms[:-1], planes):
                qmm.mul_mat_mmq(t, pl, xpp[k], m, k, out=opp[m])
            name, t, m, k = mms[-1]
            qmm.mul_mat(t, weights[-1], xs[k], m, k, out=out_local[m])

        # nodes that read the same activations (q/k/v, gate/up) share one b200_mul_mat_mmq_multi call
        prefill_step = prefill_step_per_node
        if not args.per_node:
            def pgroup(name):
                base = name.split(".")[0]
                return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
            pgroups = []
            for (name, t, m, k), pl in zip(mms[:-1], planes):
                key = (pgroup(name), name.split(".")[-1])
                if not pgroups or pgroups[-1][0] != key:
                    pgroups.append((key, k, []))
                pgroups[-1][2].append((name.split(".")[0], t, m, pl))
            opp_node = {}
            pcalls = []
            for _, k, members in pgroups:
                outs = []
                for base, t, m, pl in members:
                    if base not in opp_node:
                        opp_node[base] = torch.empty((npp, m), device=dev)
                    outs.append(opp_node[base])
                pcalls.append((qmm.MultiMatMulMMQ([t for _, t, _, _ in members], [pl for _, _, _, pl in members],
                                                  [m for _, _, m, _ in members], k, outs, npp), xpp[k]))

            def prefill_step_grouped():
                for call, x in pcalls:
                    call(x)
                name, t, m, k = mms[-1]
                qmm.mul_mat(t, weights[-1], xs[k], m, k, out=out_local[m])
            prefill_step = prefill_step_grouped

        def prefill_step_transient():   # no cached planes: b200_mul_mat rebuilds each plane in its workspace
            for (name, t, m, k), w in zip(mms[:-1], weights[:-1]):
                qmm.mul_mat(t, w, xpp[k], m, k, out=opp[m])

        b0 = qmm.launch_count()
        prefill_step()
        torch.cuda.synchronize()
        pp_launches = qmm.launch_count() - b0
        # like the decode step, the prefill step is captured once and replayed (tensor maps are kernel parameters, so
        # capture is legal); the eager figure stays beside it
        ms_pp_eager = timed(prefill_step, 8, 3)
        ms_pp = ms_pp_graph = ms_pp_eager
        if use_graph:
            gpp = torch.cuda.CUDAGraph()
            with torch.cuda.stream(stream):
                prefill_step()
                stream.synchronize()
                with torch.cuda.graph(gpp, stream=stream):
                    prefill_step()
            ms_pp_graph = timed(gpp.replay, 8, 3)
            del gpp
            # programmatic dependent launch (quantizer -> GEMM prologue overlap) is what eager stream launches give;
            # a captured graph may serialise those edges.  The step is long enough (> 5 ms, ~260 launches) not to be
            # launch-bound, so the headline is the faster of the two and both are reported.
            ms_pp = min(ms_pp_eager, ms_pp_graph)
        ms_pp_transient = timed(prefill_step_transient, 2, 1)
        ms_pp_node = timed(prefill_step_per_node, 3, 1) if prefill_step is not prefill_step_per_node else ms_pp
        flops = sum(2.0 * m * k * npp for _, _, m, k in mms[:-1])
        # int8 tensor-core denominator: measured live with a cuBLASLt int8 GEMM (torch._int_mm), else 2 x measured bf16
        int8_peak, int8_how = None, None
        try:
            A8 = torch.randint(-127, 127, (8192, 8192), dtype=torch.int8, device=dev)
            B8 = torch.randint(-127, 127, (8192, 8192), dtype=torch.int8, device=dev)
            for _ in range(3):
                torch._int_mm(A8, B8)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                torch._int_mm(A8, B8)
            e1.record()
            torch.cuda.synchronize()
            int8_peak = 2.0 * 8192 ** 3 * 10 / (e0.elapsed_time(e1) * 1e-3) / 1e12
            int8_how = "measured live: torch._int_mm 8192^3 (cuBLASLt int8), 10 back-to-back"
            del A8, B8
        except Exception as e:  # noqa: BLE001
            try:
                int8_peak = 2.0 * float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
                int8_how = "2 x MEASURED_PEAKS.json bf16_tflops (int8 probe failed: %s)" % str(e)[:60]
            except Exception:  # noqa: BLE001
                int8_peak, int8_how = 3180.0, "fallback 2 x 1.59 PFLOP/s"
        tops = flops / ms_pp / 1e9
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            bf16_burst, bf16_sust = float(mp["bf16_tflops"]), float(mp.get("bf16_tflops_sustained", 0) or 0)
        except Exception:  # noqa: BLE001
            bf16_burst, bf16_sust = 1590.0, 1400.0
        pp = {"metric": "pp512 tok/s", "value": npp * 1e3 / ms_pp, "unit": "tok/s", "ms_per_step": ms_pp,
              "steps": 8, "launch": "cuda_graph_replay" if (use_graph and ms_pp_graph <= ms_pp_eager) else "eager",
              "ms_per_step_eager": ms_pp_eager, "ms_per_step_graph": ms_pp_graph if use_graph else None,
              "launches_per_step": pp_launches,
              "scale_chunk_k": int(qmm.lib().b200_mmq_chunk()),
              "roofline": {"bound": "tensor", "achieved": tops, "peak": int8_peak, "unit": "TOP/s (int8)",
                           "frac": tops / int8_peak, "peak_source": int8_how, "nominal_int8_tops": 4500.0,
                           "frac_of_nominal": tops / 4500.0,
                           # driver-written denominator: int8 dense = 2 x the measured bf16 GEMM throughput
                           "peak_2x_bf16_burst": 2 * bf16_burst, "frac_of_2x_bf16_burst": tops / (2 * bf16_burst),
                           "peak_2x_bf16_sustained": 2 * bf16_sust or None,
                           "frac_of_2x_bf16_sustained": tops / (2 * bf16_sust) if bf16_sust else None,
                           "kernel": "k_mmq_p / k_mmq (tcgen05.mma kind::i8, TMEM accumulators)"},
              "per_node_calls": {"value": npp * 1e3 / ms_pp_node, "unit": "tok/s", "ms_per_step": ms_pp_node},
              "planes": {"cached": True, "bytes": plane_bytes, "build_seconds_once": plane_build_s},
              "transient_planes": {"value": npp * 1e3 / ms_pp_transient, "unit": "tok/s", "ms_per_step": ms_pp_transient}}
        del planes

    # ---- N > 1: pp512 with split-mode row: every rank multiplies its row slice on the tensor cores (cached planes of the
    #      slice) and the [512 x rows] result slices are all-gathered per matmul -- MB-sized messages, bandwidth-bound, so
    #      the NCCL collective is the right tool here (the fused peer-store epilogue exists for the 4-32 KB decode results)
    if world > 1 and not args.no_pp:
        npp = 512
        xpp = {k: torch.rand((npp, k), device=dev, generator=gen) * 2 - 1 for k in ks}
        sel = [(i, mm) for i, mm in enumerate(mms[:-1])]
        planes = [qmm.build_plane(mms[i][1], weights[i], slices[i][1] - slices[i][0], mms[i][3]) if weights[i] is not None else None
                  for i, _ in sel]
        # one local slice buffer and one gathered buffer per kind of matrix (attn_q, ffn_up, ...), as in the decode legs
        opp_l, opp_f = {}, {}
        for _, (name, _, m, _) in sel:
            b_ = name.split(".")[0]
            if b_ not in opp_l:
                opp_l[b_] = torch.empty((npp, maxrows[m]), device=dev)
                opp_f[b_] = torch.empty((world, npp, maxrows[m]), device=dev)

        def pgroup2(name):
            base = name.split(".")[0]
            return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
        # matmuls that read the same activations share one b200_mul_mat_mmq_multi call (activations quantized once); the
        # group's result slices are then all-gathered, and the next group starts after that (it needs the whole result)
        pg = []
        for (i, (name, t, m, k)), pl in zip(sel, planes):
            key = (pgroup2(name), name.split(".")[-1])
            if not pg or pg[-1][0] != key:
                pg.append((key, k, []))
            lo, hi, _ = slices[i]
            pg[-1][2].append((name.split(".")[0], t, hi - lo, pl, lo))
        pcalls2 = []
        for _, k, members in pg:
            own = [(b_, t, rows, pl) for b_, t, rows, pl, _ in members if pl is not None and rows > 0]
            call = None
            if own:
                call = qmm.MultiMatMulMMQ([t for _, t, _, _ in own], [pl for _, _, _, pl in own], [rows for _, _, rows, _ in own], k,
                                          [opp_l[b_][:, :rows] for b_, _, rows, _ in own], npp)
            pcalls2.append((call, xpp[k], [b_ for b_, _, _, _, _ in members]))

        def prefill_step_split():
            for call, x, bs in pcalls2:
                if call is not None:
                    call(x)
                for b_ in bs:
                    dist.all_gather_into_tensor(opp_f[b_], opp_l[b_])
        for _ in range(2):
            prefill_step_split()
        torch.cuda.synchronize()
        dist.barrier()
        ms_pp_eager = timed(prefill_step_split, 6, 2)
        ms_pp, pp_launch = ms_pp_eager, "eager"
        if use_graph:
            try:
                gpp = torch.cuda.CUDAGraph()
                with torch.cuda.stream(stream):
                    prefill_step_split()
                    stream.synchronize()
                    dist.barrier()
                    with torch.cuda.graph(gpp, stream=stream):
                        prefill_step_split()
                dist.barrier()
                ms_pp_graph = timed(gpp.replay, 6, 2)
                if ms_pp_graph < ms_pp:
                    ms_pp, pp_launch = ms_pp_graph, "cuda_graph_replay"
                del gpp
            except Exception as e:  # noqa: BLE001
                sys.stderr.write("pp512 row-split graph capture failed: %s\n" % str(e)[:200])
        # ---- the fused form: GEMM and exchange in ONE kernel (b200_mul_mat_mmq_multi_rowsplit): the persistent tensor-core
        #      kernel's epilogue stores its tiles into every rank's dst over NVLink and counts an arrival; a stream wait-value
        #      per group replaces the collective.  Checked against the all-gathered results of the NCCL leg above.
        fused_pp = None
        nccl_pp, pp_collective = None, "dist.all_gather_into_tensor per matmul (NCCL); matmuls sharing activations in one b200_mul_mat_mmq_multi call"
        if args.collective == "fused":
            from llama_cpp_dsp_b200.rowsplit import MultiMatMulMMQRowSplit, RowSplitPush, row_split_push_join
            pbase, poff = {}, 0
            for _, (name, _, m, _) in sel:
                b_ = name.split(".")[0]
                if b_ not in pbase:
                    pbase[b_] = (poff, m)
                    poff += npp * m
            symp = PeerBuffers(1, poff, len(pg), world, rank, dev)
            delta = [symp.ptr_dst[r] - symp.ptr_dst[rank] for r in range(world)]
            fcalls, ccalls = [], []
            for gi, (_, k, members) in enumerate(pg):
                ty, pls, msl, dl = [], [], [], []
                for b_, t, rows, pl, lo in members:
                    o0, mfull = pbase[b_]
                    full = symp.dst[0, o0:o0 + npp * mfull].view(npp, mfull)
                    ty.append(t), pls.append(pl), msl.append(rows)
                    dl.append(full[:, lo:lo + rows] if rows > 0 else None)
                fcalls.append((MultiMatMulMMQRowSplit(ty, pls, msl, k, dl, npp, world, rank, delta), xpp[k], symp.ptr_sync[gi]))
                own = [j for j in range(len(ty)) if msl[j] > 0 and pls[j] is not None]
                lcall = qmm.MultiMatMulMMQ([ty[j] for j in own], [pls[j] for j in own], [msl[j] for j in own], k,
                                           [dl[j] for j in own], npp) if own else None
                ccalls.append((lcall, RowSplitPush([dl[j] for j in own], npp, world, rank, delta), xpp[k], symp.ptr_sync[gi]))

            def prefill_step_fused():
                for call, x, ps in fcalls:
                    call(x, ps)
                    row_split_wait(ps, rank, world)

            # the reference's own form, one process per GPU: local GEMM, then the slices go to the peers through the copy
            # engines (b200_rowsplit_push: cudaMemcpy2DAsync per peer on side streams + arrival counters)
            def prefill_step_copy_engines():
                for lcall, push, x, ps in ccalls:
                    if lcall is not None:
                        lcall(x)
                    push(ps)
                    row_split_wait(ps, rank, world)
                row_split_push_join()
            dist.barrier()
            prefill_step_fused()
            torch.cuda.synchronize()
            dist.barrier()
            okp = True
            lastp = {}
            for (i, (name, t, m, k)) in sel:
                lastp[name.split(".")[0]] = m
            prefill_step_split()                                  # (the NCCL leg's buffers hold the last layer's results again)
            torch.cuda.synchronize()
            for b_, m in lastp.items():
                o0, _ = pbase[b_]
                full = torch.cat([opp_f[b_][r, :, :h - l] for r, (l, h) in enumerate(splits[m].ranges)], dim=1)
                got = symp.dst[0, o0:o0 + npp * m].view(npp, m)
                # same integer sums and scales; the all-gather leg may add two k halves (split-K) where the fused launch
                # never splits, so the comparison allows fp32 summation-order differences and nothing more
                okp = okp and bool(((full - got).abs().max() <= 2e-6 * full.abs().max()).item())
            flags = [None] * world
            dist.all_gather_object(flags, bool(okp))
            if all(flags):
                ms_f_eager = timed(prefill_step_fused, 6, 2)
                ms_f, f_launch = ms_f_eager, "eager"
                if use_graph:
                    try:
                        gpf = torch.cuda.CUDAGraph()
                        with torch.cuda.stream(stream):
                            prefill_step_fused()
                            stream.synchronize()
                            dist.barrier()
                            with torch.cuda.graph(gpf, stream=stream):
                                prefill_step_fused()
                        dist.barrier()
                        ms_f_graph = timed(gpf.replay, 6, 2)
                        if ms_f_graph < ms_f:
                            ms_f, f_launch = ms_f_graph, "cuda_graph_replay"
                        del gpf
                    except Exception as e:  # noqa: BLE001
                        sys.stderr.write("pp512 fused row-split graph capture failed: %s\n" % str(e)[:200])
                # copy-engine form: checked the same way, timed eagerly (side streams)
                ce_pp = None
                symp.dst.zero_()
                torch.cuda.synchronize()
                dist.barrier()
                prefill_step_copy_engines()
                torch.cuda.synchronize()
                dist.barrier()
                okc = True
                for b_, m in lastp.items():
                    o0, _ = pbase[b_]
                    full = torch.cat([opp_f[b_][r, :, :h - l] for r, (l, h) in enumerate(splits[m].ranges)], dim=1)
                    got = symp.dst[0, o0:o0 + npp * m].view(npp, m)
                    okc = okc and bool(((full - got).abs().max() <= 2e-6 * full.abs().max()).item())
                dist.all_gather_object(flags, bool(okc))
                if all(flags):
                    ms_c = timed(prefill_step_copy_engines, 6, 2)
                    ce_pp = {"value": npp * 1e3 / ms_c, "unit": "tok/s", "ms_per_step": ms_c, "launch": "eager",
                             "api": "b200_mul_mat_mmq_multi (local slices) + b200_rowsplit_push (cudaMemcpy2DAsync to every peer on side "
                                    "streams, arrival counters) + b200_rowsplit_wait per group"}
                else:
                    sys.stderr.write("pp512 copy-engine row split: results differ on some rank: %s\n" % flags)
                nccl_pp = {"value": npp * 1e3 / ms_pp, "unit": "tok/s", "ms_per_step": ms_pp, "launch": pp_launch}
                fused_pp = {"value": npp * 1e3 / ms_f, "unit": "tok/s", "ms_per_step": ms_f, "launch": f_launch,
                            "api": "b200_mul_mat_mmq_multi_rowsplit + b200_rowsplit_wait per group (GEMM epilogue stores into every rank)"}
                if ce_pp and ce_pp["ms_per_step"] < min(ms_f, ms_pp):
                    ms_pp, pp_launch, ms_pp_eager = ce_pp["ms_per_step"], "eager", ce_pp["ms_per_step"]
                    pp_collective = "copy engines: local GEMM + b200_rowsplit_push + b200_rowsplit_wait per group"
                elif ms_f < ms_pp:
                    ms_pp, pp_launch, ms_pp_eager = ms_f, f_launch, ms_f_eager
                    pp_collective = ("fused: b200_mul_mat_mmq_multi_rowsplit (the GEMM's epilogue stores its tiles into every rank over "
                                     "NVLink and counts an arrival) + b200_rowsplit_wait per group; checked against the all-gathered results")
            else:
                sys.stderr.write("pp512 fused row split: results differ from the all-gathered ones on some rank: %s\n" % flags)
            del symp, fcalls, ccalls
        flops = sum(2.0 * m * k * npp for _, (_, _, m, k) in sel)
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            bf16_burst = float(mp["bf16_tflops"])
        except Exception:  # noqa: BLE001
            bf16_burst = 1590.0
        tops = flops / ms_pp / 1e9
        pp = {"metric": "pp512 tok/s", "value": npp * 1e3 / ms_pp, "unit": "tok/s", "ms_per_step": ms_pp, "steps": 6,
              "launch": pp_launch, "ms_per_step_eager": ms_pp_eager, "nccl_all_gather_baseline": nccl_pp,
              "fused_peer_store": fused_pp, "copy_engine_push": ce_pp if fused_pp else None,
              "collective": pp_collective,
              "all_gather_bytes_per_step_per_rank": sum(4.0 * npp * maxrows[m] * (world - 1) for _, (_, _, m, _) in sel),
              "scale_chunk_k": int(qmm.lib().b200_mmq_chunk()),
              "roofline": {"bound": "tensor", "achieved": tops, "unit": "TOP/s (int8), all GPUs", "peak": world * 2 * bf16_burst,
                           "frac": tops / (world * 2 * bf16_burst),
                           "peak_source": "n_gpus x 2 x MEASURED_PEAKS.json bf16_tflops (int8 dense = 2 x bf16)"}}
        del planes, opp_l, opp_f
        torch.cuda.empty_cache()

    # ---- MUL_MAT_ID, batched decode (BASELINE config 4): Mixtral-8x7B-shaped Q4_K experts, 2 of 8 used, ids = random
    #      permutation prefix per token (tests/test-backend-ops.cpp:2121-2140), three weight copies cycled (> L2).
    #      Algorithmic bytes = distinct experts touched x expert bytes + activations + results.
    moe = None
    if world == 1 and not args.no_pp:
        import numpy as np
        hbm_peak = peaks()[0]
        n_expert, n_used = 8, 2
        rng = np.random.default_rng(0)
        moe = {"workload": "mixtral-8x7b-shaped Q4_K experts, n_used 2 of 8, CUDA-graph replay", "cases": []}
        for label, m_e, k_e, bcast in (("ffn_gate/up_exps", 14336, 4096, True), ("ffn_down_exps", 4096, 14336, False)):
            copies = [gen_weights(torch, Q4_K, n_expert * m_e, k_e, gen, dev) for _ in range(3)]
            ebytes = m_e * row_bytes(Q4_K, k_e)
            plane_e = None
            for nt in (1, 8, 32, 64, 512):
                ne11 = 1 if bcast else n_used
                bb = torch.rand((nt, ne11, k_e), device=dev, generator=gen) * 2 - 1
                ids_np = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(nt)]).astype(np.int32)
                ids_t = torch.from_numpy(ids_np).to(dev)
                oo = torch.empty((nt, n_used, m_e), device=dev)

                use_tc = bool(qmm.lib().b200_mul_mat_id_prefers_mmq(m_e, n_expert, n_used, nt))
                if use_tc and plane_e is None:      # stacked planes of the three weight copies (+1 byte per weight each)
                    plane_e = [qmm.build_plane(Q4_K, w_, n_expert * m_e, k_e) for w_ in copies]

                def moe_step():
                    for ci, w_ in enumerate(copies):
                        if use_tc:
                            qmm.mul_mat_id_mmq(Q4_K, plane_e[ci], bb, ids_t, m_e, k_e, n_expert, out=oo)
                        else:
                            qmm.mul_mat_id(Q4_K, w_, bb, ids_t, m_e, k_e, n_expert, out=oo)
                moe_run = moe_step
                if use_graph:
                    gm = torch.cuda.CUDAGraph()
                    with torch.cuda.stream(stream):
                        moe_step()
                        stream.synchronize()
                        with torch.cuda.graph(gm, stream=stream):
                            moe_step()
                    moe_run = gm.replay
                us = timed(moe_run, 10, 3) * 1e3 / len(copies)
                distinct = len(set(ids_np.flatten().tolist()))
                algo = distinct * ebytes + bb.numel() * 4 + oo.numel() * 4
                moe["cases"].append({"tensor": label, "m": m_e, "k": k_e, "n_tokens": nt, "distinct_experts": distinct,
                                     "path": "tensor cores (b200_mul_mat_id_mmq)" if use_tc else "GEMV (b200_mul_mat_id)",
                                     "us_per_call": us, "algorithmic_gbs": algo / us / 1e3,
                                     "frac_of_hbm_peak": algo / us / 1e3 / hbm_peak})
                if use_graph:
                    del gm
            del copies, plane_e
            torch.cuda.empty_cache()

    # ---- N > 1: the SAME workload on ONE GPU (rank 0 alone, the other ranks wait), so that the line carries the
    #      row-split speed-up measured in the same run on the same box, with the same grouped calls as the N = 1 bench line
    single = None
    if world > 1 and not args.no_single:
        if rank == 0:
            del weights
            torch.cuda.empty_cache()
            full = [gen_weights(torch, t, m, k, gen, dev) for _, t, m, k in mms]
            # the same grouped calls as the N = 1 bench line (matmuls sharing activations in one b200_mul_mat_multi)
            def group_of1(name):
                base = name.split(".")[0]
                return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
            g1groups, outs1 = [], {}
            for (name, t, m, k), w in zip(mms, full):
                key = (group_of1(name), name.split(".")[-1])
                if not g1groups or g1groups[-1][0] != key:
                    g1groups.append((key, k, []))
                g1groups[-1][2].append((name.split(".")[0], t, m, w))
            calls1 = []
            for _, k, members in g1groups:
                oo = []
                for base, t, m, w in members:
                    if base not in outs1:
                        outs1[base] = torch.zeros((n, m), device=dev)
                    oo.append(outs1[base])
                calls1.append((qmm.MultiMatMul([t for _, t, _, _ in members], [w for _, _, _, w in members],
                                               [m for _, _, m, _ in members], k, oo, n), xs[k]))

            def token_step_1gpu():
                for call, x in calls1:
                    call(x)
            with torch.cuda.stream(stream):
                token_step_1gpu()
                stream.synchronize()
                g1 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g1, stream=stream):
                    token_step_1gpu()
                for _ in range(3):
                    g1.replay()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                k1 = max(8, args.steps // 4)
                e0.record(stream)
                for _ in range(k1):
                    g1.replay()
                e1.record(stream)
                torch.cuda.synchronize()
            ms1 = e0.elapsed_time(e1) / k1
            single = {"value": 1e3 / ms1, "unit": "tok/s", "ms_per_step": ms1, "steps": k1, "n_gpus": 1,
                      "speedup_of_this_run": ms1 / ms_step,
                      "hbm_gbs": total_bytes_all_ranks / ms1 / 1e6,
                      "note": "same workload, all rows on rank 0's GPU, grouped calls as at N = 1, CUDA-graph replay, measured in this run"}
            del full, g1, calls1, outs1
            torch.cuda.empty_cache()
            weights = []
        dist.barrier()

    # ---- cpu baseline (rank 0, N = 1 only): the unmodified reference ggml-cpu on this host
    cpu = None
    if world == 1 and not args.no_cpu:
        threads = host_threads()
        try:
            probe = run_reference(model, 2, 1, threads)
            steps_cpu = int(max(4, min(200, 15e3 / probe["ms_per_step"])))
            r = run_reference(model, steps_cpu, 1, threads)
            cpu = {"value": 1e3 / r["ms_per_step"], "unit": "tok/s", "cores": threads, "kind": "reference",
                   "sample": "%d timed tokens (1 warm-up), each all %d MUL_MATs of %s on ggml-cpu, 3 weight copies per "
                             "shape cycled" % (steps_cpu, r["nodes"], model),
                   "ms_per_step": r["ms_per_step"], "weight_gbs": r["weight_gbs"]}
        except Exception as e:  # noqa: BLE001
            cpu = {"value": None, "unit": "tok/s", "cores": threads, "kind": "reference", "sample": "failed: " + str(e)[:160]}

    # ---- N = 1 headline run: BASELINE configs[1] (llama3-8b-q4_k_m: tg128, pp512, e2e, MUL_MAT_ID) as a sub-run
    sub8b = None
    if with_sub:
        try:
            cmd = [sys.executable, os.path.abspath(__file__), "--workload", "llama3-8b-q4_k_m", "--sub", "--steps",
                   str(args.steps), "--warmup", str(args.warmup)] + (["--no-cpu"] if args.no_cpu else [])
            out = subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900).stdout
            full = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
            sub8b = {k: full.get(k) for k in ("metric", "value", "unit", "ms_per_step", "ms_per_step_eager", "roofline",
                                               "cpu_baseline", "e2e", "e2e_ggml", "sequence", "per_node_calls", "pp512",
                                               "mul_mat_id", "gpu_launches", "clocks", "config")}
        except Exception as e:  # noqa: BLE001
            sub8b = {"error": str(e)[:300]}

    if rank == 0:
        peak, how = peaks()
        # DRAM bytes of all GEMV launches of one step, from the committed ncu pass over the same workload (None otherwise)
        traffic, traffic_src = None, None
        try:
            for tf in ("profiles/r02/decode_step_dram_traffic_%s.json" % model, "profiles/r01/decode_step_dram_traffic.json"):
                if not os.path.exists(os.path.join(ROOT, tf)):
                    continue
                tj = json.load(open(os.path.join(ROOT, tf)))
                if world == 1 and tj["algorithmic_bytes_per_step"] == step_bytes:
                    traffic = tj["gemv_dram_bytes_read"] + tj["gemv_dram_bytes_write"]
                    traffic_src = "ncu dram__bytes_read.sum + dram__bytes_write.sum over the step's GEMV launches (%s)" % tf
                    break
        except Exception:  # noqa: BLE001
            pass
        achieved = step_bytes / ms_step / 1e6   # GB/s per GPU (rank 0's slice)
        line = {
            "metric": metric_name(model), "value": 1e3 / ms_step, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": {"workload": model + (" decode, split-mode row over %d GPUs, %s" % (world, collective) if world > 1
                                            else " decode"),
                       "matmuls_per_step": len(mms), "batch": n, "l2": "inputs_exceed_l2 (%.2f GB of weights streamed "
                       "per step per GPU)" % (weight_bytes_local / 1e9),
                       "launch": "cuda_graph_replay" if use_graph else "eager", "calls": grouping},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": how, "kernel": "k_gemv (all GEMV launches of the step)",
                         "algorithmic_bytes_per_step": step_bytes, "launches_per_step": launches_per_step},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
            "ms_per_step_eager": ms_eager,
            "per_node_calls": per_node,
            "sequence": sequence,
            "sequence_row_split": seq_split,
            "e2e_ggml": e2e_ggml,
            "llama3_8b_q4_k_m": sub8b,
            "pp512": pp,
            "mul_mat_id": moe,
            "single_gpu_same_workload": single,
            "nccl_all_gather_baseline": nccl_leg,
            "fused_row_split": fused_info if nccl_leg else None,
            "algorithmic_bytes_per_token_all_gpus": total_bytes_all_ranks,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        # CUDA graphs that captured NCCL collectives make destroy_process_group hang at exit (seen on the 2-GPU box):
        # every rank has printed / synchronized, so leave without the teardown
        os._exit(0)


if __name__ == "__main__":
    main()
