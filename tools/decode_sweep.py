"""Whole-token decode sweep over env knobs, in one process (weights generated once): every quantized matmul of a
Llama-3-8B-shaped Q4_K_M token through b200_mul_mat_multi per src1 group, CUDA-graph replay; results are checked bit for bit against the first
configuration.  B200_QMM_LIB selects an A/B build of the library.
usage: python tools/decode_sweep.py [model] KNOB=v,KNOB=v ...   (each argument after the model is one configuration)"""
import os
import sys
import zlib

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200pkg  # noqa: E402
import bench  # noqa: E402

qmm = b200pkg.load()


def main():
    args = sys.argv[1:]
    model = "llama3-8b-q4_k_m"
    if args and args[0] in bench.MODELS:
        model = args.pop(0)
    configs = [dict(kv.split("=") for kv in a.split(",") if kv) for a in args] or [{}]
    dev = torch.device("cuda", 0)
    mms = bench.matmul_list(model)
    weights = []
    for name, t, m, k in mms:
        g = torch.Generator(device=dev)
        g.manual_seed(zlib.crc32(("%s/0" % name).encode()))
        weights.append(bench.gen_weights(torch, t, m, k, g, dev))
    xs = {k: torch.rand((1, k), device=dev) * 2 - 1 for k in {k for _, _, _, k in mms}}

    def group_of(name):
        base = name.split(".")[0]
        return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
    groups = []
    for (name, t, m, k), w in zip(mms, weights):
        key = (group_of(name), name.split(".")[-1])
        if not groups or groups[-1][0] != key:
            groups.append((key, k, []))
        groups[-1][2].append((name.split(".")[0], t, m, w))
    out_node, calls = {}, []
    for _, k, members in groups:
        outs = []
        for base, t, m, w in members:
            out_node.setdefault(base, torch.zeros((1, m), device=dev))
            outs.append(out_node[base])
        calls.append((qmm.MultiMatMul([t for _, t, _, _ in members], [w for _, _, _, w in members],
                                      [m for _, _, m, _ in members], k, outs, 1), xs[k]))
    step_bytes = sum(bench.gemv_bytes(t, m, k, 1) for _, t, m, k in mms)
    stream = torch.cuda.Stream(device=dev)
    ref = None
    for cfg in configs:
        knobs = dict(cfg)
        for kn in list(os.environ):
            if kn.startswith("B200_GEMV") or kn in ("B200_PDL",):
                os.environ.pop(kn)
        for kn, v in knobs.items():
            os.environ[kn] = v
        qmm.reload_env()

        def step():
            for call, x in calls:
                call(x)
        with torch.cuda.stream(stream):
            step()
            stream.synchronize()
            res = torch.cat([o.flatten() for o in out_node.values()]).clone()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                step()
            for _ in range(5):
                g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(40):
                g.replay()
            e1.record(stream)
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 40
        same = "ref" if ref is None else ("bit-identical" if torch.equal(res, ref) else "DIFFERS max %g" % float((res - ref).abs().max()))
        if ref is None:
            ref = res
        print("%-60s %7.3f ms  %7.1f tok/s  %7.1f GB/s  [%s]" % (cfg, ms, 1e3 / ms, step_bytes / ms / 1e6, same), flush=True)
        del g


if __name__ == "__main__":
    main()
