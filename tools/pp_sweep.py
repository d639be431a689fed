"""pp512 sweep over env knobs in one process: the 224 per-layer matmuls of a Llama-3-8B-shaped Q4_K_M prefill at n=512
through b200_mul_mat_mmq_multi per src1 group (cached planes), CUDA-graph replay.
usage: python tools/pp_sweep.py KNOB=v,KNOB=v ...   (each argument is one configuration; LAYERS=n limits the model)"""
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
    configs = [dict(kv.split("=") for kv in a.split(",") if kv) for a in sys.argv[1:]] or [{}]
    layers = int(os.environ.get("LAYERS", "32"))
    npp = int(os.environ.get("NPP", "512"))
    dev = torch.device("cuda", 0)
    mms = bench.matmul_list("llama3-8b-q4_k_m")[:-1][:7 * layers]
    planes = []
    for name, t, m, k in mms:
        g = torch.Generator(device=dev)
        g.manual_seed(zlib.crc32(("%s/0" % name).encode()))
        w = bench.gen_weights(torch, t, m, k, g, dev)
        planes.append(qmm.build_plane(t, w, m, k))
        del w
    xpp = {k: torch.rand((npp, k), device=dev) * 2 - 1 for k in {k for _, _, _, k in mms}}

    def pgroup(name):
        base = name.split(".")[0]
        return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)
    groups = []
    for (name, t, m, k), pl in zip(mms, planes):
        key = (pgroup(name), name.split(".")[-1])
        if not groups or groups[-1][0] != key:
            groups.append((key, k, []))
        groups[-1][2].append((name.split(".")[0], t, m, pl))
    out_node, calls = {}, []
    for _, k, members in groups:
        outs = []
        for base, t, m, pl in members:
            out_node.setdefault(base, torch.empty((npp, m), device=dev))
            outs.append(out_node[base])
        calls.append((qmm.MultiMatMulMMQ([t for _, t, _, _ in members], [pl for _, _, _, pl in members],
                                         [m for _, _, m, _ in members], k, outs, npp), xpp[k]))
    flops = sum(2.0 * m * k * npp for _, _, m, k in mms)
    stream = torch.cuda.Stream(device=dev)
    ref = None
    for cfg in configs:
        for kn in list(os.environ):
            if kn.startswith("B200_MMQ"):
                os.environ.pop(kn)
        for kn, v in cfg.items():
            os.environ[kn] = v
        qmm.reload_env()

        def step():
            for call, x in calls:
                call(x)
        with torch.cuda.stream(stream):
            step()
            stream.synchronize()
            res = torch.cat([o.flatten()[:4096] for o in out_node.values()]).clone()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                step()
            for _ in range(2):
                g.replay()
            torch.cuda.synchronize()
            reps = int(os.environ.get("REPS", "8"))
            with bench.ClockSampler(0) as clk:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(reps):
                    g.replay()
                e1.record(stream)
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            try:
                watts = clk.nv.nvmlDeviceGetPowerUsage(clk.h) / 1e3
            except Exception:  # noqa: BLE001
                watts = float("nan")
            cs = clk.summary()
            clocks = "sm %s MHz (min %s) %s %.0f W" % (cs["sm_mhz"], min(clk.samples) if clk.samples else None, cs["reasons"], watts)
        same = "ref" if ref is None else ("bit-identical" if torch.equal(res, ref) else "DIFFERS max %g" % float((res - ref).abs().max()))
        if ref is None:
            ref = res
        print("%-50s %7.3f ms  %8.0f tok/s  %7.1f TOP/s  [%s]  %s" % (cfg, ms, npp * 1e3 / ms, flops / ms / 1e9, same, clocks), flush=True)
        del g


if __name__ == "__main__":
    main()
