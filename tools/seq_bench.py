"""8B decode token: ONE persistent sequence launch (b200_sequence_run) vs grouped calls (b200_mul_mat_multi), both as
CUDA-graph replays over the same weights; checks bit-identity first.  usage: python tools/seq_bench.py [model] [steps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200pkg  # noqa: E402
import bench  # noqa: E402

qmm = b200pkg.load()
model = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b-q4_k_m"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev)
gen.manual_seed(1)
mms = bench.matmul_list(model)
ks = sorted({k for _, _, _, k in mms})
xs = {k: torch.rand((1, k), device=dev, generator=gen) * 2 - 1 for k in ks}


def group_of(name):
    base = name.split(".")[0]
    return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)


groups = []
for name, t, m, k in mms:
    key = (group_of(name), name.split(".")[-1])
    if not groups or groups[-1][0] != key:
        groups.append((key, k, []))
    groups[-1][2].append((name.split(".")[0], t, m, bench.gen_weights(torch, t, m, k, gen, dev)))
out_a, out_b, calls, seq_groups = {}, {}, [], []
for _, k, members in groups:
    oa, ob = [], []
    for base, t, m, w in members:
        out_a.setdefault(base, torch.zeros((1, m), device=dev))
        out_b.setdefault(base, torch.zeros((1, m), device=dev))
        oa.append(out_a[base]), ob.append(out_b[base])
    calls.append((qmm.MultiMatMul([t for _, t, _, _ in members], [w for _, _, _, w in members],
                                  [m for _, _, m, _ in members], k, oa, 1), xs[k]))
    seq_groups.append((xs[k], k, [(t, w, m, o) for (_, t, m, w), o in zip(members, ob)], os.environ.get("SEQ_INDEPENDENT") is None))
seq = qmm.Sequence(seq_groups)
step_bytes = sum(bench.gemv_bytes(t, m, k, 1) for _, t, m, k in mms)


def grouped():
    for c, x in calls:
        c(x)


grouped()
seq()
torch.cuda.synchronize()
print("sequence status:", seq.status(), " groups:", len(seq_groups))
bad = [b for b in out_a if not torch.equal(out_a[b], out_b[b])]
print("bit-identical to grouped calls:", not bad, bad)
stream = torch.cuda.Stream()


def timed(fn, n):
    with torch.cuda.stream(stream):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def graphed(fn):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        fn()
        stream.synchronize()
        with torch.cuda.graph(g, stream=stream):
            fn()
    return g.replay


peak = bench.peaks()[0]
for label, fn in (("grouped calls, graph replay", graphed(grouped)), ("sequence, eager launch", seq),
                  ("sequence, graph replay", graphed(seq))):
    ms = timed(fn, steps)
    print("%-30s %.4f ms/token  %7.1f tok/s  %6.1f GB/s  %.3f of %.0f GB/s" % (label, ms, 1e3 / ms, step_bytes / ms / 1e6,
                                                                              step_bytes / ms / 1e6 / peak, peak))
print("sequence status after timing:", seq.status())
bad = [b for b in out_a if not torch.equal(out_a[b], out_b[b])]
print("bit-identical after timing:", not bad, bad)
