"""Phase timeline of the sequence kernel (b200_sequence_run) on the 8B decode token, from in-kernel globaltimer stamps:
per group and CTA  0 boundary reached, 1 earlier groups complete everywhere, 2 activations quantized, 3 first weights in
hand, 4 team 0 done.  usage: python tools/seq_trace.py [first_group] [ngroups_to_print] [model]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200pkg  # noqa: E402
import bench  # noqa: E402

qmm = b200pkg.load()
first = int(sys.argv[1]) if len(sys.argv) > 1 else 40
count = int(sys.argv[2]) if len(sys.argv) > 2 else 12
model = sys.argv[3] if len(sys.argv) > 3 else "llama3-8b-q4_k_m"
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev)
gen.manual_seed(1)
mms = bench.matmul_list(model)
xs = {k: torch.rand((1, k), device=dev, generator=gen) * 2 - 1 for k in sorted({k for _, _, _, k in mms})}


def group_of(name):
    base = name.split(".")[0]
    return {"attn_q": "qkv", "attn_k": "qkv", "attn_v": "qkv", "ffn_gate": "gu", "ffn_up": "gu"}.get(base, base)


groups, names = [], []
for name, t, m, k in mms:
    key = (group_of(name), name.split(".")[-1])
    if not names or names[-1] != key:
        names.append(key)
        groups.append((xs[k], k, [], True))
    groups[-1][2].append((t, bench.gen_weights(torch, t, m, k, gen, dev), m, torch.zeros(m, device=dev)))
seq = qmm.Sequence(groups)
ng, ncta = len(groups), qmm.lib().b200_sm_count(0)
trace = torch.zeros((ng, ncta, 8), dtype=torch.int64, device=dev)
for _ in range(3):
    seq()
torch.cuda.synchronize()
qmm.check(qmm.lib().b200_debug_sequence_trace(seq.plan.data_ptr(), trace.data_ptr(), None))
seq()
seq()
torch.cuda.synchronize()
qmm.check(qmm.lib().b200_debug_sequence_trace(seq.plan.data_ptr(), None, None))
tr = trace.cpu().numpy().astype(np.int64)
t0 = tr[first][:, 0][tr[first][:, 0] > 0].min()
print("group  kind       MB   | boundary reached (min/med/max) | +others done | +quantized | +first weights | +team0 done | period")
prev = None
for g in range(first, min(ng, first + count)):
    a = tr[g]
    ok = a[:, 0] > 0
    a = a[ok]
    mb = sum(bench.gemv_bytes(t, m, groups[g][1], 1) for t, _, m, _ in groups[g][2]) / 1e6
    b0 = a[:, 0] - t0
    d = [np.median(a[:, i + 1] - a[:, i]) for i in range(4)]
    end = (a[:, 4] - t0).max()
    print("%4d  %-8s %6.1f | %7d %7d %7d | %8d | %8d | %8d | %8d | %s" % (
        g, names[g][0], mb, b0.min(), int(np.median(b0)), b0.max(), d[0], d[1], d[2], d[3],
        "" if prev is None else "%d  (%.0f GB/s)" % (end - prev, mb * 1e6 / max(1, end - prev))))
    prev = end
print("finer: group | team0 done -> all warps at the barrier (6-4) | -> published (7-6) | -> next group's boundary stamp (0'-7) | "
      "quantized -> units in registers (5-2) | -> first stage in hand (3-5)")
for g in range(first, min(ng - 1, first + count)):
    a, b = tr[g], tr[g + 1]
    ok = (a[:, 0] > 0) & (b[:, 0] > 0) & (a[:, 6] > 0)
    a, b = a[ok], b[ok]
    print("%4d %-8s | %6d | %6d | %6d | %6d | %6d" % (g, names[g][0], np.median(a[:, 6] - a[:, 4]), np.median(a[:, 7] - a[:, 6]),
                                                    np.median(b[:, 0] - a[:, 7]), np.median(a[:, 5] - a[:, 2]), np.median(a[:, 3] - a[:, 5])))
tot = (tr[ng - 1][:, 4].max() - tr[0][:, 0][tr[0][:, 0] > 0].min()) / 1e3
print("first boundary -> last team done: %.1f us for %d groups" % (tot, ng))
