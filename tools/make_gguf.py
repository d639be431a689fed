"""Random-init GGUF of a Llama-shaped model for the reference's llama-bench / llama-cli (BASELINE configs 1, 2, 5): valid
blocks written directly in the quantized formats (random payload, sane fp16 scales -- the generator of tests/qref.py),
the Q4_K_M tensor-type recipe of src/llama-quant.cpp:129-131,166-168,235-236 (or plain Q4_0), a synthetic SPM vocabulary.
There is no network for real checkpoints; llama-bench feeds random token ids anyway.

  python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf [--model llama3-8b-q4_k_m] [--layers N]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gguf  # noqa: E402

Q4_0, Q4_K, Q6_K = 2, 12, 14
BLOCK = {Q4_0: 32, Q4_K: 256, Q6_K: 256}
TSIZE = {Q4_0: 18, Q4_K: 144, Q6_K: 210}
GTYPE = {Q4_0: gguf.GGMLQuantizationType.Q4_0, Q4_K: gguf.GGMLQuantizationType.Q4_K, Q6_K: gguf.GGMLQuantizationType.Q6_K}
MODELS = {
    "llama3-8b-q4_k_m": dict(n_layer=32, n_embd=4096, n_ff=14336, n_head=32, n_head_kv=8, vocab=128256, recipe="q4_k_m",
                             rope_base=500000.0),
    "llama3-70b-q4_k_m": dict(n_layer=80, n_embd=8192, n_ff=28672, n_head=64, n_head_kv=8, vocab=128256, recipe="q4_k_m",
                              rope_base=500000.0),
    "llama2-7b-q4_0": dict(n_layer=32, n_embd=4096, n_ff=11008, n_head=32, n_head_kv=32, vocab=32000, recipe="q4_0",
                           rope_base=10000.0),
}


def use_more_bits(i, n):
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def random_blocks(t, nrows, k, rng, scale=0.02):
    nb = nrows * (k // BLOCK[t])
    raw = rng.integers(0, 256, size=(nb, TSIZE[t]), dtype=np.uint8)

    def put_f16(col, vals):
        raw[:, col:col + 2] = np.asarray(vals, dtype=np.float16).view(np.uint8).reshape(nb, 2)

    sgn = rng.choice([-1.0, 1.0], size=nb)
    if t == Q4_0:
        put_f16(0, sgn * rng.uniform(0.1, 1.0, nb) * scale / 4)
    elif t == Q4_K:
        put_f16(0, rng.uniform(0.1, 1.0, nb) * scale / 200)
        put_f16(2, rng.uniform(0.1, 1.0, nb) * scale / 60)
    elif t == Q6_K:
        put_f16(208, sgn * rng.uniform(0.1, 1.0, nb) * scale / 2000)
    return raw.reshape(nrows, -1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--model", default="llama3-8b-q4_k_m", choices=list(MODELS))
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (smaller file for smoke runs)")
    args = ap.parse_args()
    c = dict(MODELS[args.model])
    if args.layers:
        c["n_layer"] = args.layers
    L, E, F, V = c["n_layer"], c["n_embd"], c["n_ff"], c["vocab"]
    KV = E // c["n_head"] * c["n_head_kv"]
    rng = np.random.default_rng(1234)
    w = gguf.GGUFWriter(args.out, "llama")
    w.add_name(args.model + " (random-init, synthetic)")
    w.add_block_count(L)
    w.add_context_length(8192)
    w.add_embedding_length(E)
    w.add_feed_forward_length(F)
    w.add_head_count(c["n_head"])
    w.add_head_count_kv(c["n_head_kv"])
    w.add_rope_dimension_count(E // c["n_head"])
    w.add_rope_freq_base(c["rope_base"])
    w.add_layer_norm_rms_eps(1e-5)
    w.add_vocab_size(V)
    w.add_file_type(15 if c["recipe"] == "q4_k_m" else 2)      # LLAMA_FTYPE_MOSTLY_Q4_K_M / MOSTLY_Q4_0
    w.add_quantization_version(2)
    # synthetic SentencePiece-style vocabulary: <unk>, <s>, </s>, 256 byte tokens, then dummy pieces
    toks = [b"<unk>", b"<s>", b"</s>"] + [b"<0x%02X>" % i for i in range(256)] + [b"\xe2\x96\x81t%d" % i for i in range(V - 259)]
    types = [2, 3, 3] + [6] * 256 + [1] * (V - 259)                    # UNKNOWN, CONTROL, CONTROL, BYTE..., NORMAL...
    w.add_tokenizer_model("llama")
    w.add_token_list(toks)
    w.add_token_scores([0.0] * 259 + [-float(i) for i in range(V - 259)])
    w.add_token_types(types)
    w.add_bos_token_id(1)
    w.add_eos_token_id(2)
    w.add_unk_token_id(0)
    w.add_add_bos_token(True)
    w.add_add_eos_token(False)

    images = {}

    def blocks(t, rows, k):
        key = (t, rows, k)
        if key not in images:                                          # one random image per distinct (type, shape)
            images[key] = random_blocks(t, rows, k, rng)
        return images[key]

    def add_q(name, t, rows, k):
        w.add_tensor(name, blocks(t, rows, k), raw_dtype=GTYPE[t])

    def add_f32(name, n):
        w.add_tensor(name, np.ones(n, dtype=np.float32))

    base = Q4_K if c["recipe"] == "q4_k_m" else Q4_0
    add_q("token_embd.weight", base, V, E)
    for i in range(L):
        hi = (Q6_K if use_more_bits(i, L) else Q4_K) if c["recipe"] == "q4_k_m" else Q4_0
        add_f32("blk.%d.attn_norm.weight" % i, E)
        add_q("blk.%d.attn_q.weight" % i, base, E, E)
        add_q("blk.%d.attn_k.weight" % i, base, KV, E)
        add_q("blk.%d.attn_v.weight" % i, hi, KV, E)
        add_q("blk.%d.attn_output.weight" % i, base, E, E)
        add_f32("blk.%d.ffn_norm.weight" % i, E)
        add_q("blk.%d.ffn_gate.weight" % i, base, F, E)
        add_q("blk.%d.ffn_up.weight" % i, base, F, E)
        add_q("blk.%d.ffn_down.weight" % i, hi, E, F)
    add_f32("output_norm.weight", E)
    add_q("output.weight", Q6_K if c["recipe"] == "q4_k_m" else Q4_0, V, E)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    print("wrote %s: %.2f GB, %d layers" % (args.out, os.path.getsize(args.out) / 1e9, L))


if __name__ == "__main__":
    main()
