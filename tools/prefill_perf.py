"""2k-token prefill of the synthetic Llama-3-8B through the product path (BASELINE.json configs[2]: "decode + 2k prefill"):
M = 2048 GEMMs, rotary, reshape_and_cache, prefill attention over the paged cache, last-token logits.  Prints one JSON
line per quantisation: ms per prefill, tokens/s, the M = 2048 GEMM share timed alone, attention share timed alone."""
import json
import sys

import torch

sys.path.insert(0, ".")
from nm_vllm_b200 import decoder, load_extension  # noqa: E402

load_extension()
DEV = "cuda:0"
T = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
quants = sys.argv[2].split(",") if len(sys.argv) > 2 else ["w4a16", "w4a16_24"]


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


for quant in quants:
    dtype = torch.float16 if quant == "w4a16_24" else torch.bfloat16
    m = decoder.SyntheticLlama(decoder.LLAMA3_8B, quant=quant, dtype=dtype, device=DEV, num_blocks=T // 16 + 8, max_pos=T + 16)
    toks = torch.randint(0, m.shape.vocab, (T,), device=DEV)
    bt = torch.randperm(T // 16 + 8)[:T // 16].to(torch.int32).view(1, -1)
    pos, meta = m.prefill_meta([T], [0], bt)
    ms = timed(lambda: m.forward(toks, pos, meta))
    # shares, each timed alone over the 32 layers
    x = {K: torch.randn((T, K), device=DEV, dtype=dtype) for K in (4096, 14336)}
    lin = [L[k] for L in m.layers for k in ("qkv", "o", "gate_up", "down")]
    gemm_ms = timed(lambda: [q(x[q.K]) for q in lin])
    flops = 2.0 * T * sum(q.K * q.N for q in lin)
    from nm_vllm_b200.attention import PagedAttention
    q = torch.randn((T, m.H, 128), device=DEV, dtype=dtype)
    k = torch.randn((T, m.Hkv, 128), device=DEV, dtype=dtype)
    kc, vc = m.kv[0]
    attn_ms = timed(lambda: [PagedAttention.forward_prefix(q, k, k, kc, vc, meta["block_tables"], meta["query_start_loc"],
                                                           meta["seq_lens"], meta["context_lens"], T, None, None)
                             for _ in range(m.nlayers)])
    attn_flops = m.nlayers * 4.0 * m.H * 128 * T * T / 2
    print(json.dumps({"workload": "llama-3-8b %s prefill, 1 x %d tokens" % (quant, T), "ms_per_prefill": ms,
                      "prefill_tok_s": T / ms * 1e3, "gemm_ms_alone": gemm_ms, "gemm_tflops": flops / gemm_ms / 1e9,
                      "attention_ms_alone": attn_ms, "attention_tflops_causal": attn_flops / attn_ms / 1e9}), flush=True)
    del m
    torch.cuda.empty_cache()
