"""ORACLE / test infrastructure (never on the product path): BASELINE.json configs[0] -- OPT-125m geometry, random-init,
batch 8 x 128-token prompt / 128 decode tokens -- on HOST cores, as the reference's CPU backend runs it: torch linears
(vllm/executor/cpu_executor.py:19-155 only runs unquantized bf16 models), torch SDPA for the prompt
(vllm/attention/backends/torch_sdpa.py:140-230), and the reference's own csrc/cpu paged_attention_v1 /
reshape_and_cache (oracle/_ref, compiled unmodified) for the cache write and the decode steps (falls back to the torch
restatement when that library is not loadable).  With --w4a16 the linears are the "torch dequant W4A16 g128 linear
reference" the config names: (q - 8) * s dequantised per call, then matmul.  Usage: python -m oracle.cpu_opt125m"""
import argparse
import json
import os
import time

import torch
import torch.nn.functional as F

from . import ops_ref, refcpu
from .cpu_decode import _attention_torch

HID, FFN, LAYERS, HEADS, D, VOCAB, BS = 768, 3072, 12, 12, 64, 50272, 16


class Lin:
    def __init__(self, k, n, g, w4, dtype):
        self.w4 = w4
        if w4:
            self.q = torch.randint(0, 16, (k, n), generator=g, dtype=torch.uint8)
            self.s = (torch.rand((k // 128, n), generator=g) * 0.002 + 0.003).to(dtype)
        else:
            self.w = (torch.randn((k, n), generator=g) * 0.02).to(dtype)

    def __call__(self, x):
        if self.w4:   # dequant-then-matmul (tests/kernels/test_marlin_gemm.py:172-179)
            k, n = self.q.shape
            w = ((self.q.to(self.s.dtype) - 8).view(k // 128, 128, n) * self.s.view(k // 128, 1, n)).view(k, n)
            return torch.matmul(x, w)
        return torch.matmul(x, self.w)


def run(batch=8, prompt=128, decode=128, w4=False, dtype=torch.bfloat16, threads=None, seed=0):
    if threads:
        torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(seed)
    layers = [dict(qkv=Lin(HID, 3 * HID, g, w4, dtype), o=Lin(HID, HID, g, w4, dtype), fc1=Lin(HID, FFN, g, w4, dtype),
                   fc2=Lin(FFN, HID, g, w4, dtype), ln1=torch.ones(HID, dtype=dtype), ln2=torch.ones(HID, dtype=dtype))
              for _ in range(LAYERS)]
    embed = (torch.randn((VOCAB, HID), generator=g) * 0.02).to(dtype)
    pos_emb = (torch.randn((prompt + decode + 2, HID), generator=g) * 0.02).to(dtype)
    total = prompt + decode
    nb_seq = (total + BS - 1) // BS
    NB = batch * nb_seq
    kcs = [torch.zeros((NB, HEADS, D // 8, BS, 8), dtype=dtype) for _ in range(LAYERS)]
    vcs = [torch.zeros((NB, HEADS, D, BS), dtype=dtype) for _ in range(LAYERS)]
    bt = torch.arange(NB, dtype=torch.int32).view(batch, nb_seq)
    use_ref = refcpu.available() and dtype in (torch.bfloat16, torch.float32)
    scale = float(D ** -0.5)

    def write_cache(li, k, v, positions):   # k, v [tokens, HEADS, D]; positions [tokens] (seq-major)
        seq = torch.arange(batch).repeat_interleave(positions.numel() // batch)
        slots = bt[seq, (positions // BS).long()].long() * BS + positions % BS
        if use_ref:
            refcpu.reshape_and_cache(k, v, kcs[li], vcs[li], slots)
        else:
            ops_ref.reshape_and_cache(k, v, kcs[li], vcs[li], slots)

    def layer_fwd(li, L, h, positions, ctx_len, is_prompt):
        x = F.layer_norm(h, (HID,), L["ln1"])
        q, k, v = L["qkv"](x).split([HID, HID, HID], dim=-1)
        T = h.shape[0]
        q, k, v = q.view(T, HEADS, D), k.reshape(T, HEADS, D).contiguous(), v.reshape(T, HEADS, D).contiguous()
        write_cache(li, k, v, positions)
        if is_prompt:
            qb, kb, vb = (t.view(batch, -1, HEADS, D).transpose(1, 2) for t in (q, k, v))
            a = F.scaled_dot_product_attention(qb, kb, vb, is_causal=True).transpose(1, 2).reshape(T, HID)
        else:
            lens = torch.full((batch,), ctx_len, dtype=torch.int32)
            if use_ref:
                a = torch.empty_like(q)
                refcpu.paged_attention_v1(a, q.contiguous(), kcs[li], vcs[li], HEADS, scale, bt, lens, BS, ctx_len)
            else:
                a = _attention_torch(q, kcs[li], vcs[li], bt, lens, scale)
            a = a.reshape(T, HID)
        h = h + L["o"](a)
        x = F.layer_norm(h, (HID,), L["ln2"])
        return h + L["fc2"](F.relu(L["fc1"](x)))

    def step(tokens, positions, ctx_len, is_prompt):
        h = embed[tokens] + pos_emb[positions]
        for li, L in enumerate(layers):
            h = layer_fwd(li, L, h, positions, ctx_len, is_prompt)
        last = h.view(batch, -1, HID)[:, -1]
        return torch.argmax(torch.matmul(last, embed.t()), dim=-1)

    toks = torch.randint(0, VOCAB, (batch, prompt), generator=g)
    t0 = time.perf_counter()
    nxt = step(toks.reshape(-1), torch.arange(prompt).repeat(batch), prompt, True)
    t_prefill = time.perf_counter() - t0
    t0 = time.perf_counter()
    for i in range(decode):
        pos = torch.full((batch,), prompt + i, dtype=torch.long)
        nxt = step(nxt, pos, prompt + i + 1, False)
    t_decode = time.perf_counter() - t0
    return {"config": "OPT-125m geometry random-init, batch %d x %d prompt / %d decode, %s, %s linears" %
                      (batch, prompt, decode, str(dtype).replace("torch.", ""), "torch dequant W4A16 g128" if w4 else "bf16"),
            "prefill_s": t_prefill, "decode_s": t_decode, "decode_tok_s": batch * decode / t_decode,
            "total_tok_s": batch * (prompt + decode) / (t_prefill + t_decode), "threads": torch.get_num_threads(),
            "host_cpu_count": os.cpu_count(),
            "attention": "reference csrc/cpu ops (oracle/_ref)" if use_ref else "torch restatement"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--w4a16", action="store_true")
    ap.add_argument("--decode", type=int, default=128)
    a = ap.parse_args()
    print(json.dumps(run(w4=a.w4a16, decode=a.decode, threads=os.cpu_count())))
