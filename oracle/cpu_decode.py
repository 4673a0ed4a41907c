"""ORACLE / test infrastructure (never on the product path): the decode step on HOST cores, used only by bench.py's
`cpu_baseline` leg and `--impl reference` arm, and by tests as a checker of the GPU decoder.

What is timed is the reference's own CPU arithmetic for the path:
  * linears: `torch.matmul(a, w_ref)` on the dequantised weights -- the oracle the reference's tests use for
    gptq_marlin_gemm (tests/kernels/test_marlin_gemm.py:172-179; w_ref = (q - 8) * s, quant_utils.py:39-106).  The
    reference has no quantized CPU linear (vllm/executor/cpu_executor.py only runs unquantized models), so this is the
    most favourable CPU statement of it: weights already dequantised to bf16, nothing unpacked per call;
  * attention + cache write: the reference's csrc/cpu/attention.cpp / cache.cpp compiled unmodified
    (oracle/_ref/_refcpu_C.so, oracle/Makefile) when this host can load it (kind "reference"), else a torch restatement
    (kind "port");
  * rmsnorm / rotary / silu_and_mul: torch restatements from oracle/ops_ref.py.
One layer (+ the lm_head once) is timed and scaled to the model's layer count: the layers are identical in shape, the
sample keeps the CPU leg within seconds.
"""
import time

import torch

from . import ops_ref, refcpu


def _attention_torch(q, kc, vc, block_tables, seq_lens, scale):
    """Vectorised restatement of ref_single_query_cached_kv_attention (tests/kernels/test_attention.py:63-116) for
    equal-length sequences: gather blocks through the table, fp32 softmax, P rounded to the value dtype before P.V."""
    S, H, D = q.shape
    NB, Hkv, Dx, BS, x = kc.shape
    L = int(seq_lens[0])
    nb = (L + BS - 1) // BS
    bt = block_tables[:, :nb].long()
    k = kc[bt]                                         # [S, nb, Hkv, D/x, BS, x]
    k = k.permute(0, 2, 1, 4, 3, 5).reshape(S, Hkv, nb * BS, D)[:, :, :L]
    v = vc[bt].permute(0, 2, 1, 4, 3).reshape(S, Hkv, nb * BS, D)[:, :, :L]
    G = H // Hkv
    qf = q.float().view(S, Hkv, G, D)
    logits = torch.einsum("skgd,sktd->skgt", qf, k.float()) * scale
    p = torch.softmax(logits, dim=-1).to(v.dtype).float()
    out = torch.einsum("skgt,sktd->skgd", p, v.float())
    return out.reshape(S, H, D).to(q.dtype)


class CpuDecodeLayer:
    """One decoder layer (+ lm_head) of `shape` on CPU with random dequantised weights."""

    def __init__(self, shape, batch, ctx, dtype=torch.bfloat16, block_size=16, num_blocks=None, seed=0, tp=1):
        s = shape
        self.s, self.B, self.ctx, self.dtype, self.BS = s, batch, ctx, dtype, block_size
        g = torch.Generator().manual_seed(seed)
        H, Hkv, D = s.heads // tp, max(1, s.kv_heads // tp), s.head_dim
        self.H, self.Hkv, self.D = H, Hkv, D
        inter = s.inter // tp

        def w(k, n):
            return (torch.randn((k, n), generator=g) * 0.02).to(dtype)

        self.w_qkv, self.w_o = w(s.hidden, (H + 2 * Hkv) * D), w(H * D, s.hidden)
        self.w_gu, self.w_down = w(s.hidden, 2 * inter), w(inter, s.hidden)
        self.ln = torch.ones(s.hidden, dtype=dtype)
        nb_seq = (ctx + block_size - 1) // block_size
        num_blocks = num_blocks or max(batch * nb_seq, 64)
        self.kc = ((torch.rand((num_blocks, Hkv, D // 8, block_size, 8), generator=g) * 2 - 1) * D ** -0.5).to(dtype)
        self.vc = ((torch.rand((num_blocks, Hkv, D, block_size), generator=g) * 2 - 1) * D ** -0.5).to(dtype)
        self.bt = torch.randint(0, num_blocks, (batch, nb_seq), generator=g, dtype=torch.int32)
        self.lens = torch.full((batch,), ctx, dtype=torch.int32)
        self.slots = self.bt[:, (ctx - 1) // block_size].long() * block_size + (ctx - 1) % block_size
        self.pos = torch.full((batch,), ctx - 1, dtype=torch.long)
        inv = 1.0 / (s.rope_theta ** (torch.arange(0, D, 2, dtype=torch.float32) / D))
        fr = torch.outer(torch.arange(ctx + 1, dtype=torch.float32), inv)
        self.cos_sin = torch.cat([fr.cos(), fr.sin()], dim=-1).to(dtype)
        self.x = torch.randn((batch, s.hidden), generator=g).to(dtype)
        self.use_ref = refcpu.available() and block_size == 16 and dtype in (torch.bfloat16, torch.float32)
        self.lm_head = None

    def kind(self):
        return "reference" if self.use_ref else "port"

    def layer(self, h, residual):
        s, H, Hkv, D = self.s, self.H, self.Hkv, self.D
        x, residual = ops_ref.fused_add_rms_norm(h, residual, self.ln, s.eps)
        qkv = torch.matmul(x, self.w_qkv)
        q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
        q, k = ops_ref.rotary_embedding(self.pos, q.contiguous(), k.contiguous(), D, self.cos_sin, True)
        q, k, v = q.view(-1, H, D), k.view(-1, Hkv, D), v.reshape(-1, Hkv, D)
        if self.use_ref:
            refcpu.reshape_and_cache(k, v, self.kc, self.vc, self.slots)
            out = torch.empty_like(q)
            refcpu.paged_attention_v1(out, q, self.kc, self.vc, Hkv, float(D ** -0.5), self.bt, self.lens, self.BS,
                                      self.ctx)
        else:
            ops_ref.reshape_and_cache(k, v, self.kc, self.vc, self.slots)
            out = _attention_torch(q, self.kc, self.vc, self.bt, self.lens, float(D ** -0.5))
        h = torch.matmul(out.view(-1, H * D), self.w_o)
        x, residual = ops_ref.fused_add_rms_norm(h, residual, self.ln, s.eps)
        gu = torch.matmul(x, self.w_gu)
        act = ops_ref.silu_and_mul(gu)
        return torch.matmul(act, self.w_down), residual

    def head(self, h):
        if self.lm_head is None:
            g = torch.Generator().manual_seed(1)
            self.lm_head = (torch.randn((self.s.vocab, self.s.hidden), generator=g) * 0.02).to(self.dtype)
        return torch.argmax(torch.matmul(h, self.lm_head.t()), dim=-1)


def time_decode(shape, batch, ctx, dtype=torch.bfloat16, layer_reps=3, threads=None, tp=1):
    """Returns dict(step_s, layer_s, head_s, cores, kind, sample): one layer timed `layer_reps` times (best), lm_head
    once after a warm-up, step_s = layers * layer_s + head_s."""
    if threads:
        torch.set_num_threads(threads)
    cores = torch.get_num_threads()
    L = CpuDecodeLayer(shape, batch, ctx, dtype, tp=tp)
    h, res = L.x, torch.zeros_like(L.x)
    L.layer(h, res)  # warm-up
    ts = []
    for _ in range(layer_reps):
        t0 = time.perf_counter()
        L.layer(h, res)
        ts.append(time.perf_counter() - t0)
    layer_s = min(ts)
    L.head(h)
    t0 = time.perf_counter()
    L.head(h)
    head_s = time.perf_counter() - t0
    step_s = shape.layers * layer_s + head_s
    return dict(step_s=step_s, layer_s=layer_s, head_s=head_s, cores=cores, kind=L.kind(),
                sample="1 of %d decoder layers x%d reps (best) + lm_head once, batch %d ctx %d, scaled to %d layers; "
                       "linears = torch.matmul on dequantised %s weights, attention = %s" %
                       (shape.layers, layer_reps, batch, ctx, shape.layers, str(dtype).replace("torch.", ""),
                        "reference csrc/cpu paged_attention_v1 (oracle/_ref)" if L.use_ref else "torch restatement"))
