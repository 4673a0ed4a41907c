"""Synthetic Llama decode step built ONLY from the hot-path operators (bench / smoke / e2e harness).

What the reference does per decode step (LlamaDecoderLayer.forward, vllm/model_executor/models/llama.py:
input_layernorm -> qkv_proj -> rotary_emb -> Attention (reshape_and_cache + paged_attention) -> o_proj ->
post_attention_layernorm -> gate_up_proj -> SiluAndMul -> down_proj), with random-init weights in the quantized
layouts the LinearMethod classes consume (packed int32 / int16 tensors must be synthesised, any
bit pattern is a valid symmetric W4 weight).  Tensor parallelism follows vllm/model_executor/layers/linear.py:
column-parallel qkv / gate_up shard N, row-parallel o / down shard K and are followed by one all-reduce
(linear.py:779-803); attention heads are sharded, kv heads replicated when tp > Hkv (llama.py:107-117).

`DecodeEngine` is the public entry the end-to-end number is measured through: host (pinned) step inputs in, sampled
token ids out, with the whole device step captured in one CUDA graph (model_runner.py:910-1118 does the same).
"""
import os
from dataclasses import dataclass
from typing import Optional

import torch

from . import _custom_ops as ops

PARTITION_SIZE = 512  # vllm/attention/ops/paged_attn.py:10


@dataclass
class LlamaShape:
    name: str
    hidden: int
    inter: int
    heads: int
    kv_heads: int
    head_dim: int
    layers: int
    vocab: int
    rope_theta: float = 500000.0
    eps: float = 1e-5


LLAMA3_8B = LlamaShape("llama-3-8b", 4096, 14336, 32, 8, 128, 32, 128256)
LLAMA3_70B = LlamaShape("llama-3-70b", 8192, 28672, 64, 8, 128, 80, 128256)
TINY = LlamaShape("tiny-llama", 512, 1024, 8, 2, 64, 2, 1024)
MINI = LlamaShape("mini-llama-d128", 1024, 2048, 8, 2, 128, 2, 1024)   # head 128 + GQA: the fused attention fast path


def _rand_i32(shape, gen, device):
    return torch.randint(-2**31, 2**31 - 1, shape, generator=gen, dtype=torch.int32, device=device)


class QuantLinear:
    """One quantized linear in the layout its LinearMethod consumes.  kinds:
       w4a16     gptq_marlin   (qweight [K/16, 2N] int32, scales [K/g, N], workspace)     gptq_marlin.py:355-466
       w4a16_24  gptq_marlin_24 (qweight [K/32, 2N] int32, meta [K/32, 2N] int16, fp16)  gptq_marlin_24.py:140-291
       fp8/int8  cutlass_scaled_mm with per-tensor weight scale + dynamic activation quant fp8.py:240-379
       bf16      unquantized (lm_head default: gptq_marlin.py:154-157)"""

    fuse_silu_max_m = int(os.environ.get("B200_FUSE_SILU_MAX_M", "64"))

    def __init__(self, kind, K, N, dtype, gen, device, group=128):
        self.kind, self.K, self.N, self.dtype = kind, K, N, dtype
        dev = device
        if kind == "w4a16":
            # a GPTQ checkpoint as the reference loads it (qweight [K/8, N] int32 row-packed, scales [K/g, N], sequential
            # g_idx): GPTQMarlinLinearMethod then repacks it into the Marlin layout on its first apply() -- our
            # gptq_marlin_repack kernel + marlin_permute_scales run on the product path, nothing is fabricated in kernel layout
            from .quantization import GPTQMarlinConfig, GPTQMarlinLinearMethod
            self.method = GPTQMarlinLinearMethod(GPTQMarlinConfig(4, group, False, True, False))
            self.layer = torch.nn.Module()
            with torch.device(dev):
                self.method.create_weights(self.layer, K, [N], K, N, dtype)
            self.layer.qweight.data.copy_(_rand_i32((K // 8, N), gen, dev))
            ng = K // group if group > 0 else 1
            self.layer.scales.data.copy_((torch.rand((ng, N), generator=gen, device=dev) * 0.002 + 0.003).to(dtype))
            self.layer.g_idx.data.copy_((torch.arange(K, device=dev) // (group if group > 0 else K)).to(torch.int32))
        elif kind == "w4a16_24":
            self.qweight = _rand_i32((K // 32, N * 2), gen, dev)
            # each 4-bit quad of the metadata must be one of the six legal 2:4 codes (format_24.py:96-102)
            legal = torch.tensor([0x4, 0x8, 0x9, 0xC, 0xD, 0xE], dtype=torch.int32, device=dev)
            quads = legal[torch.randint(0, 6, (K // 32, N * 2, 4), generator=gen, device=dev)]
            m = quads[..., 0] | (quads[..., 1] << 4) | (quads[..., 2] << 8) | (quads[..., 3] << 12)
            self.meta = m.to(torch.int16)
            ng = K // group if group > 0 else 1
            self.scales = (torch.rand((ng, N), generator=gen, device=dev) * 0.002 + 0.003).to(torch.float16)
            self.workspace = torch.zeros((N // 128) * 64, dtype=torch.int32, device=dev)
        elif kind in ("fp8", "fp8_static", "int8"):
            # fp8: dynamic per-tensor activation scale (Fp8Config activation_scheme "dynamic", two-kernel quantiser);
            # fp8_static: activation_scheme "static" (input_scale from the checkpoint) -> the quantiser fuses into the
            # producing norm / activation kernel; int8: compressed-tensors W8A8 dynamic per-token, fuses as well
            self.input_scale = torch.full((1,), 0.05, dtype=torch.float32, device=dev) if kind == "fp8_static" else None
            if kind in ("fp8", "fp8_static"):
                w = (torch.randn((N, K), generator=gen, device=dev) * 0.02 / 0.0005).clamp(-448, 448)
                self.weight = w.to(torch.float8_e4m3fn).t()      # column-major [K, N] view (fp8.py:262-270)
            else:
                self.weight = torch.randint(-127, 128, (N, K), generator=gen, device=dev, dtype=torch.int8).t()
            self.weight_scale = torch.full((1,), 0.0005, dtype=torch.float32, device=dev)
        elif kind == "bf16":
            self.weight = (torch.randn((N, K), generator=gen, device=dev) * 0.02).to(dtype)
        else:
            raise ValueError(kind)

    def __getattr__(self, name):
        # Marlin-layout tensors of the w4a16 kind live on the LinearMethod's layer (valid after the first call)
        if name in ("qweight", "scales") and "layer" in self.__dict__:
            return getattr(self.__dict__["layer"], name)
        raise AttributeError(name)

    def weight_bytes(self):
        K, N = self.K, self.N
        if self.kind == "w4a16":
            return K * N // 2 + self.layer.scales.numel() * 2
        if self.kind == "w4a16_24":
            return K * N // 4 + K * N // 8 + self.scales.numel() * 2
        if self.kind in ("fp8", "fp8_static", "int8"):
            return K * N
        return K * N * 2

    def __call__(self, x):
        M = x.shape[0]
        if self.kind == "w4a16":
            return self.method.apply(self.layer, x)
        if self.kind == "w4a16_24":
            return ops.gptq_marlin_24_gemm(x, self.qweight, self.meta, self.scales, self.workspace, 4, M, self.N,
                                           self.K)
        if self.kind in ("fp8", "fp8_static"):
            xq, xs = ops.scaled_fp8_quant(x, self.input_scale)
            return ops.cutlass_scaled_mm(xq, self.weight, xs, self.weight_scale, self.dtype)
        if self.kind == "int8":
            xq, xs = ops.scaled_int8_quant(x)
            return ops.cutlass_scaled_mm(xq, self.weight, xs, self.weight_scale, self.dtype)
        return torch.matmul(x, self.weight.t())

    def silu_and_mul(self, x):
        """gate_up layers: SiluAndMul(self(x)); one launch for the W4A16 kind (GEMM epilogue), two otherwise."""
        # measured on B200 (profiles/r2_silu_fusion_ab.txt): +2.3 .. 4.6 % tok/s at every batch 1..64 once the activation math
        # is split over the gate AND up threads of the epilogue (gate threads alone stalled the accumulator hand-off: -1 .. 2 %
        # at batch 32 / 64); prefill-sized M keeps the two launches (not measured)
        if self.kind == "w4a16" and x.shape[0] <= self.fuse_silu_max_m and self.method.can_fuse_silu_and_mul(self.layer):
            return self.method.apply_silu_and_mul(self.layer, x)
        gu = self(x)
        act = torch.empty((gu.shape[0], gu.shape[1] // 2), dtype=gu.dtype, device=gu.device)
        ops.silu_and_mul(act, gu)
        return act

    # ---- W8A8 with the activation quantiser fused into the producing kernel (fp8_static / int8) ----------------------
    @property
    def fused_quant_kind(self):
        return {"fp8_static": "fp8", "int8": "int8_token", "fp8": "fp8_dynamic"}.get(self.kind)   # fp8_dynamic: opt-in, see SyntheticLlama.fuse_dynamic_fp8

    def mm_quantized(self, xq, xs):
        return ops.cutlass_scaled_mm(xq, self.weight, xs, self.weight_scale, self.dtype)


class SyntheticLlama:
    """Random-init Llama decoder (one TP rank) over a paged KV cache; `forward` is one decode step for `batch` tokens."""

    def __init__(self, shape: LlamaShape, quant="w4a16", dtype=torch.bfloat16, device="cuda:0", tp_size=1, tp_rank=0,
                 kv_cache_dtype="auto", kv_scale=1.0, block_size=16, num_blocks=4096, max_pos=8192, seed=0,
                 group=None, layers=None):
        self.shape, self.quant, self.dtype, self.device = shape, quant, dtype, device
        self.tp, self.rank = tp_size, tp_rank
        self.kv_cache_dtype, self.kv_scale, self.block_size, self.num_blocks = kv_cache_dtype, kv_scale, block_size, num_blocks
        self.group = group  # torch.distributed process group (None -> default) when tp > 1
        self.custom_ar = None  # CustomAllreduce once enable_custom_all_reduce() has been called
        self.n_custom_ar_calls = 0  # stand-alone all-reduce kernel launches (bench.py counts them per step)
        self.fuse_gemm_all_reduce = True  # row-parallel W4A16 GEMMs all-reduce in their epilogue when custom_ar is on
        # W8A8: activation quantiser (static fp8, per-token int8, dynamic per-tensor fp8) fused into the norm / silu kernels
        self.fuse_act_quant = os.environ.get("B200_FUSE_ACT_QUANT", "1") != "0"
        # dynamic per-tensor fp8 (kind "fp8_dynamic": grid-wide maximum inside the norm / silu launch) is OFF by default:
        # measured on the 8B fp8 step it is slower than the three short kernels it replaces (5.02 vs 4.91 ms per step, 322 vs
        # 514 launches) -- the in-kernel grid barrier costs more than two PDL-overlapped launches (profiles/r2_fp8_dynamic_fusion_ab.txt)
        self.fuse_dynamic_fp8 = os.environ.get("B200_FUSE_DYNAMIC_FP8", "0") == "1"
        self.fuse_attention = True  # rope + cache write + paged attention in one launch when the shape is on the fast path
        self.fuse_silu = os.environ.get("B200_NO_FUSE_SILU", "0") != "1"   # SiluAndMul in the epilogue of the W4A16 gate_up GEMM
        s = shape
        self.nlayers = layers if layers is not None else s.layers
        assert s.heads % tp_size == 0
        self.H = s.heads // tp_size
        self.Hkv = max(1, s.kv_heads // tp_size)     # replicated when tp > Hkv (llama.py:107-117)
        D = s.head_dim
        self.q_size, self.kv_size = self.H * D, self.Hkv * D
        gen = torch.Generator(device=device).manual_seed(seed + 7919 * tp_rank)
        inter = s.inter // tp_size
        self.layers = []
        for _ in range(self.nlayers):
            L = dict(
                ln1=torch.ones(s.hidden, dtype=dtype, device=device),
                ln2=torch.ones(s.hidden, dtype=dtype, device=device),
                qkv=QuantLinear(quant, s.hidden, self.q_size + 2 * self.kv_size, dtype, gen, device),
                o=QuantLinear(quant, self.q_size, s.hidden, dtype, gen, device),
                gate_up=QuantLinear(quant, s.hidden, 2 * inter, dtype, gen, device),
                down=QuantLinear(quant, inter, s.hidden, dtype, gen, device),
            )
            self.layers.append(L)
        self.norm = torch.ones(s.hidden, dtype=dtype, device=device)
        g0 = torch.Generator(device=device).manual_seed(seed)  # embeddings / lm_head identical on every rank
        self.embed = (torch.randn((s.vocab, s.hidden), generator=g0, device=device) * 0.02).to(dtype)
        assert s.vocab % tp_size == 0
        vshard = s.vocab // tp_size
        full_head = QuantLinear("bf16", s.hidden, s.vocab, dtype, g0, device)
        full_head.weight = full_head.weight[tp_rank * vshard:(tp_rank + 1) * vshard].contiguous()
        full_head.N = vshard
        self.lm_head = full_head
        # rotary cos/sin cache: vllm/model_executor/layers/rotary_embedding.py (_compute_cos_sin_cache)
        inv = 1.0 / (s.rope_theta ** (torch.arange(0, D, 2, dtype=torch.float32, device=device) / D))
        fr = torch.outer(torch.arange(max_pos, dtype=torch.float32, device=device), inv)
        self.cos_sin = torch.cat([fr.cos(), fr.sin()], dim=-1).to(dtype)
        # paged KV cache, one flat tensor per layer (PagedAttention.get_kv_cache_shape: paged_attn.py:37-43)
        cache_t = dtype if kv_cache_dtype == "auto" else torch.uint8
        x = 16 // torch.tensor([], dtype=cache_t).element_size()
        self.kv = []
        for _ in range(self.nlayers):
            kc = torch.empty((num_blocks, self.Hkv, D // x, block_size, x), dtype=cache_t, device=device)
            vc = torch.empty((num_blocks, self.Hkv, D, block_size), dtype=cache_t, device=device)
            if cache_t == torch.uint8:
                kc.random_(0, 120, generator=gen); vc.random_(0, 120, generator=gen)
            else:
                kc.uniform_(-D ** -0.5, D ** -0.5, generator=gen); vc.uniform_(-D ** -0.5, D ** -0.5, generator=gen)
            self.kv.append((kc, vc))
        self.scale = float(D ** -0.5)

    # ---- byte accounting ------------------------------------------------------------------------------
    def linear_weight_bytes(self):
        return sum(L[k].weight_bytes() for L in self.layers for k in ("qkv", "o", "gate_up", "down"))

    def kv_read_bytes(self, seq_lens):
        es = self.kv[0][0].element_size()
        return int(sum(seq_lens)) * self.Hkv * self.shape.head_dim * 2 * es * self.nlayers

    def enable_custom_all_reduce(self, max_size=8192 * 1024):
        """Route the row-parallel all-reduces through the one-shot NVLink kernel (csrc/all_reduce.cu); tensors it does
        not take (too large / unaligned) still go to NCCL, like tensor_model_parallel_all_reduce in the reference
        (parallel_state.py:273-293)."""
        from .custom_all_reduce import CustomAllreduce
        if self.tp > 1 and self.custom_ar is None:
            self.custom_ar = CustomAllreduce(self.group, torch.device(self.device), max_size)
        return self.custom_ar

    def _row_parallel(self, lin, x):
        """RowParallelLinear.forward (linear.py:779-803): local GEMM over this rank's K shard, then the sum over ranks.
        With the custom all-reduce enabled the W4A16 GEMM exchanges its output tiles with the peers inside its own
        epilogue (one kernel, csrc/marlin_gemm_kernel.cuh `FUSE`); otherwise GEMM followed by an all-reduce."""
        if (self.tp > 1 and self.custom_ar is not None and self.fuse_gemm_all_reduce and lin.kind == "w4a16"
                and self.custom_ar.fuse_next_gemm(x.shape[0], lin.N)):
            return lin(x)
        return self._all_reduce(lin(x))

    def _all_reduce(self, t):
        if self.tp > 1:
            if self.custom_ar is not None and self.custom_ar.should_custom_ar(t):
                self.n_custom_ar_calls += 1
                return self.custom_ar.all_reduce_(t)
            torch.distributed.all_reduce(t, group=self.group)
        return t

    def _attention(self, li, q, k, v, positions, meta):
        """rotary_embedding -> reshape_and_cache -> paged_attention_v1/v2 (llama.py LlamaAttention.forward +
        PagedAttention.write_to_paged_cache / forward_decode), or the one-launch fusion of the three."""
        kc, vc = self.kv[li]
        B, D = q.shape[0], self.shape.head_dim
        if "query_start_loc" in meta:
            # prefill (prompt or chunk): rotary -> cache write -> attention over (cached context, the chunk itself);
            # the prefill branch of the reference's backends (xformers.py:262-300 -> PagedAttention.forward_prefix)
            from .attention import PagedAttention
            ops.rotary_embedding(positions, q.view(B, self.H * D), k.view(B, self.Hkv * D), D, self.cos_sin, True)
            ops.reshape_and_cache(k, v, kc, vc, meta["slot_mapping"], self.kv_cache_dtype, self.kv_scale)
            out = PagedAttention.forward_prefix(q, k, v, kc, vc, meta["block_tables"], meta["query_start_loc"],
                                                meta["seq_lens"], meta["context_lens"], meta["max_query_len"], None, None,
                                                self.kv_cache_dtype, self.kv_scale, self.scale)
            return out.view(B, self.H * D)
        out = torch.empty((B, self.H, D), dtype=q.dtype, device=q.device)
        max_len = meta["max_seq_len"]
        nparts = (max_len + PARTITION_SIZE - 1) // PARTITION_SIZE
        # v1/v2 heuristic of PagedAttention.forward_decode (paged_attn.py:120-121)
        use_v1 = max_len <= 8192 and (nparts == 1 or B * self.H > 512)
        tmp = es = ml = None
        if not use_v1:
            tmp = torch.empty((B, self.H, nparts, D), dtype=q.dtype, device=q.device)
            es = torch.empty((B, self.H, nparts), dtype=torch.float32, device=q.device)
            ml = torch.empty_like(es)
        if self.fuse_attention and D == 128 and self.block_size == 16 and ops.rope_cache_paged_attention(
                out, es, ml, tmp, q, k, v, kc, vc, positions, self.cos_sin, meta["slot_mapping"], self.Hkv, self.scale,
                meta["block_tables"], meta["seq_lens"], self.block_size, max_len, self.kv_cache_dtype, self.kv_scale):
            return out.view(B, self.H * D)
        # rotary_embedding takes [tokens, heads * head_size] (pos_encoding_kernels.cu:130-141): the 2-D views of the same memory
        ops.rotary_embedding(positions, q.view(B, self.H * D), k.view(B, self.Hkv * D), D, self.cos_sin, True)
        ops.reshape_and_cache(k, v, kc, vc, meta["slot_mapping"], self.kv_cache_dtype, self.kv_scale)
        if use_v1:
            ops.paged_attention_v1(out, q, kc, vc, self.Hkv, self.scale, meta["block_tables"], meta["seq_lens"],
                                   self.block_size, max_len, None, self.kv_cache_dtype, self.kv_scale)
        else:
            ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, self.Hkv, self.scale, meta["block_tables"],
                                   meta["seq_lens"], self.block_size, max_len, None, self.kv_cache_dtype,
                                   self.kv_scale)
        return out.view(B, self.H * D)

    def forward(self, token_ids, positions, meta, return_logits=False):
        """One step over `token_ids` [T].  Decode: T = batch, `meta` = slot_mapping / block_tables / seq_lens / max_seq_len.
        Prefill: `meta` additionally holds query_start_loc [S + 1], context_lens [S], max_query_len (seq_lens = context +
        query length); logits are then computed for the LAST token of every sequence only (logits_processor.py
        _prune_hidden_states)."""
        s = self.shape
        D = s.head_dim
        h = self.embed.index_select(0, token_ids)
        residual = None
        fq = self.layers[0]["qkv"].fused_quant_kind if self.fuse_act_quant else None
        if fq == "fp8_dynamic" and (not self.fuse_dynamic_fp8 or token_ids.shape[0] > ops.FUSED_DYNAMIC_FP8_MAX_TOKENS):
            fq = None   # off by default (measured slower); prefill-sized batches cannot take it at all (co-residency)
        for li, L in enumerate(self.layers):
            if fq is not None:
                # norm (+ residual add) and the activation quantiser of qkv_proj in one kernel
                first = residual is None
                if first:
                    residual = h
                xq, xs = ops.rms_norm_quant(h, L["ln1"], s.eps, fq, L["qkv"].input_scale, None if first else residual)
                qkv = L["qkv"].mm_quantized(xq, xs)
            else:
                if residual is None:
                    residual = h
                    x = torch.empty_like(h)
                    ops.rms_norm(x, h, L["ln1"], s.eps)
                else:
                    ops.fused_add_rms_norm(h, residual, L["ln1"], s.eps)
                    x = h
                qkv = L["qkv"](x)
            q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
            a = self._attention(li, q.view(-1, self.H, D), k.view(-1, self.Hkv, D), v.view(-1, self.Hkv, D), positions, meta)
            h = self._row_parallel(L["o"], a)
            if fq is not None:
                xq, xs = ops.rms_norm_quant(h, L["ln2"], s.eps, fq, L["gate_up"].input_scale, residual)
                gu = L["gate_up"].mm_quantized(xq, xs)
                aq, asc = ops.silu_and_mul_quant(gu, fq, L["down"].input_scale)
                h = self._all_reduce(L["down"].mm_quantized(aq, asc))
                continue
            ops.fused_add_rms_norm(h, residual, L["ln2"], s.eps)
            if self.fuse_silu:
                act = L["gate_up"].silu_and_mul(h)
            else:
                gu = L["gate_up"](h)
                act = torch.empty((gu.shape[0], gu.shape[1] // 2), dtype=gu.dtype, device=gu.device)
                ops.silu_and_mul(act, gu)
            h = self._row_parallel(L["down"], act)
        ops.fused_add_rms_norm(h, residual, self.norm, s.eps)
        if "query_start_loc" in meta:
            h = h.index_select(0, (meta["query_start_loc"][1:] - 1).long())
        logits = self.lm_head(h)
        if self.tp > 1:  # vocab-parallel logits gathered once per step
            parts = [torch.empty_like(logits) for _ in range(self.tp)]
            torch.distributed.all_gather(parts, logits, group=self.group)
            logits = torch.cat(parts, dim=-1)
        if return_logits:
            return logits
        return ops.greedy_sample(logits)

    def prefill_meta(self, query_lens, context_lens, block_tables):
        """Device metadata of a prefill call: sequence b contributes query_lens[b] new tokens at positions context_lens[b]..,
        stored through its row of `block_tables` (int32 [S, max_blocks], CPU or device).  Returns (positions, meta)."""
        dev = self.device
        bt = block_tables.to(dev)
        pos, slots, starts = [], [], [0]
        btc = block_tables.cpu()
        for b, (n, c) in enumerate(zip(query_lens, context_lens)):
            p = torch.arange(c, c + n)
            pos.append(p)
            slots.append(btc[b, (p // self.block_size).long()].long() * self.block_size + p % self.block_size)
            starts.append(starts[-1] + n)
        meta = dict(slot_mapping=torch.cat(slots).to(dev), block_tables=bt,
                    query_start_loc=torch.tensor(starts, dtype=torch.int32, device=dev),
                    seq_lens=torch.tensor([n + c for n, c in zip(query_lens, context_lens)], dtype=torch.int32, device=dev),
                    context_lens=torch.tensor(list(context_lens), dtype=torch.int32, device=dev),
                    max_query_len=max(query_lens))
        return torch.cat(pos).to(dev), meta


class DecodeEngine:
    """Public end-to-end entry: `step(host_inputs) -> host token ids`.

    Mirrors the decode branch of the reference's ModelRunner (vllm/worker/model_runner.py:630-645 input buffers,
    :910-1118 CUDA-graph capture/replay): step inputs are assembled on the host, copied into persistent device buffers,
    one CUDA graph replays the whole step, and the sampled token ids come back to the host."""

    def __init__(self, model: SyntheticLlama, batch: int, max_ctx: int, use_graph=True):
        self.m, self.B, self.max_ctx = model, batch, max_ctx
        dev = model.device
        self.max_blocks = (max_ctx + model.block_size - 1) // model.block_size
        B = batch
        self.d_tok = torch.zeros(B, dtype=torch.long, device=dev)
        self.d_pos = torch.zeros(B, dtype=torch.long, device=dev)
        self.d_slot = torch.full((B,), -1, dtype=torch.long, device=dev)   # _PAD_SLOT_ID until inputs are loaded (model_runner.py:60)
        self.d_lens = torch.ones(B, dtype=torch.int32, device=dev)
        self.d_bt = torch.zeros((B, self.max_blocks), dtype=torch.int32, device=dev)
        self.d_out = torch.zeros(B, dtype=torch.long, device=dev)
        self.h_out = torch.zeros(B, dtype=torch.long).pin_memory()
        self.use_graph = use_graph
        self.graph = None
        self.graph_max_len = None
        self.launches_per_step = None

    def host_inputs(self, ctx_len: int, seed=0):
        """Pinned host tensors of one synthetic decode step: every sequence has `ctx_len` tokens in the cache (incl. the
        one being written), random block tables WITHOUT collisions: like the reference's block manager, no two sequences
        of a batch own the same physical block (benchmark_paged_attention.py:62-70 draws with collisions, which makes a
        step's result depend on kernel order as soon as it writes the cache)."""
        m, B = self.m, self.B
        g = torch.Generator().manual_seed(seed)
        nb = (ctx_len + m.block_size - 1) // m.block_size
        bt = torch.zeros((B, self.max_blocks), dtype=torch.int32)
        assert B * nb <= m.num_blocks, "KV cache too small for batch x context"
        bt[:, :nb] = torch.randperm(m.num_blocks, generator=g)[:B * nb].to(torch.int32).view(B, nb)
        pos = torch.full((B,), ctx_len - 1, dtype=torch.long)
        slot = bt[:, (ctx_len - 1) // m.block_size].long() * m.block_size + (ctx_len - 1) % m.block_size
        tok = torch.randint(0, m.shape.vocab, (B,), generator=g, dtype=torch.long)
        lens = torch.full((B,), ctx_len, dtype=torch.int32)
        return dict(token_ids=tok.pin_memory(), positions=pos.pin_memory(), slot_mapping=slot.pin_memory(),
                    seq_lens=lens.pin_memory(), block_tables=bt.pin_memory(), max_seq_len=ctx_len)

    @staticmethod
    def h2d_bytes(hi):
        return sum(v.numel() * v.element_size() for v in hi.values() if isinstance(v, torch.Tensor))

    def _device_step(self, max_len):
        meta = dict(slot_mapping=self.d_slot, block_tables=self.d_bt, seq_lens=self.d_lens, max_seq_len=max_len)
        self.d_out.copy_(self.m.forward(self.d_tok, self.d_pos, meta))

    def load_inputs(self, hi):
        self.d_tok.copy_(hi["token_ids"], non_blocking=True)
        self.d_pos.copy_(hi["positions"], non_blocking=True)
        self.d_slot.copy_(hi["slot_mapping"], non_blocking=True)
        self.d_lens.copy_(hi["seq_lens"], non_blocking=True)
        self.d_bt.copy_(hi["block_tables"], non_blocking=True)

    def prepare(self, max_len):
        """Eager warm-up (reserves kernel scratch) then capture the step for this max_seq_len bucket."""
        if self.graph is not None and self.graph_max_len == max_len:
            return
        self._device_step(max_len)
        torch.cuda.synchronize()
        if not self.use_graph:
            return
        st = torch.cuda.Stream(device=self.m.device)
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            self._device_step(max_len)
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._device_step(max_len)
        self.graph_max_len = max_len
        torch.cuda.synchronize()

    def run_device(self):
        """Device-only step (inputs already resident)."""
        if self.graph is not None:
            self.graph.replay()
        else:
            self._device_step(self.graph_max_len)

    def step(self, hi):
        """Host in -> host out (synchronous)."""
        # inputs first: the eager warm-up / graph capture inside prepare() then runs on real metadata (its cache writes
        # are the ones the step makes anyway) instead of whatever the device buffers held
        self.load_inputs(hi)
        if self.use_graph:
            self.prepare(hi["max_seq_len"])
        else:
            self.graph_max_len = hi["max_seq_len"]
        self.run_device()
        self.h_out.copy_(self.d_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.h_out
