"""ORACLE (test infrastructure only): CPU restatements of the reference's *test oracles* and operator arithmetic for the
floating-point ops of the hot path.  Each function cites what it follows.  All run on CPU tensors."""
import math
from typing import Optional

import numpy as np
import torch


# ---- fp8 codec: csrc/quantization/fp8/nvidia/quant_utils.cuh:294-360 (dequant), :458-487 (quant, SATFINITE) ---------
_FP8 = {"fp8": torch.float8_e4m3fn, "fp8_e4m3": torch.float8_e4m3fn, "auto": torch.float8_e4m3fn,
        "fp8_e5m2": torch.float8_e5m2}
_FP8_MAX = {torch.float8_e4m3fn: 448.0, torch.float8_e5m2: 57344.0}


def fp8_quant(x: torch.Tensor, scale: float, kind: str = "fp8") -> torch.Tensor:
    """uint8 storage of cvt_satfinite(float(x) / scale)."""
    dt = _FP8[kind]
    v = x.float() / scale
    v = torch.clamp(v, -_FP8_MAX[dt], _FP8_MAX[dt])  # SATFINITE: saturate instead of inf/nan for finite inputs
    return v.to(dt).view(torch.uint8)


def fp8_dequant(u8: torch.Tensor, scale: float, out_dtype: torch.dtype, kind: str = "fp8") -> torch.Tensor:
    """T(float(half(fp8)) * scale)."""
    dt = _FP8[kind]
    return (u8.view(dt).to(torch.float16).float() * scale).to(out_dtype)


# ---- reshape_and_cache: python loop of tests/kernels/test_cache.py:179-189, kernel csrc/cache_kernels.cu:153-204 ----
def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype="auto", kv_scale=1.0):
    T, H, D = key.shape
    NB, _, Dx, BS, x = key_cache.shape
    for t in range(T):
        slot = int(slot_mapping[t])
        if slot < 0:
            continue
        blk, off = divmod(slot, BS)
        k = key[t]
        v = value[t]
        if kv_cache_dtype != "auto":
            k = fp8_quant(k, kv_scale, kv_cache_dtype)
            v = fp8_quant(v, kv_scale, kv_cache_dtype)
        key_cache[blk, :, :, off, :] = k.reshape(H, Dx, x)
        value_cache[blk, :, :, off] = v


# ---- paged attention: tests/kernels/test_attention.py:47-116 (ref_single_query_cached_kv_attention) -----------------
def paged_attention(query, key_cache, value_cache, block_tables, seq_lens, scale, alibi_slopes=None,
                    kv_cache_dtype="auto", kv_scale=1.0, p_dtype=None):
    """query [S,H,D]; caches in the paged layouts; returns [S,H,D] in query.dtype.  fp32 math; probabilities are rounded
    to the value dtype before P.V exactly where the reference test does (`.to(value.dtype)`, test_attention.py:58)."""
    S, H, D = query.shape
    NB, Hkv, Dx, BS, x = key_cache.shape
    G = H // Hkv
    out = torch.empty_like(query)
    dt = query.dtype
    p_dtype = p_dtype or dt
    for i in range(S):
        L = int(seq_lens[i])
        if L == 0:
            out[i] = 0
            continue
        nb = (L + BS - 1) // BS
        blocks = block_tables[i, :nb].long()
        k = key_cache[blocks]            # [nb, Hkv, Dx, BS, x]
        v = value_cache[blocks]          # [nb, Hkv, D, BS]
        if kv_cache_dtype != "auto":
            k = fp8_dequant(k, kv_scale, dt, kv_cache_dtype)
            v = fp8_dequant(v, kv_scale, dt, kv_cache_dtype)
        k = k.permute(0, 3, 1, 2, 4).reshape(nb * BS, Hkv, D)[:L]
        v = v.permute(0, 3, 1, 2).reshape(nb * BS, Hkv, D)[:L]
        if G > 1:
            k = torch.repeat_interleave(k, G, dim=1)
            v = torch.repeat_interleave(v, G, dim=1)
        w = scale * torch.einsum("hd,khd->hk", query[i].float(), k.float())
        if alibi_slopes is not None:
            pos = (torch.arange(L) - L + 1).float()
            w = w + alibi_slopes.float().view(-1, 1) * pos.view(1, -1)
        p = torch.softmax(w, dim=-1).to(p_dtype)
        out[i] = torch.einsum("hk,khd->hd", p.float(), v.float()).to(dt)
    return out


# ---- W8A8: tests/kernels/test_cutlass.py:36-48 (baseline_scaled_mm) -------------------------------------------------
def scaled_mm(a, b, scale_a, scale_b, out_dtype, bias=None):
    out = (scale_a * (scale_b * (a.to(torch.float32) @ b.to(torch.float32)))).to(out_dtype)
    if bias is not None:
        out = out + bias
    return out


# ---- activation quantizers ----------------------------------------------------------------------------------------
def scaled_fp8_quant(x: torch.Tensor, scale: Optional[torch.Tensor] = None):
    """csrc/quantization/fp8/common.cu:24-30,39-72; oracle of tests/quantization/test_fp8.py:58-94 (per_tensor_quantize)."""
    if scale is None:
        scale = (x.abs().max().float() / 448.0).reshape(1)
    inv = (1.0 / scale.float())
    q = torch.clamp(x.float() * inv, -448.0, 448.0).to(torch.float8_e4m3fn)
    return q, scale


def scaled_int8_quant(x: torch.Tensor, scale: Optional[torch.Tensor] = None):
    """int8_quant_kernels.cu:30-71; oracle of tests/kernels/test_int8_quant.py:33-46,64-71 (round-half-even, saturate)."""
    if scale is not None:
        q = torch.clamp(torch.round(x.float() / scale.float()), -128, 127).to(torch.int8)
        return q, scale
    amax = x.abs().amax(dim=-1, keepdim=True).float()
    s = amax / 127.0
    q = torch.clamp(torch.round(x.float() * (127.0 / amax)), -128, 127).to(torch.int8)
    return q, s


# ---- quantized linear oracles: tests/kernels/test_marlin_gemm.py:172-179, :208-227 -----------------------------------
def marlin_gemm_ref(a: torch.Tensor, w_ref: torch.Tensor) -> torch.Tensor:
    """fp32 dequant-then-matmul (north_star's <= 1e-2 oracle)."""
    return a.float() @ w_ref.float()


def rel_mean_err(out: torch.Tensor, ref: torch.Tensor) -> float:
    """compute_max_diff of marlin_utils.py:208-210."""
    return float(torch.mean(torch.abs(out.float() - ref.float())) / torch.mean(torch.abs(ref.float())))


# ---- glue: csrc/layernorm_kernels.cu:22-47,201-243; activation_kernels.cu:14-40; pos_encoding_kernels.cu:10-93 ------
def rms_norm(x, weight, eps):
    v = x.float()
    inv = torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + eps)
    return (v * inv).to(x.dtype) * weight


def fused_add_rms_norm(x, residual, weight, eps):
    z = x + residual
    return rms_norm(z, weight, eps), z


def silu_and_mul(x):
    d = x.shape[-1] // 2
    a = x[..., :d].float()
    return (a / (1.0 + torch.exp(-a))).to(x.dtype) * x[..., d:]


def rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox):
    dt = query.dtype
    rot = cos_sin_cache.shape[1]
    emb = rot // 2

    def rope(t):
        T = t.shape[0]
        th = t.reshape(T, -1, head_size).clone()
        cs = cos_sin_cache[positions]  # [T, rot]
        cos = cs[:, :emb].unsqueeze(1)
        sin = cs[:, emb:].unsqueeze(1)
        if is_neox:
            x = th[..., :emb]
            y = th[..., emb:rot]
        else:
            x = th[..., 0:rot:2]
            y = th[..., 1:rot:2]
        nx = (x * cos).to(dt) - (y * sin).to(dt)
        ny = (y * cos).to(dt) + (x * sin).to(dt)
        if is_neox:
            th[..., :emb] = nx
            th[..., emb:rot] = ny
        else:
            th[..., 0:rot:2] = nx
            th[..., 1:rot:2] = ny
        return th.reshape(t.shape)

    return rope(query), rope(key)


# ---- prefill attention over the paged cache ---------------------------------------------------------------------------
def context_attention_ref(q, k, v, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len, alibi_slopes=None,
                          sliding_window=0):
    """ORACLE of `context_attention_fwd` (vllm/attention/ops/prefix_prefill.py:674-799).  The reference kernel is Triton (it
    cannot run here), so this restates what ITS OWN test checks it against (tests/kernels/test_prefix_prefill.py:157-196):
    plain softmax attention with xformers' BlockDiagonalCausalFromBottomRightMask (+ make_local_attention_from_bottomright
    for the sliding window) over the concatenation [cached context (gathered through the block table), new tokens], GQA by
    repeating kv heads, scale 1/sqrt(head_size); ALiBi as in the kernel (:551-557): slope * (key_pos - query_pos).
    Parity for this row is pinned by that definition, not by stored vectors.  q [T, H, D]; k, v [T, Hkv, D]; caches in the
    vLLM layouts; returns [T, H, D] in q's dtype (fp32 softmax, probabilities rounded to the value dtype before P.V)."""
    T, H, D = q.shape
    Hkv = k.shape[1]
    G = H // Hkv
    BS = v_cache.shape[3]
    out = torch.zeros_like(q)
    scale = 1.0 / (D ** 0.5)
    for b in range(len(b_seq_len)):
        ctx, L = int(b_ctx_len[b]), int(b_seq_len[b])
        ql, st = L - ctx, int(b_start_loc[b])
        if ql <= 0:
            continue
        nb = (ctx + BS - 1) // BS
        blocks = b_loc[b, :nb].long()
        kc = k_cache[blocks].permute(0, 3, 1, 2, 4).reshape(nb * BS, Hkv, D)[:ctx]
        vc = v_cache[blocks].permute(0, 3, 1, 2).reshape(nb * BS, Hkv, D)[:ctx]
        kk = torch.cat([kc, k[st:st + ql]], 0).float().repeat_interleave(G, dim=1)      # [L, H, D]
        vv = torch.cat([vc, v[st:st + ql]], 0).repeat_interleave(G, dim=1)
        w = scale * torch.einsum("qhd,khd->hqk", q[st:st + ql].float(), kk)              # [H, ql, L]
        qpos = ctx + torch.arange(ql).view(-1, 1)
        kpos = torch.arange(L).view(1, -1)
        if alibi_slopes is not None:
            w = w + alibi_slopes.float().view(-1, 1, 1) * (kpos - qpos).float()
        mask = kpos <= qpos
        if sliding_window and sliding_window > 0:
            mask = mask & (qpos - kpos < sliding_window)
        w = w.masked_fill(~mask, float("-inf"))
        p = torch.softmax(w, dim=-1).to(q.dtype)
        out[st:st + ql] = torch.einsum("hqk,khd->qhd", p.float(), vv.float()).to(q.dtype)
    return out
