"""ORACLE / test infrastructure (never on the product path): ONE decode step of nm_vllm_b200.decoder.SyntheticLlama
restated on the CPU from the oracle's own pieces, on the SAME weights (packed Marlin tensors unpacked with
quant_layouts.marlin_unpack, scales un-permuted) and a copy of the KV cache.  Used by tests/test_gpu_decode_e2e.py and
__graft_entry__.smoke() as the checker of the chained GPU kernels.  Reference call chain: LlamaDecoderLayer.forward
(vllm/model_executor/models/llama.py:200-230) -> gptq_marlin apply / rotary_emb / PagedAttention.forward_decode."""
import torch

from . import ops_ref, quant_layouts as ql


def _dense(lin):
    """bf16 [K, N] weight the Marlin linear represents: (q - 8) * s per group of 128 (quant_utils.py:39-106)."""
    K, N = lin.K, lin.N
    q = ql.marlin_unpack(lin.qweight.cpu(), K, N, 4).float() - 8.0
    _, scale_perm, scale_perm_single = ql.marlin_perm_tables(4)
    s = ql.unpermute_scales(lin.scales.cpu(), K, N, 128, scale_perm, scale_perm_single).float()
    return (q.view(K // 128, 128, N) * s.view(K // 128, 1, N)).view(K, N).to(lin.dtype)


def _mm(x, w):
    return torch.matmul(x.float(), w.float()).to(x.dtype)


def decode_step_logits(m, hi, kv0):
    """m: SyntheticLlama (W4A16, tp 1); hi: DecodeEngine.host_inputs(...); kv0: CPU copies of m.kv taken BEFORE the GPU
    step (updated in place with this step's K/V).  Returns fp32 logits [B, vocab]."""
    s, D = m.shape, m.shape.head_dim
    tok, pos = hi["token_ids"], hi["positions"]
    h = m.embed.cpu()[tok]
    residual = None
    cos_sin = m.cos_sin.cpu()
    for li, L in enumerate(m.layers):
        if residual is None:
            residual = h
            x = ops_ref.rms_norm(h, L["ln1"].cpu(), s.eps)
        else:
            x, residual = ops_ref.fused_add_rms_norm(h, residual, L["ln1"].cpu(), s.eps)
        qkv = _mm(x, _dense(L["qkv"]))
        q, k, v = qkv.split([m.q_size, m.kv_size, m.kv_size], dim=-1)
        q, k = ops_ref.rotary_embedding(pos, q, k, D, cos_sin, True)
        kc, vc = kv0[li]
        ops_ref.reshape_and_cache(k.reshape(-1, m.Hkv, D), v.reshape(-1, m.Hkv, D), kc, vc, hi["slot_mapping"])
        a = ops_ref.paged_attention(q.reshape(-1, m.H, D), kc, vc, hi["block_tables"], hi["seq_lens"], m.scale)
        o = _mm(a.reshape(-1, m.H * D), _dense(L["o"]))
        x, residual = ops_ref.fused_add_rms_norm(o, residual, L["ln2"].cpu(), s.eps)
        act = ops_ref.silu_and_mul(_mm(x, _dense(L["gate_up"])))
        h = _mm(act, _dense(L["down"]))
    x, _ = ops_ref.fused_add_rms_norm(h, residual, m.norm.cpu(), s.eps)
    ref_logits = torch.matmul(x.float(), m.lm_head.weight.cpu().float().t())

    return ref_logits
