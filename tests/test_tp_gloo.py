"""Tensor-parallel host logic on CPU, world_size 2 over gloo (the N > 1 path of the decode step: SURVEY 8e).

Each rank shards a full-precision layer the way the reference's Column/Row/QKV parallel linears do (nm_vllm_b200/tp.py),
quantises ITS shard through the oracle (Marlin W4 g128), computes the oracle's dequant-matmul on it and all-reduces the
row-parallel partial sums; the result must equal the unsharded quantised layer."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nm_vllm_b200 import tp
from oracle import ops_ref, quant_layouts as ql


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        torch.manual_seed(0)  # same full-size tensors on every rank
        hidden, inter, H, Hkv, D, M, g = 256, 512, 4, 2, 64, 5, 128
        x = torch.randn(M, hidden)
        w_qkv = torch.randn(hidden, (H + 2 * Hkv) * D) * 0.05
        w_gu = torch.randn(hidden, 2 * inter) * 0.05
        w_down = torch.randn(inter, hidden) * 0.05

        def q_ref(w):  # oracle quantise -> dequantised fp32 reference weight + the packed tensors
            w_ref, packed, scales, *_ = ql.marlin_quantize(w, 4, g)
            return w_ref.float(), packed, scales

        # ---- column-parallel (merged gate_up) -> act -> row-parallel (down) + all-reduce ---------------------------
        full_gu, _, _ = q_ref(w_gu)
        full_down, _, s_full = q_ref(w_down)
        ref_act = ops_ref.silu_and_mul(x @ full_gu)
        ref_out = ref_act @ full_down

        my_gu, _, _ = q_ref(tp.shard_merged_column(w_gu, [inter, inter], world, rank))
        my_down, packed, s_my = q_ref(tp.shard_row(w_down, world, rank, group_size=g))
        # the shard's packed tensor / scales have the shard's own shapes and the scale rows are the full tensor's rows
        assert packed.shape == (inter // world // 16, hidden * 16 // 8)
        lo, hi = tp.shard_range(inter, world, rank)
        _, sp, sps = ql.marlin_perm_tables(4)
        un = lambda s_, k: ql.unpermute_scales(s_, k, hidden, g, sp, sps)
        torch.testing.assert_close(un(s_my, inter // world), un(s_full, inter)[lo // g:hi // g])
        act = ops_ref.silu_and_mul(x @ my_gu)
        torch.testing.assert_close(act, tp.shard_row_input(ref_act, world, rank), rtol=1e-5, atol=1e-5)
        out = tp.all_reduce_(act @ my_down)
        torch.testing.assert_close(out, ref_out, rtol=1e-4, atol=1e-4)

        # ---- qkv head sharding: local q/k/v columns are the full layer's columns of this rank's heads -----------------
        full_qkv, _, _ = q_ref(w_qkv)
        my_qkv, _, _ = q_ref(tp.shard_qkv(w_qkv, H, Hkv, D, world, rank))
        y_full = x @ full_qkv
        qf, kf, vf = y_full.split([H * D, Hkv * D, Hkv * D], dim=1)
        qlo, qhi = tp.shard_range(H, world, rank)
        klo, khi = tp.kv_head_range(Hkv, world, rank)
        expect = torch.cat([qf[:, qlo * D:qhi * D], kf[:, klo * D:khi * D], vf[:, klo * D:khi * D]], dim=1)
        torch.testing.assert_close(x @ my_qkv, expect, rtol=1e-5, atol=1e-5)

        # kv-head replication when tp > num_kv_heads: ranks 0,1 of tp=2 share the single kv head
        assert tp.kv_head_range(1, world, rank) == (0, 1)
        # vocab-parallel logits: all_gather of the shards == full logits -> same argmax on every rank
        lm = torch.randn(hidden, 64)
        parts = [torch.empty(M, 64 // world) for _ in range(world)]
        dist.all_gather(parts, x @ tp.shard_column(lm, world, rank))
        assert torch.equal(torch.cat(parts, 1).argmax(-1), (x @ lm).argmax(-1))
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))


@pytest.mark.timeout(180)
def test_tp2_row_column_qkv_over_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(30)
    assert all(r[1] == "ok" for r in res), res


def test_shard_helpers_single_process():
    w = torch.arange(8 * 12, dtype=torch.float32).reshape(8, 12)
    assert torch.equal(torch.cat([tp.shard_column(w, 2, r) for r in range(2)], 1), w)
    assert torch.equal(torch.cat([tp.shard_row(w, 4, r) for r in range(4)], 0), w)
    with pytest.raises(AssertionError):
        tp.shard_row(torch.zeros(384, 4), 2, 0, group_size=128)   # 192 rows do not hold whole groups of 128
    assert tp.kv_head_range(8, 2, 1) == (4, 8)
    assert tp.kv_head_range(8, 16, 5) == (2, 3)
