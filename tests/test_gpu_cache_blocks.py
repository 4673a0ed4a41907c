"""GPU parity of the KV block maintenance ops (copy_blocks, swap_blocks, reshape_and_cache_flash) against CPU
restatements of the reference semantics (tests/kernels/test_cache.py:36-112 copy_blocks, :205-278 flash, :281-358
swap_blocks).  Pure byte moves: every comparison is bit-exact."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops(built_extension):
    from nm_vllm_b200 import _custom_ops
    return _custom_ops


def _caches(num_layers, NB, H, D, BS, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    ks, vs = [], []
    for _ in range(num_layers):
        if dtype == torch.uint8:
            ks.append(torch.randint(0, 255, (NB, H, D // x, BS, x), generator=g, dtype=torch.uint8))
            vs.append(torch.randint(0, 255, (NB, H, D, BS), generator=g, dtype=torch.uint8))
        else:
            ks.append(torch.randn((NB, H, D // x, BS, x), generator=g).to(dtype))
            vs.append(torch.randn((NB, H, D, BS), generator=g).to(dtype))
    return ks, vs


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32, torch.uint8])
@pytest.mark.parametrize("geom", [(1, 8, 64, 8, 64, 4), (5, 8, 128, 16, 200, 37), (3, 1, 80, 32, 80, 25), (2, 8, 256, 16, 30, 1)])
def test_copy_blocks(ops, dtype, geom):
    L, H, D, BS, NB, nmap = geom
    random.seed(0)
    src = random.sample(range(NB), nmap)
    rest = list(set(range(NB)) - set(src))
    dst = random.sample(rest, 2 * nmap)
    mapping = []
    for i in range(nmap):   # every source block is copied to two destinations (test_cache.py:70-80)
        mapping += [(src[i], dst[2 * i]), (src[i], dst[2 * i + 1])]
    ks, vs = _caches(L, NB, H, D, BS, dtype, seed=3)
    dks, dvs = [k.to(DEV) for k in ks], [v.to(DEV) for v in vs]
    ops.copy_blocks(dks, dvs, torch.tensor(mapping, dtype=torch.int64, device=DEV).view(-1, 2))
    for s, d in mapping:
        for k in ks:
            k[d].copy_(k[s])
        for v in vs:
            v[d].copy_(v[s])
    for a, b in zip(dks + dvs, ks + vs):
        assert torch.equal(a.cpu(), b)


def test_copy_blocks_empty(ops):
    ks, vs = _caches(2, 8, 2, 64, 16, torch.float16, seed=1)
    dks, dvs = [k.to(DEV) for k in ks], [v.to(DEV) for v in vs]
    ops.copy_blocks(dks, dvs, torch.empty((0, 2), dtype=torch.int64, device=DEV))
    ops.copy_blocks([], [], torch.empty((0, 2), dtype=torch.int64, device=DEV))
    assert torch.equal(dks[0].cpu(), ks[0])


@pytest.mark.parametrize("direction", [("cuda", "cpu"), ("cpu", "cuda"), ("cuda", "cuda")])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.uint8])
@pytest.mark.parametrize("contiguous_runs", [False, True])
def test_swap_blocks(ops, direction, dtype, contiguous_runs):
    H, D, BS, NB, nmap = 8, 128, 16, 96, 40
    random.seed(1)
    if contiguous_runs:   # runs of consecutive blocks are merged into one copy each: same bytes
        mapping = [(10 + i, 50 + i) for i in range(12)] + [(30 + i, 5 + i) for i in range(7)] + [(2, 90)]
    else:
        src = random.sample(range(NB), nmap)
        dst = random.sample(range(NB), nmap) if direction[0] != direction[1] else \
            random.sample(sorted(set(range(NB)) - set(src)), nmap)
        mapping = list(zip(src, dst))
    (ks,), (vs,) = _caches(1, NB, H, D, BS, dtype, seed=7)
    (kd,), (vd,) = _caches(1, NB, H, D, BS, dtype, seed=8)
    for s_cpu, d_cpu in ((ks, kd), (vs, vd)):
        s = s_cpu.to(DEV) if direction[0] == "cuda" else s_cpu.clone().pin_memory()
        d = d_cpu.to(DEV) if direction[1] == "cuda" else d_cpu.clone().pin_memory()
        ops.swap_blocks(s, d, torch.tensor(mapping, dtype=torch.int64, device="cpu").view(-1, 2))
        torch.cuda.synchronize()
        ref = d_cpu.clone()
        for a, b in mapping:
            ref[b].copy_(s_cpu[a])
        assert torch.equal(d.cpu(), ref)
        assert torch.equal(s.cpu(), s_cpu)


def test_swap_blocks_rejects_gpu_mapping(ops):
    (ks,), _ = _caches(1, 8, 2, 64, 16, torch.float16, seed=1)
    with pytest.raises(RuntimeError, match="block_mapping must be on CPU"):
        ops.swap_blocks(ks.to(DEV), ks.to(DEV), torch.zeros((1, 2), dtype=torch.int64, device=DEV))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("geom", [(7, 2, 128, 16, 16), (42, 8, 64, 32, 40), (3, 1, 80, 8, 9), (64, 8, 128, 16, 256), (5, 3, 20, 16, 4)])
def test_reshape_and_cache_flash(ops, dtype, geom):
    T, H, D, BS, NB = geom
    g = torch.Generator().manual_seed(5)
    slots = torch.tensor(random.Random(2).sample(range(NB * BS), T), dtype=torch.int64)
    slots[T // 2] = -1   # padded token: skipped (cache_kernels.cu:218-221)
    qkv = torch.randn((T, 3, H, D), generator=g).to(dtype)
    kc = torch.randn((NB, BS, H, D), generator=g).to(dtype)
    vc = torch.randn((NB, BS, H, D), generator=g).to(dtype)
    dqkv, dkc, dvc = qkv.to(DEV), kc.to(DEV), vc.to(DEV)
    ops.reshape_and_cache_flash(dqkv[:, 1], dqkv[:, 2], dkc, dvc, slots.to(DEV), "auto")   # strided token rows
    for t in range(T):
        s = int(slots[t])
        if s < 0:
            continue
        kc[s // BS, s % BS] = qkv[t, 1]
        vc[s // BS, s % BS] = qkv[t, 2]
    assert torch.equal(dkc.cpu(), kc)
    assert torch.equal(dvc.cpu(), vc)


def test_reshape_and_cache_flash_rejects_fp8(ops):
    k = torch.zeros((2, 2, 64), dtype=torch.float16, device=DEV)
    c = torch.zeros((4, 16, 2, 64), dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.reshape_and_cache_flash(k, k, c, c, torch.zeros(2, dtype=torch.int64, device=DEV), "fp8")
