"""Product-side 2:4 (Sparse-Marlin) layout helpers: what a checkpoint exporter / the synthetic harness needs to produce the
tensors `gptq_marlin_24_gemm` consumes.  Mirrors the reference's Python utilities
(vllm/model_executor/layers/quantization/utils/format_24.py:21-177 `sparse_semi_structured_from_dense_cutlass`,
`mask_creator` :279-308; utils/marlin_utils.py:144-205 `compress_quantized_24_weight`, `marlin_24_quantize`;
utils/marlin_24_perms.py:9-50) in plain torch ops that run on the GPU; every function is bit-exact against the golden
vectors generated from the reference's own code (tests/test_quantization_cpu.py).

Format recap.  Along K every group of 4 weights keeps 2.  Per group the two kept positions (p0 < p1) are stored as
`p0 | p1 << 2`, four groups per int16; the kept values go, in order, into a [N, K/2] matrix.  Groups with fewer than two
non-zeros pad with a zero-valued slot: the slot choice below is the reference's (it makes the code unique per mask).
The int16 metadata matrix [N, K/16] is then scattered into the operand order of the sm80 sparse tensor-core instruction
(rows interleaved in blocks of 32 / 64, 2x2 transposes inside 8-row groups) and REINTERPRETED as [K/32, 2N]."""
from typing import Tuple

import torch


def mask_2_of_4(w: torch.Tensor) -> torch.Tensor:
    """Boolean mask of `w` [..., 4k] keeping the two largest magnitudes of every 4 consecutive elements of the last dim
    (ties resolved like torch.argsort: the earlier element counts as smaller)."""
    g = w.detach().abs().reshape(-1, 4)
    drop = torch.argsort(g, dim=1)[:, :2]
    keep = torch.ones_like(g, dtype=torch.bool)
    keep.scatter_(1, drop, False)
    return keep.reshape(w.shape)


def _position_lut(device) -> torch.Tensor:
    """[16, 2] (p0, p1) for every non-zero mask b0 | b1 << 1 | b2 << 2 | b3 << 3 of a group.  Two non-zeros: their
    positions.  Otherwise the reference's padding rule (format_24.py:96-138): p0 = 0 if b0 else (1 if b1 else 2);
    p1 = 1 if b0 and b1, else 3 if b3 (or nothing usable below it), else 2 -- written out from the boolean form
    p1 = (e0 | e2 | b3) | ((e1 | ~b1) << 1) with e0 = b0 & b1, e1 = ~b0 & b1, e2 = ~b0 & ~b1."""
    rows = []
    for mask in range(16):
        b0, b1, b3 = bool(mask & 1), bool(mask & 2), bool(mask & 8)
        e0, e1, e2 = b0 and b1, (not b0) and b1, (not b0) and (not b1)
        p0 = (1 if e1 else 0) | ((1 if e2 else 0) << 1)
        p1 = (1 if (e0 or e2 or b3) else 0) | ((1 if (e1 or not b1) else 0) << 1)
        rows.append((p0, p1))
    return torch.tensor(rows, dtype=torch.int64, device=device)


def _meta_destinations(rows: int, cols: int, device) -> torch.Tensor:
    """Flat destination of meta[r, c] ([rows, cols] int16) in the sparse-MMA operand order (format_24.py:21-46)."""
    r = torch.arange(rows, device=device).unsqueeze(1).expand(rows, cols)
    c = torch.arange(cols, device=device).unsqueeze(0).expand(rows, cols)
    # row shuffle inside blocks of 64 rows: (r % 2) * 2 + (r % 8) / 4 picks the slot inside a quad, rows 2-3 of every 4 move up by
    # 32, and each 8-row group advances by 4
    rr = (r // 64) * 64 + (r % 2) * 2 + (r % 8) // 4 + ((r % 32) % 4) // 2 * 32 + ((r % 64) // 8) * 4
    # 2 x 2 transpose of (row parity, column parity)
    up = ((rr % 2 == 0) & (c % 2 == 1)).to(torch.int64)
    down = ((rr % 2 == 1) & (c % 2 == 0)).to(torch.int64)
    rr = rr + up - down
    cc = c - up + down
    return ((cc // 2) * rows * 2 + rr * 2 + (cc % 2)).reshape(-1)


def compress_2_of_4(dense: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """dense [N, K] with zeros at the pruned positions -> (values [N, K/2], metadata int16 [N, K/16] in operand order)."""
    n, k = dense.shape
    assert n % 32 == 0 and k % 16 == 0, "2:4 compression needs N % 32 == 0 and K % 16 == 0"
    g = dense.reshape(n, k // 4, 4)
    nz = (g != 0).to(torch.int64)
    code = nz[..., 0] | (nz[..., 1] << 1) | (nz[..., 2] << 2) | (nz[..., 3] << 3)
    pos = _position_lut(dense.device)[code]                      # [n, k/4, 2]
    vals = torch.gather(g, 2, pos).reshape(n, k // 2)
    quad = (pos[..., 0] | (pos[..., 1] << 2)).reshape(n, k // 16, 4)
    word = quad[..., 0] | (quad[..., 1] << 4) | (quad[..., 2] << 8) | (quad[..., 3] << 12)
    word = torch.where(word >= 2 ** 15, word - 2 ** 16, word).to(torch.int16)
    out = torch.empty(n * (k // 16), dtype=torch.int16, device=dense.device)
    out.scatter_(0, _meta_destinations(n, k // 16, dense.device), word.reshape(-1))
    return vals, out.reshape(n, k // 16)


def compress_quantized_24_weight(q_24: torch.Tensor, size_k: int, size_n: int, num_bits: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """q_24 [K, N] unsigned quantised weights whose pruned slots hold the zero point 2^(bits-1) -> (q_comp [K/2, N],
    meta int16 [K/32, 2N]).  The zero point is removed first so that pruned == 0, and put back on the kept values."""
    assert q_24.shape == (size_k, size_n)
    zp = (1 << num_bits) // 2
    vals, meta = compress_2_of_4((q_24.to(torch.int64) - zp).t().contiguous())
    q_comp = (vals.t().contiguous() + zp).to(torch.int32)
    return q_comp, meta.reshape(-1).reshape(size_k // 32, size_n * 2)      # reinterpretation, not a transpose


def get_perms_24(num_bits: int):
    """(weight permutation [1024], scale permutation [64], channelwise scale permutation [64]) of the Marlin-24 tiles: a
    tile is 16 compressed rows x 64 columns; lane l = 4 col + q of a warp owns compressed rows {2q, 2q+1, 2q+8, 2q+9} of
    the column pairs it feeds to mma.sp, four int32 per lane, nibbles interleaved [0,2,4,6,1,3,5,7] (bytes [0,2,1,3])."""
    pf = 32 // num_bits
    order = [0, 2, 4, 6, 1, 3, 5, 7] if num_bits == 4 else [0, 2, 1, 3]
    perm = []
    for lane in range(32):
        col, q = divmod(lane, 4)
        base = []
        for block in (0, 1):
            for row in (2 * q, 2 * q + 1, 2 * q + 8, 2 * q + 9):
                base.append(16 * row + (col // 2) * 256 + 8 * (col % 2) + 4 * block)
        for j in range(4):
            shifted = [b + j for b in base]
            for s0 in range(0, 8, pf):
                perm.extend(shifted[s0 + o] for o in order)
    scale_perm = [8 * i + j for i in range(8) for j in (0, 4, 1, 5, 2, 6, 3, 7)]
    scale_perm_single = [8 * i + j for i in range(8) for j in range(8)]
    return torch.tensor(perm, dtype=torch.long), scale_perm, scale_perm_single


def marlin_24_pack(q_comp: torch.Tensor, num_bits: int) -> torch.Tensor:
    """q_comp [K/2, N] -> Sparse-Marlin weights int32 [K/32, N * 16 / pf]: 16 x 16 blocks laid out row-major per tile row, the
    1024-entry tile permutation, then pf values per int32 (value i in bits [bits * i, ...))."""
    kc, n = q_comp.shape
    pf, tile = 32 // num_bits, 16
    assert kc % tile == 0 and n % tile == 0
    perm = get_perms_24(num_bits)[0].to(q_comp.device)
    t = q_comp.reshape(kc // tile, tile, n // tile, tile).permute(0, 2, 1, 3).reshape(kc // tile, n * tile)
    t = t.reshape(-1, perm.numel())[:, perm].reshape(kc // tile, n * tile).to(torch.int64)
    t = t.reshape(kc // tile, n * tile // pf, pf)
    out = torch.zeros(t.shape[:2], dtype=torch.int64, device=q_comp.device)
    for i in range(pf):
        out |= t[..., i] << (num_bits * i)
    out &= 0xFFFFFFFF
    return torch.where(out >= 2 ** 31, out - 2 ** 32, out).to(torch.int32)


def marlin_24_permute_scales(s: torch.Tensor, size_k: int, size_n: int, group_size: int, num_bits: int = 4) -> torch.Tensor:
    _, grouped, single = get_perms_24(num_bits)
    perm = grouped if (group_size < size_k and group_size != -1) else single
    idx = torch.tensor(perm, dtype=torch.long, device=s.device)
    return s.reshape(-1, len(perm)).index_select(1, idx).reshape(-1, size_n).contiguous()


def marlin_24_quantize(w: torch.Tensor, num_bits: int, group_size: int):
    """fp16 weights [K, N] -> (w_24_ref [K, N], q_packed, meta, scales) ready for gptq_marlin_24_gemm: prune 2 of 4 along K,
    symmetric per-group quantisation, compression, Marlin-24 packing (marlin_utils.py:167-205)."""
    k, n = w.shape
    group = k if group_size == -1 else group_size
    w24 = torch.where(mask_2_of_4(w.t().contiguous()).t(), w, torch.zeros_like(w))
    max_q = 2 ** num_bits - 1
    half = (max_q + 1) // 2
    wg = w24.reshape(k // group, group, n)
    s = wg.abs().amax(dim=1, keepdim=True) * (2.0 / max_q)
    q = torch.clamp(torch.round(wg / s).to(torch.int32) + half, 0, max_q)
    w_ref = ((q - half).to(torch.float16) * s).reshape(k, n)
    q_comp, meta = compress_quantized_24_weight(q.reshape(k, n), k, n, num_bits)
    return (w_ref, marlin_24_pack(q_comp, num_bits), meta,
            marlin_24_permute_scales(s.reshape(k // group, n), k, n, group_size, num_bits))
