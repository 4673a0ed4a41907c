"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement, in numpy / torch-CPU, of the reference's weight-quantisation and Marlin layout helpers.  Pinned
against the reference's own Python helpers run in the authoring container (oracle/make_golden.py writes the vectors
under tests/golden/, tests/test_oracle_golden.py replays them).

Reference functions restated (all under /root/reference/vllm/model_executor/layers/quantization/utils/):
  quant_utils.py:39-106   quantize_weights      -> quantize_weights
  quant_utils.py:109-122  sort_weights          -> sort_weights
  quant_utils.py:125-146  gptq_pack             -> gptq_pack
  marlin_perms.py:9-43    get_perms             -> marlin_perm_tables
  marlin_utils.py:25-57   marlin_weights        -> marlin_pack
  marlin_utils.py:60-69   marlin_permute_scales -> permute_scales
  marlin_24_perms.py:9-50 get_perms_24          -> marlin24_perm_tables
  format_24.py:21-46      meta reordering       -> meta_scatter_offsets
  format_24.py:52-177     2:4 compression       -> compress_24
  format_24.py:184-276    2:4 decompression     -> decompress_24
  format_24.py:279-308    mask_creator          -> mask_24
  marlin_utils.py:144-205 compress_quantized_24_weight / marlin_24_quantize -> same names
"""
import numpy as np
import torch

SUPPORTED_BITS = (4, 8)


def pack_factor(num_bits: int) -> int:
    assert num_bits in SUPPORTED_BITS
    return 32 // num_bits


# ------------------------------------------------------------------------------------------------ quantisation
def quantize_weights(w: torch.Tensor, num_bits: int, group_size: int, act_order: bool = False, seed: int = 0,
                     rand_perm: torch.Tensor = None):
    """Symmetric per-group quantisation.  Returns (w_ref, q_w int32 [K,N], s [K/g,N], g_idx, rand_perm).
    act_order simulates GPTQ desc_act with a random row permutation (pass `rand_perm` to replay a given one)."""
    K, N = w.shape
    if group_size == -1:
        group_size = K
    assert group_size <= K and K % group_size == 0
    max_q = 2**num_bits - 1
    half = (max_q + 1) // 2
    wg = w.reshape(K // group_size, group_size, N)
    s = wg.abs().amax(dim=1, keepdim=True)  # [G,1,N]
    s = s * (2.0 / max_q)
    q = torch.round(wg / s).to(torch.int32) + half
    q = torch.clamp(q, 0, max_q)
    w_ref = (q - half).to(torch.float16) * s  # dtype promotion as in the reference (half * w.dtype)
    q = q.reshape(K, N).contiguous()
    w_ref = w_ref.reshape(K, N).contiguous()
    s = s.reshape(K // group_size, N).contiguous()
    g_idx = torch.empty(0, dtype=torch.int32)
    if act_order:
        assert group_size < K
        g_idx = (torch.arange(K, dtype=torch.int32) // group_size)
        if rand_perm is None:
            rand_perm = torch.randperm(K, generator=torch.Generator().manual_seed(seed))
        rand_perm = rand_perm.long()
        g_idx = g_idx[rand_perm].contiguous()
        q = q[rand_perm, :].contiguous()
        w_ref = w_ref[rand_perm, :].contiguous()
    else:
        rand_perm = torch.empty(0, dtype=torch.int64)
    return w_ref, q, s, g_idx, rand_perm


def sort_weights(q_w: torch.Tensor, g_idx: torch.Tensor):
    order = torch.argsort(g_idx).to(torch.int32)
    return q_w[order.long(), :].contiguous(), g_idx[order.long()].contiguous(), order


def gptq_pack(q_w: torch.Tensor, num_bits: int) -> torch.Tensor:
    """[K,N] ints -> [K/pf, N] int32, row r holds k = pf*r + i in bits [num_bits*i, ...)."""
    K, N = q_w.shape
    pf = pack_factor(num_bits)
    assert K % pf == 0
    v = q_w.numpy().astype(np.uint32).reshape(K // pf, pf, N)
    out = np.zeros((K // pf, N), dtype=np.uint32)
    for i in range(pf):
        out |= v[:, i, :] << np.uint32(num_bits * i)
    return torch.from_numpy(out.view(np.int32).copy())


# ------------------------------------------------------------------------------------------------ Marlin (dense)
def marlin_perm_tables(num_bits: int):
    """(perm[1024], scale_perm[64], scale_perm_single[32]).

    A 16(k) x 64(n) tile is stored so that fragment lane l = 4c + q (c = column inside an 8-wide half tile, q = k-pair)
    of n16-tile j holds k in {2q, 2q+1, 2q+8, 2q+9} for n = 16j + c (half 0) and n = 16j + c + 8 (half 1).
    Element index inside the source chunk (4 n16-tiles of 16x16, row-major each) is 256 j + 16 k + n%16.
    """
    pf = pack_factor(num_bits)
    order = [0, 2, 4, 6, 1, 3, 5, 7] if num_bits == 4 else [0, 2, 1, 3]
    perm = np.empty(1024, dtype=np.int64)
    pos = 0
    for lane in range(32):
        c, q = divmod(lane, 4)
        for j in range(4):
            entries = []
            for half in range(2):
                for krow in (2 * q, 2 * q + 1, 2 * q + 8, 2 * q + 9):
                    entries.append(256 * j + 16 * krow + c + 8 * half)
            for g in range(0, 8, pf):
                grp = entries[g:g + pf]
                for o in order:
                    perm[pos] = grp[o]
                    pos += 1
    scale_perm = [i + 8 * j for i in range(8) for j in range(8)]
    scale_perm_single = [2 * i + j for i in range(4) for j in (0, 1, 8, 9, 16, 17, 24, 25)]
    return perm, scale_perm, scale_perm_single


def _tile_permute(q_w: torch.Tensor, perm: np.ndarray) -> np.ndarray:
    K, N = q_w.shape
    assert K % 16 == 0 and N % 16 == 0
    v = q_w.numpy().reshape(K // 16, 16, N // 16, 16).transpose(0, 2, 1, 3).reshape(K // 16, N * 16)
    v = v.reshape(-1, perm.size)[:, perm].reshape(K // 16, N * 16)
    return v


def _pack_rows(v: np.ndarray, num_bits: int) -> torch.Tensor:
    pf = pack_factor(num_bits)
    v = v.astype(np.uint32)
    out = np.zeros((v.shape[0], v.shape[1] // pf), dtype=np.uint32)
    for i in range(pf):
        out |= v[:, i::pf] << np.uint32(num_bits * i)
    return torch.from_numpy(out.view(np.int32).copy())


def marlin_pack(q_w: torch.Tensor, num_bits: int, perm: np.ndarray = None) -> torch.Tensor:
    """[K,N] ints -> Marlin tile layout [K/16, N*16/pf] int32."""
    if perm is None:
        perm = marlin_perm_tables(num_bits)[0]
    return _pack_rows(_tile_permute(q_w, perm), num_bits)


def marlin_unpack(packed: torch.Tensor, size_k: int, size_n: int, num_bits: int, perm: np.ndarray = None) -> torch.Tensor:
    """Inverse of marlin_pack (used by the GEMM oracle to prove a kernel's reading of the layout)."""
    if perm is None:
        perm = marlin_perm_tables(num_bits)[0]
    pf = pack_factor(num_bits)
    p = packed.numpy().view(np.uint32)
    vals = np.zeros((p.shape[0], p.shape[1] * pf), dtype=np.int32)
    mask = np.uint32((1 << num_bits) - 1)
    for i in range(pf):
        vals[:, i::pf] = ((p >> np.uint32(num_bits * i)) & mask).astype(np.int32)
    inv = np.empty_like(perm)
    inv[perm] = np.arange(perm.size)
    v = vals.reshape(-1, perm.size)[:, inv].reshape(size_k // 16, size_n // 16, 16, 16)
    return torch.from_numpy(v.transpose(0, 2, 1, 3).reshape(size_k, size_n).copy())


def permute_scales(s: torch.Tensor, size_k: int, size_n: int, group_size: int, scale_perm, scale_perm_single):
    if group_size < size_k and group_size != -1:
        s = s.reshape(-1, len(scale_perm))[:, scale_perm]
    else:
        s = s.reshape(-1, len(scale_perm_single))[:, scale_perm_single]
    return s.reshape(-1, size_n).contiguous()


def unpermute_scales(s: torch.Tensor, size_k: int, size_n: int, group_size: int, scale_perm, scale_perm_single):
    p = scale_perm if (group_size < size_k and group_size != -1) else scale_perm_single
    inv = np.empty(len(p), dtype=np.int64)
    inv[np.asarray(p)] = np.arange(len(p))
    return s.reshape(-1, len(p))[:, inv].reshape(-1, size_n).contiguous()


def marlin_quantize(w: torch.Tensor, num_bits: int, group_size: int, act_order: bool = False, seed: int = 0,
                    rand_perm: torch.Tensor = None):
    """-> (w_ref, marlin_q_w, marlin_s, g_idx, sort_indices, rand_perm)  (marlin_utils.py:72-107)"""
    K, N = w.shape
    if group_size == -1:
        group_size = K
    w_ref, q_w, s, g_idx, rand_perm = quantize_weights(w, num_bits, group_size, act_order, seed, rand_perm)
    sort_indices = torch.empty(0, dtype=torch.int32)
    if act_order:
        q_w, g_idx, sort_indices = sort_weights(q_w, g_idx)
    perm, sp, sps = marlin_perm_tables(num_bits)
    return (w_ref, marlin_pack(q_w, num_bits, perm), permute_scales(s, K, N, group_size, sp, sps), g_idx, sort_indices,
            rand_perm)


# ------------------------------------------------------------------------------------------------ Marlin-24 (2:4)
def marlin24_perm_tables(num_bits: int):
    """(perm[1024], scale_perm[64], scale_perm_single[64]) for the [32(k, dense) = 16 compressed rows, 64] tiles."""
    pf = pack_factor(num_bits)
    order = [0, 2, 4, 6, 1, 3, 5, 7] if num_bits == 4 else [0, 2, 1, 3]
    perm = np.empty(1024, dtype=np.int64)
    pos = 0
    for lane in range(32):
        col, q = divmod(lane, 4)
        col_o = col // 2
        entries = []
        for block in range(2):
            for krow in (2 * q, 2 * q + 1, 2 * q + 8, 2 * q + 9):
                entries.append(16 * krow + col_o * 256 + 8 * (col % 2) + 4 * block)
        for j in range(4):
            shifted = [e + j for e in entries]
            for g in range(0, 8, pf):
                grp = shifted[g:g + pf]
                for o in order:
                    perm[pos] = grp[o]
                    pos += 1
    scale_perm = [8 * i + j for i in range(8) for j in (0, 4, 1, 5, 2, 6, 3, 7)]
    scale_perm_single = [8 * i + j for i in range(8) for j in range(8)]
    return perm, scale_perm, scale_perm_single


def mask_24(w_t: torch.Tensor) -> torch.Tensor:
    """Keep the 2 largest |w| of every 4 consecutive elements (row-major order of the given tensor)."""
    g = w_t.detach().abs().reshape(-1, 4)
    drop = torch.argsort(g, dim=1)[:, :2]
    m = torch.ones_like(g)
    m.scatter_(1, drop, 0.0)
    return m.reshape(w_t.shape)


def inject_24(w: torch.Tensor):
    mask = mask_24(w.t().contiguous()).t().bool()
    return (mask * w).contiguous(), mask.contiguous()


def meta_scatter_offsets(m: int, meta_ncols: int, meta_itemsize: int = 2) -> torch.Tensor:
    """Destination offset of meta[row, col] in the reordered (CUTLASS sm80 ColumnMajorInterleaved<2>) buffer."""
    r = torch.arange(m)[:, None].repeat(1, meta_ncols)
    c = torch.arange(meta_ncols).repeat(m, 1)
    gx = 64
    gy = 32 if meta_itemsize == 2 else 16
    r = (r // gx * gx + (r % 2) * 2 + (r % 8) // 4 + ((r % gy) % 4) // 2 * 32 + ((r % gx) // 8) * 4)
    tr = ((r % 2 == 0) & (c % 2 == 1)).to(torch.int8)
    bl = ((r % 2 == 1) & (c % 2 == 0)).to(torch.int8)
    r = r + (tr - bl)
    c = c - (tr - bl)
    return ((c // 2) * m * 2 + r * 2 + (c % 2)).reshape(-1)


def compress_24(dense: torch.Tensor):
    """dense [m, k] (int32 or 16-bit float; zeros mark pruned slots) -> (sparse [m, k/2], meta_reordered int16 [m, k/16])."""
    m, k = dense.shape
    assert m % 32 == 0 and k % 16 == 0
    d4 = dense.reshape(-1, k // 4, 4)
    nz = d4 != 0
    m0, m1, m3 = nz[..., 0], nz[..., 1], nz[..., 3]
    e0 = m0 & m1
    e1 = ~m0 & m1
    e2 = ~m0 & ~m1
    idx0 = e1.to(torch.int64) | (e2.to(torch.int64) << 1)
    idx1 = (e0 | e2 | m3).to(torch.int64) | ((e1 | ~m1).to(torch.int64) << 1)
    s0 = d4.gather(-1, idx0.unsqueeze(-1))
    s1 = d4.gather(-1, idx1.unsqueeze(-1))
    sparse = torch.stack((s0, s1), dim=-1).reshape(m, k // 2)
    quad = (idx0 | (idx1 << 2)).reshape(m, k // 16, 4).to(torch.int16)
    meta = quad[..., 0] | (quad[..., 1] << 4) | (quad[..., 2] << 8) | (quad[..., 3] << 12)
    out = meta.new_empty(m * (k // 16))
    out.scatter_(0, meta_scatter_offsets(m, k // 16, 2), meta.reshape(-1))
    return sparse, out.reshape(m, k // 16)


def decompress_24(sparse: torch.Tensor, meta_reordered: torch.Tensor) -> torch.Tensor:
    m, kh = sparse.shape
    ncols = meta_reordered.shape[1]
    assert ncols * 16 == 2 * kh
    meta = torch.gather(meta_reordered.reshape(-1), 0, meta_scatter_offsets(m, ncols, 2)).reshape(m, ncols)
    mi = meta.to(torch.int64) & 0xFFFF
    idx = torch.stack([(mi >> (2 * i)) & 3 for i in range(8)], dim=-1)  # [m, ncols, 8]
    base = (torch.arange(m * ncols * 4) * 4).reshape(-1, 1).repeat(1, 2).reshape(-1)
    dense = torch.zeros(m * 2 * kh, dtype=sparse.dtype)
    dense.scatter_(0, idx.reshape(-1) + base, sparse.reshape(-1))
    return dense.reshape(m, 2 * kh)


def compress_quantized_24_weight(q_24: torch.Tensor, size_k: int, size_n: int, num_bits: int):
    zp = (1 << num_bits) // 2
    comp, meta = compress_24((q_24 - zp).t().contiguous())
    comp = comp.t().contiguous() + zp
    # "resize without moving data": [N, K/16] int16 buffer reinterpreted as [K/32, 2N]
    meta = meta.reshape(-1)[:].reshape(meta.shape[1] // 2, meta.shape[0] * 2)
    return comp, meta


def marlin_24_quantize(w: torch.Tensor, num_bits: int, group_size: int):
    """-> (w_24_ref, marlin_24_q_w_comp [K/32, N*16/pf], meta int16 [K/32, 2N], marlin_24_s)  (marlin_utils.py:167-205)"""
    K, N = w.shape
    if group_size == -1:
        group_size = K
    w_24, _ = inject_24(w)
    w_24_ref, q_w_24, s, _, _ = quantize_weights(w_24, num_bits, group_size, False)
    comp, meta = compress_quantized_24_weight(q_w_24, K, N, num_bits)
    perm, sp, sps = marlin24_perm_tables(num_bits)
    q_packed = marlin_pack(comp, num_bits, perm)
    s_perm = permute_scales(s, K, N, group_size, sp, sps)
    return w_24_ref, q_packed, meta, s_perm
