"""Host side of the one-shot NVLink all-reduce (csrc/all_reduce.cu), mirroring the reference's
`vllm/distributed/device_communicators/custom_all_reduce.py:35-270` (`CustomAllreduce`): same constructor arguments,
`should_custom_ar`, `custom_all_reduce` (returns None when the tensor does not qualify so the caller falls back to
NCCL, :241-262), `all_reduce`, `close`.

Differences that follow from the B200 design: the peer-visible memory is ONE region per rank
allocated by the library (not torch tensors + per-input registration), inputs are pushed into the peers' regions by
the kernel itself, so there is no `register_buffer` / `register_graph_buffers` step and `capture()` is a no-op context
manager kept for call-site compatibility.
"""
import contextlib
import ctypes
from typing import Optional

import torch
import torch.distributed as dist

from . import capi

_DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


class CustomAllreduce:
    _SUPPORTED_WORLD_SIZES = [2, 4, 6, 8]   # custom_all_reduce.py:37

    def __init__(self, group=None, device=None, max_size: int = 8192 * 1024) -> None:
        self.disabled = True
        self._ctx = None
        self._own = None
        self._peers = []
        self.group = group
        self.rank = dist.get_rank(group)
        self.world_size = dist.get_world_size(group)
        if self.world_size == 1:
            return
        if self.world_size not in self._SUPPORTED_WORLD_SIZES:
            raise ValueError("CustomAllreduce: world size %d not in %s" % (self.world_size, self._SUPPORTED_WORLD_SIZES))
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        self.max_size = int(max_size)
        lib = capi.lib()
        with torch.cuda.device(self.device):
            nbytes = ctypes.c_int64()
            self._check(lib.b200_ar_region_bytes(self.world_size, ctypes.c_int64(self.max_size), ctypes.byref(nbytes)))
            ptr = ctypes.c_void_p()
            handle = ctypes.create_string_buffer(64)
            self._check(lib.b200_ar_alloc(ctypes.c_int64(nbytes.value), ctypes.byref(ptr), handle))
            self._own = ptr.value
            handles = [None] * self.world_size
            dist.all_gather_object(handles, bytes(handle.raw), group=group)
            regions = (ctypes.c_void_p * self.world_size)()
            for r, h in enumerate(handles):
                if r == self.rank:
                    regions[r] = self._own
                else:
                    pp = ctypes.c_void_p()
                    self._check(lib.b200_ar_open(ctypes.c_char_p(h), ctypes.byref(pp)))
                    self._peers.append(pp.value)
                    regions[r] = pp.value
            ctx = ctypes.c_int64()
            self._check(lib.b200_ar_init(self.rank, self.world_size, regions, ctypes.c_int64(self.max_size),
                                         ctypes.byref(ctx)))
            self._ctx = ctx.value
        dist.barrier(group=group)   # every rank has mapped every region before the first kernel pushes into them
        self.disabled = False

    @staticmethod
    def _check(rc):
        if rc != 0:
            raise RuntimeError("custom all-reduce: " + capi.last_error())

    @contextlib.contextmanager
    def capture(self):
        yield

    def should_custom_ar(self, inp: torch.Tensor) -> bool:
        """csrc/custom_all_reduce.cu:32-50: byte size multiple of 16, within max_size, contiguous."""
        if self.disabled or inp.dtype not in _DT or not inp.is_cuda:
            return False
        nbytes = inp.numel() * inp.element_size()
        return nbytes % 16 == 0 and 0 < nbytes <= self.max_size and inp.is_contiguous() and inp.data_ptr() % 16 == 0

    def all_reduce(self, inp: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if out is None:
            out = torch.empty_like(inp)
        # the kernel works on raw pointers: refuse what should_custom_ar would refuse instead of reducing the wrong memory
        if not self.should_custom_ar(inp) or out.shape != inp.shape or out.dtype != inp.dtype or not out.is_contiguous() \
                or out.device != inp.device:
            raise ValueError("CustomAllreduce.all_reduce: tensor does not qualify (dtype / contiguity / size / device); "
                             "use custom_all_reduce() which returns None in that case")
        if inp.device != self.device:
            raise ValueError("CustomAllreduce.all_reduce: tensor is on %s, the communicator on %s" % (inp.device, self.device))
        st = torch.cuda.current_stream(inp.device).cuda_stream
        self._check(capi.lib().b200_ar_all_reduce(ctypes.c_int64(self._ctx), ctypes.c_void_p(inp.data_ptr()),
                                                  ctypes.c_void_p(out.data_ptr()), ctypes.c_int64(inp.numel()),
                                                  _DT[inp.dtype], ctypes.c_void_p(st)))
        return out

    def fuse_next_gemm(self, m: int, n: int) -> bool:
        """Ask the NEXT gptq_marlin_gemm issued by this thread ([m, n] output) to all-reduce its output inside its own
        epilogue.  False when the shape does not qualify: the caller then runs GEMM + all_reduce as two kernels."""
        if self.disabled:
            return False
        return capi.lib().b200_ar_fuse_next_gemm(ctypes.c_int64(self._ctx), ctypes.c_int64(m), ctypes.c_int64(n)) == 0

    def all_reduce_(self, inp: torch.Tensor) -> torch.Tensor:
        """In place (what the decode step uses: the GEMM output is overwritten by the sum)."""
        return self.all_reduce(inp, inp)

    def custom_all_reduce(self, input: torch.Tensor) -> Optional[torch.Tensor]:
        if not self.should_custom_ar(input):
            return None
        return self.all_reduce(input)

    def close(self):
        if self._ctx is not None:
            lib = capi.lib()
            lib.b200_ar_dispose(ctypes.c_int64(self._ctx))
            self._ctx = None
            for p in self._peers:
                lib.b200_ar_close(ctypes.c_void_p(p))
            self._peers = []
            if self._own:
                lib.b200_ar_free(ctypes.c_void_p(self._own))
                self._own = None
        self.disabled = True
