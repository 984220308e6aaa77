"""Small-vector all-reduce over NVLink peer memory (nf_allreduce_small_f32).

torch is used only to allocate and map the exchange buffers (symmetric memory: every rank gets a
device pointer to every peer's buffer); the exchange itself is one libnfcore kernel per step.  When the
peers cannot be mapped (no P2P, more than 8 ranks, vector too long) callers fall back to NCCL.
"""
import ctypes

import torch

from . import _lib
from ._lib import NFError


class SmallAllReduce:
    def __init__(self, group=None, device=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        lib = _lib.load()
        self.group = dist.group.WORLD if group is None else group
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        nbytes = lib.nf_allreduce_small_buffer_bytes(self.world)
        if nbytes == 0:
            raise NFError('SmallAllReduce supports at most 8 ranks')
        self.max_count = lib.nf_allreduce_small_max_count()
        self.buf = symm_mem.empty((nbytes + 3) // 4, dtype=torch.float32, device=self.device)
        self.buf.zero_()
        torch.cuda.synchronize(self.device)
        self.handle = symm_mem.rendezvous(self.buf, self.group)
        ptrs = list(self.handle.buffer_ptrs)
        self._ptrs = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(p) for p in ptrs])
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.seq = 0
        dist.barrier(self.group)

    def __call__(self, src, dst=None):
        """dst (default: in place) = sum over ranks of src; enqueued on the current stream."""
        if src.dtype != torch.float32 or not src.is_contiguous() or src.numel() > self.max_count:
            raise NFError('SmallAllReduce: float32 contiguous vector of at most %d elements' % self.max_count)
        dst = src if dst is None else dst
        self.seq += 1
        st = torch.cuda.current_stream(self.device).cuda_stream
        rc = _lib.load().nf_allreduce_small_f32(self._ptrs, self.world, self.rank, ctypes.c_void_p(src.data_ptr()),
                                                ctypes.c_void_p(dst.data_ptr()), src.numel(), ctypes.c_uint64(self.seq),
                                                ctypes.c_void_p(self.status.data_ptr()), ctypes.c_void_p(st))
        _lib.check(rc)
        return dst

    def counted(self, vec, n, n_local):
        """In place: vec[:n] = sums over ranks, vec[n:n+3] = (n_hi, n_lo, status) -- nf_allreduce_small_counted_f32.
        The caller reads vec back and passes it to `global_count`."""
        if vec.dtype != torch.float32 or not vec.is_contiguous() or vec.numel() < n + 3 or n > self.max_count:
            raise NFError('SmallAllReduce.counted: float32 contiguous vector of n + 3 <= %d elements' % (self.max_count + 3))
        self.seq += 1
        st = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.load().nf_allreduce_small_counted_f32(self._ptrs, self.world, self.rank, ctypes.c_void_p(vec.data_ptr()), n,
                                                              int(n_local), ctypes.c_uint64(self.seq), ctypes.c_void_p(st)))
        return vec

    @staticmethod
    def global_count(tail):
        """(n_hi, n_lo, status) as read back from the device -> global sample count; raises when a peer did not arrive."""
        if float(tail[2]) != 0.0:
            raise NFError('peer exchange: a rank did not arrive within the time limit; the sums are not valid')
        return int(round(float(tail[0]))) * (1 << 20) + int(round(float(tail[1])))

    def next_seq(self):
        self.seq += 1
        return self.seq

    def check(self):
        """Synchronising: raises if a peer failed to arrive in some earlier call."""
        if int(self.status.item()) != 0:
            raise NFError('SmallAllReduce: a peer did not arrive (bounded spin expired)')
