"""GPU tests of the sample-sharded exchange (csrc/nf_allreduce.cu and the last block of the packed-pair kernel).

world = 1 runs everywhere: the exchange degenerates to "push into my own slot, wait for my own flag, add one slot",
which exercises the slot layout, both slot sets, the sequence logic, the count and the status word through the C ABI.
The multi-rank run (tools/dist_check.py under torchrun) needs at least two GPUs and is skipped otherwise.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _own_buffer(lib, dev):
    nbytes = lib.nf_allreduce_small_buffer_bytes(1)
    buf = torch.zeros((nbytes + 3) // 4, dtype=torch.float32, device=dev)
    ptrs = (ctypes.c_void_p * 1)(ctypes.c_void_p(buf.data_ptr()))
    return buf, ptrs


def test_small_allreduce_world1_through_the_abi():
    from nfcore import _lib
    lib = _lib.load()
    dev = torch.device('cuda:0')
    buf, ptrs = _own_buffer(lib, dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    nmax = lib.nf_allreduce_small_max_count()
    assert nmax >= 389 + 2                                   # K = 64, D = 2 result vector fits
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    g = torch.Generator(device=dev).manual_seed(0)
    for seq, n in enumerate((195, 1, nmax, 0, 389), start=1):
        src = torch.randn(max(n, 1), device=dev, generator=g)
        dst = torch.full((max(n, 1),), float('nan'), device=dev)
        assert lib.nf_allreduce_small_f32(ptrs, 1, 0, ctypes.c_void_p(src.data_ptr()), ctypes.c_void_p(dst.data_ptr()), n,
                                          ctypes.c_uint64(seq), ctypes.c_void_p(status.data_ptr()), st) == 0
        torch.cuda.synchronize()
        assert torch.equal(dst[:n], src[:n]) and int(status.item()) == 0
    # in place, with the sample count riding along; counts above 2^24 stay exact
    for seq, (n, cnt) in enumerate(((195, 1 << 20), (7, 123456789), (389, (1 << 26) + 3)), start=6):
        vec = torch.randn(n + 3, device=dev, generator=g)
        ref = vec.clone()
        assert lib.nf_allreduce_small_counted_f32(ptrs, 1, 0, ctypes.c_void_p(vec.data_ptr()), n, cnt, ctypes.c_uint64(seq), st) == 0
        torch.cuda.synchronize()
        assert torch.equal(vec[:n], ref[:n])
        tail = vec[n:].cpu().numpy()
        assert int(tail[0]) * (1 << 20) + int(tail[1]) == cnt and tail[2] == 0.0
    # argument checks: too long, sequence number 0, bad rank
    v = torch.zeros(nmax + 8, device=dev)
    assert lib.nf_allreduce_small_f32(ptrs, 1, 0, ctypes.c_void_p(v.data_ptr()), ctypes.c_void_p(v.data_ptr()), nmax + 1,
                                      ctypes.c_uint64(9), ctypes.c_void_p(status.data_ptr()), st) == -1
    assert lib.nf_allreduce_small_counted_f32(ptrs, 1, 0, ctypes.c_void_p(v.data_ptr()), 4, 1, ctypes.c_uint64(0), st) == -1
    assert lib.nf_allreduce_small_counted_f32(ptrs, 1, 1, ctypes.c_void_p(v.data_ptr()), 4, 1, ctypes.c_uint64(9), st) == -1


@pytest.mark.parametrize('K,N', [(32, 1 << 18), (16, 200001), (64, 160000), (8, 5000)])
def test_exchange_entry_world1_equals_plain_entry(K, N):
    """nf_planar_elbo_fwd_bwd_exchange_f32 with one rank returns bitwise what nf_planar_elbo_fwd_bwd_f32 returns, plus
    the sample count and a clear status -- on the single-launch kernel (large N) and on the two-launch path (N = 5000)."""
    from nfcore import _lib, ops, semantics as sem
    lib = _lib.load()
    dev = torch.device('cuda:0')
    D = 2
    rs = np.random.RandomState(K)
    W = 0.1 + 0.01 * rs.randn(K, D); U = 0.1 + 0.01 * rs.randn(K, D); B = 0.1 * np.ones(K)
    S, c = sem.forward(W, U)
    a = [torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (W, S, B, c)]
    eps = torch.randn(N, D, dtype=torch.float32, device=dev, generator=torch.Generator(device=dev).manual_seed(2))
    plain = ops.planar_elbo(eps, *a, 'p3')
    ref = torch.cat([plain[0], plain[1].reshape(-1), plain[2].reshape(-1), plain[3], plain[4]])
    buf, ptrs = _own_buffer(lib, dev)

    class OneRank:                      # what nfcore.peer.SmallAllReduce provides, without a process group
        world, rank, seq = 1, 0, 0
        _ptrs = ptrs

        def next_seq(self):
            self.seq += 1
            return self.seq
    ar = OneRank()
    nvec = 3 + 2 * K * D + 2 * K
    for rep in range(3):                # three calls: both slot sets, increasing sequence numbers
        out = torch.full((nvec + 3,), float('nan'), dtype=torch.float32, device=dev)
        ops.planar_elbo_exchange(eps, *a, 'p3', None, 1e-7, out, ar)
        torch.cuda.synchronize()
        assert torch.equal(out[:nvec], ref), float((out[:nvec] - ref).abs().max())
        tail = out[nvec:].cpu().numpy()
        assert int(tail[0]) * (1 << 20) + int(tail[1]) == N and tail[2] == 0.0


def test_sharded_objective_on_two_or_more_gpus():
    """tools/dist_check.py under torchrun: the sharded objective (exchange inside the kernel's last block, the
    two-launch path, NCCL for float64) equals the single-GPU objective on the concatenated samples; 50 rounds of the
    peer exchange equal NCCL."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip('needs at least two GPUs')
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world), '--master-addr', '127.0.0.1',
           '--master-port', '29533', os.path.join(ROOT, 'tools', 'dist_check.py')]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(r.stdout[-4000:])
    assert r.returncode == 0, r.stdout[-4000:]
