"""Check and time the tcgen05 GEMM against a float64 matmul (tuning / bring-up aid)."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import torch, nfcore
lib = nfcore.load()
dev = 'cuda:0'
p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
torch.manual_seed(0)
for (M, N, K, act) in [(128, 64, 32, 0), (128, 64, 64, 0), (256, 128, 96, 1), (4096, 64, 784, 1), (4096, 1376, 64, 0), (1000, 100, 40, 2), (64, 784, 4096, 0), (1376, 64, 4096, 0)]:
    A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev) / K ** 0.5; bias = torch.randn(N, device=dev)
    C = torch.full((M, N), float('nan'), device=dev); CT = torch.full((N, M), float('nan'), device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.nf_gemm_tn_tf32x3(p(A), K, p(B), K, M, N, K, p(C), N, p(CT), M, p(bias), act, st)
    torch.cuda.synchronize()
    ref = A.double() @ B.double().T + bias.double()
    if act == 1: ref = torch.relu(ref)
    if act == 2: ref = torch.sigmoid(ref)
    err = float((C.double() - ref).abs().max() / ref.abs().max())
    errT = float((CT.double().T - ref).abs().max() / ref.abs().max())
    tf32 = float(((A @ B.T + bias).double() - (A.double() @ B.double().T + bias.double())).abs().max() / ref.abs().max()) if act == 0 else float('nan')
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): lib.nf_gemm_tn_tf32x3(p(A), K, p(B), K, M, N, K, p(C), N, None, 0, p(bias), act, st)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    print('M=%5d N=%5d K=%5d act=%d rc=%d  rel.err C %.2e  CT %.2e  (torch fp32 matmul err %.1e)  %.1f us  %.2f TFLOP/s' %
          (M, N, K, act, rc, err, errT, tf32, us, 2.0 * M * N * K / us / 1e6), flush=True)
