import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import torch, nfcore
lib = nfcore.load(); dev = 'cuda:0'
p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
M, N, K = [int(v) for v in sys.argv[1:4]]
A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev); bias = torch.randn(N, device=dev)
C = torch.empty(M, N, device=dev); CT = torch.empty(N, M, device=dev)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3): lib.nf_gemm_tn_tf32x3(p(A), K, p(B), K, M, N, K, p(C), N, p(CT), M, p(bias), 1, st)
torch.cuda.synchronize()
