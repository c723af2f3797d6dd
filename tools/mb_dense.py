import torch, time, sys
sys.path.insert(0, '/root/repo')
from gcnn_keras_b200 import functional as F
torch.backends.cuda.matmul.allow_tf32 = False
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b)/n*1e3
for (M,K,U) in [(18432,128,128),(404544,128,128),(404544,20,384),(2000000,128,128),(16384,640,1050)]:
    x=torch.randn(M,K,device='cuda'); w=torch.randn(K,U,device='cuda'); g=torch.randn(M,U,device='cuda')
    r=[t(lambda: F.raw_dense_tc(0,x,w)), t(lambda: x@w), t(lambda: F.raw_dense_tc(1,g,w)), t(lambda: g@w.t()), t(lambda: F.raw_dense_tc(2,x,g)), t(lambda: F.raw_gemm_tn(x,g)), t(lambda: x.t()@g)]
    print(M,K,U, "NN tc %.1f cublas %.1f | NT tc %.1f cublas %.1f | TN tc %.1f simt %.1f cublas %.1f us" % tuple(r), "NN GB/s %.0f TFLOPs %.1f" % ((M*K+M*U)*4/r[0]/1e3, 2*M*K*U/r[0]/1e6))
