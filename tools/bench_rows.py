import sys; sys.path.insert(0,'/root/repo')
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit
M,H=16384,4096
dy=torch.randn(M,H,device='cuda').bfloat16(); x=torch.randn(M,H,device='cuda').bfloat16(); w=torch.ones(H,device='cuda').bfloat16()
rstd=torch.rand(M,device='cuda')+0.5; dres=torch.randn(M,H,device='cuda').bfloat16(); out=torch.empty_like(x)
flush=torch.empty(256<<20,device='cuda',dtype=torch.uint8)
def f(): flush.zero_(); ops.rmsnorm_bwd(dy,x,w,rstd,dres=dres,out=out)
t=timeit(f,iters=20)-timeit(lambda: flush.zero_(),iters=20)
print('rmsnorm_bwd us %.1f GB/s %.0f'%(t*1e3, 4*M*H*2/t/1e6))
gu=torch.randn(M,2*14336,device='cuda').bfloat16(); act=torch.empty(M,14336,device='cuda').bfloat16()
def g(): flush.zero_(); ops.swiglu_fwd(gu,out=act)
t=timeit(g,iters=20)-timeit(lambda: flush.zero_(),iters=20)
print('swiglu_fwd us %.1f GB/s %.0f'%(t*1e3, 3*M*14336*2/t/1e6))
