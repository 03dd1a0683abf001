"""A few launches of the HBM-bound row kernels named by north_star, at the benchmark's sizes, for one ncu --set full pass:
rmsnorm_bwd (M=16384, H=4096), hrf_pool_fwd/bwd (B=8, S=2048, H=4096), pearson_update ([5632, 250000] fp32), se_pool."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200 import ops
from phantom_vlb_b200.litmodule import VLBLitModule

M, H = 16384, 4096
dy = torch.randn(M, H, device="cuda").bfloat16(); x = torch.randn(M, H, device="cuda").bfloat16()
w = torch.ones(H, device="cuda").bfloat16(); rstd = torch.rand(M, device="cuda") + 0.5
dres = torch.randn(M, H, device="cuda").bfloat16(); out = torch.empty_like(x)
for _ in range(2):
    ops.rmsnorm_bwd(dy, x, w, rstd, dres=dres, out=out)
xs = torch.randn(64 * 576, 4096, device="cuda").bfloat16()
for _ in range(2):
    ops.se_pool(xs, 64, 576)
N, P = 5632, 250000
a = torch.randn(N, P, device="cuda"); b = torch.randn(N, P, device="cuda")
for _ in range(2):
    ops.pearson(a, b)
del a, b
m = VLBLitModule(lit_config())
m.configure_model(device=torch.device("cuda", 0))
batch = {k: v.cuda() for k, v in host_batches(m.model_config, 8, 1, 1234)[0].items()}
m.engine.train_step(batch, 1e-4)          # hrf_pool_fwd / finalize / bwd inside one step at B=8
torch.cuda.synchronize()
print("ok")
