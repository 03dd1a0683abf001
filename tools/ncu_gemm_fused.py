"""Two launches each of the SwiGLU-fused CTA-pair GEMMs at the 7B decoder shapes (M = B*2048, LoRA k-block included), for
`ncu -k regex:gemm2 --set full` captures (profiles/r01d_ncu_full_gemm2_swiglu.json)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
M = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
H, I, r = 4096, 14336, 16
g = torch.Generator(device="cuda").manual_seed(0)
rn = lambda *s: torch.randn(s, device="cuda", generator=g)
x = rn(M, H).bfloat16()
wgu = (rn(2 * I, H) * 0.02).bfloat16()
xa = rn(M, 2 * r).bfloat16()
bcat = torch.zeros(2 * I, 64, device="cuda", dtype=torch.bfloat16)
bcat[:, :2 * r] = (rn(2 * I, 2 * r) * 0.02).bfloat16()
xa64 = torch.zeros(M, 64, device="cuda", dtype=torch.bfloat16)
xa64[:, :2 * r] = xa
gu = torch.empty(M, 2 * I, device="cuda", dtype=torch.bfloat16)
act = torch.empty(M, I, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    ops.gemm_gateup_swiglu(x, wgu, gu, act, a2=xa64, b2=bcat)
torch.cuda.synchronize()
wd = (rn(H, I) * 0.02).bfloat16()
t = rn(M, r).bfloat16()
a_d = (rn(r, I) * 0.02).bfloat16()
dgu = torch.empty(M, 2 * I, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    ops.gemm_down_dx_swiglu_bwd(x, wd, gu, dgu, a2=t, b2=a_d)
torch.cuda.synchronize()
