"""A few launches of the masked skinny LoRA GEMMs for ncu: python tools/ncu_lora_down.py [K] [n_adapters]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops

K = int(sys.argv[1]) if len(sys.argv) > 1 else 14336
na = int(sys.argv[2]) if len(sys.argv) > 2 else 1
M = 16384
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
A = (torch.randn(16 * na, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
t = torch.randn(M, 16 * na, device="cuda", generator=g).bfloat16()
bits = torch.empty((na, K // 64, M), device="cuda", dtype=torch.int64)
for a in range(na):
    ops.dropout_bits(M, K, 0.1, 3 + a, bits[a])
bl = [bits[a] for a in range(na)]
xa = torch.empty(M, 16 * na, device="cuda", dtype=torch.bfloat16)
dA = torch.zeros(16 * na, K, device="cuda")
for _ in range(3):
    ops.lora_down_masked(x, A, xa, bl, 2.0)
    ops.lora_grad_a_masked(x, t, dA, bl, 1.0)
torch.cuda.synchronize()
print("ok")
