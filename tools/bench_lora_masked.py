"""Masked skinny LoRA GEMMs (csrc/lora_dropout.cu) vs their unmasked tcgen05 equivalents and the HBM floor, M = 16384 tokens."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import peaks
from phantom_vlb_b200 import ops

M = 16384
_, hbm, _ = peaks()


def timeit(fn, it=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3


g = torch.Generator(device="cuda").manual_seed(0)
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
for K, na in ((4096, 3), (4096, 2), (4096, 1), (14336, 1)):
    xs = [torch.randn(M, K, device="cuda", generator=g).bfloat16() for _ in range(3)]      # rotate inputs: > L2
    A = (torch.randn(16 * na, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
    t = torch.randn(M, 16 * na, device="cuda", generator=g).bfloat16()
    bits = torch.empty((na, K // 64, M), device="cuda", dtype=torch.int64)
    for a in range(na):
        ops.dropout_bits(M, K, 0.1, 3 + a, bits[a])
    bl = [bits[a] for a in range(na)]
    xa = torch.empty(M, 16 * na, device="cuda", dtype=torch.bfloat16)
    dA = torch.zeros(16 * na, K, device="cuda")
    i = [0]

    def nxt():
        i[0] += 1
        return xs[i[0] % 3]
    floor = (M * K * 2 + na * M * K / 8) / hbm / 1e3
    res = dict(K=K, n_adapters=na, hbm_floor_us=round(floor, 1),
               bits_us=round(timeit(lambda: ops.dropout_bits(M, K, 0.1, 7, bits[0])), 1),
               down_masked_us=round(timeit(lambda: ops.lora_down_masked(nxt(), A, xa, bl, 2.0)), 1),
               down_plain_us=round(timeit(lambda: ops.gemm(nxt(), A, xa, alpha=2.0)), 1),
               grad_a_masked_us=round(timeit(lambda: ops.lora_grad_a_masked(nxt(), t, dA, bl, 1.0)), 1),
               grad_a_plain_us=round(timeit(lambda: ops.gemm(t, nxt(), dA, a_mn=True, b_mn=True, out_f32=True, accumulate=True)), 1))
    print(json.dumps(res), flush=True)
    del xs
