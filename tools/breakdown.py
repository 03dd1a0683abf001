"""Per-entry-point time breakdown of one 7B training step (CUDA events around every C-ABI call, in situ).
usage: python tools/breakdown.py [B] [warmup] [lora_dropout] [head_dropout]   -- prints a table; GEMMs are classed by shape role."""
import collections
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200 import lib
from phantom_vlb_b200.litmodule import VLBLitModule

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
warm = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg = lit_config(lora_dropout=float(sys.argv[3]) if len(sys.argv) > 3 else 0.0,
                 dropout_rate=float(sys.argv[4]) if len(sys.argv) > 4 else 0.0)
m = VLBLitModule(cfg)
m.configure_model(device=torch.device("cuda", 0))
batch = {k: v.cuda() for k, v in host_batches(m.model_config, B, 1, 1234)[0].items()}
for _ in range(warm):
    m.engine.train_step(batch, 1e-4)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    m.engine.train_step(batch, 1e-4)
e1.record()
torch.cuda.synchronize()
print("plain step: %.2f ms" % (e0.elapsed_time(e1) / 3))
lib.profile_begin()
m.engine.train_step(batch, 1e-4)
rows = lib.profile_end()
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for name, ms, a in rows:
    fl = 0.0
    if name == "vlb_gemm_dx_lora_masked":
        name += "[N=%d,K=%d,na=%d]" % (a[7], a[8], a[13])
    if name in ("vlb_lora_down_masked", "vlb_lora_grad_a_masked"):
        name += "[K=%d,na=%d]" % (a[7], a[8])
    if name == "vlb_gemm_bf16":
        M, N, K, K2, f32 = a[6], a[7], a[8], a[13] if a[9] else 0, a[21]
        fl = 2.0 * M * N * (K + K2)
        if f32 and M <= 64:
            name = "gemm:lora_grad[M<=64,f32]"
        elif N <= 64:
            name = "gemm:xA[N<=64]"
        elif M <= 64:
            name = "gemm:skinnyM"
        elif K <= 64:
            name = "gemm:K<=64"
        else:
            name = "gemm:big"
    agg[name][0] += 1
    agg[name][1] += ms
    agg[name][2] += fl
tot = sum(v[1] for v in agg.values())
print("instrumented sum: %.2f ms over %d calls" % (tot, len(rows)))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    tf = " %7.1f TF/s" % (v[2] / v[1] / 1e9) if v[2] else ""
    print("%-30s %5d %9.3f ms %5.1f%%%s" % (k, v[0], v[1], 100 * v[1] / tot, tf))
