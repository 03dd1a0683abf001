"""Long-K pair GEMMs (down projection forward K=14336, gate/up dX K=28672) with and without the serpentine tile / K order
(VLB_GEMM_NO_SERP toggled between launches): CUDA-event timings of 10 back-to-back launches each; under ncu (-k regex:gemm2
--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum) the DRAM bytes per launch."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops

M = 16384
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for name, N, K, bmn in (("d", 4096, 14336, False), ("d_gu", 4096, 28672, True)):
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn((K, N) if bmn else (N, K), device="cuda") * 0.02).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for serp in (1, 0, 1, 0):
        if serp:
            os.environ.pop("VLB_GEMM_NO_SERP", None)
        else:
            os.environ["VLB_GEMM_NO_SERP"] = "1"
        ops.gemm(a, w, out, b_mn=bmn)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            ops.gemm(a, w, out, b_mn=bmn)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print("%s K=%d serpentine=%d: %.3f ms  %.0f TF/s" % (name, K, serp, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
    del a, w, out
