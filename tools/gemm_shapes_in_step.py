"""Per-shape GEMM throughput inside one instrumented 7B train step (CUDA events around every GEMM launch)."""
import collections, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200.litmodule import VLBLitModule
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
m = VLBLitModule(lit_config())
m.configure_model(device=torch.device("cuda", 0))
batch = {k: v.cuda() for k, v in host_batches(m.model_config, B, 1, 1234)[0].items()}
eng = m.engine
for _ in range(2):
    eng.train_step(batch, 1e-4)
torch.cuda.synchronize()
eng.profile = []
eng.train_step(batch, 1e-4)
torch.cuda.synchronize()
agg = collections.OrderedDict()
for e0, e1, fl, shp in eng.profile:
    t = e0.elapsed_time(e1)
    a = agg.setdefault(shp, [0, 0.0, 0.0])
    a[0] += 1; a[1] += t; a[2] += fl
eng.profile = None
tot = sum(a[1] for a in agg.values())
print("total GEMM ms %.1f" % tot)
for shp, (n, t, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print("M=%6d N=%6d K=%6d K2=%3d  n=%4d  ms=%7.2f  %5.1f%%  %7.1f TF/s" % (shp + (n, t, 100 * t / tot, fl / t / 1e9)))
