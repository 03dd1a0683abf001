"""Forward-only determinism check at full size: the forward has no atomics, so every repetition must be bit-identical."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200.litmodule import VLBLitModule
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
m = VLBLitModule(lit_config())
m.configure_model(device=torch.device("cuda", 0))
batch = {k: v.cuda() for k, v in host_batches(m.model_config, B, 1, 1234)[0].items()}
ref = None
bad = 0
for i in range(n):
    out = m.validation_step(batch)
    p = out["brain_preds"].float().clone()
    if ref is None:
        ref = p
        print("run 0 loss", float(out["loss"]), "nan", bool(torch.isnan(p).any()))
    elif not torch.equal(p, ref):
        bad += 1
        d = (p - ref).abs()
        print("run", i, "differs: max", float(d.max()), "nan", bool(torch.isnan(p).any()), "loss", float(out["loss"]))
print("done, differing runs:", bad, "of", n - 1)
