"""Losses of a few consecutive 7B train steps on one fixed batch (sanity: finite, decreasing)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200.litmodule import VLBLitModule
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4
m = VLBLitModule(lit_config())
m.configure_model(device=torch.device("cuda", 0))
batch = {k: v.cuda() for k, v in host_batches(m.model_config, B, 1, 1234)[0].items()}
ls = []
for i in range(n):
    out = m.engine.train_step(batch, 1e-4)
    ls.append(float(out[0]))
torch.cuda.synchronize()
print("losses", ls, "grad sumsq", float(m.engine.T.sumsq))
