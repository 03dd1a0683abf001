"""One full 7B training step between cudaProfilerStart/Stop, for ncu --profile-from-start off."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200.litmodule import VLBLitModule

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
warm = int(sys.argv[2]) if len(sys.argv) > 2 else 2
cfg = lit_config()
m = VLBLitModule(cfg)
m.configure_model(device=torch.device("cuda", 0))
batch = {k: v.cuda() for k, v in host_batches(m.model_config, B, 1, 1234)[0].items()}
for _ in range(warm):
    m.engine.train_step(batch, 1e-4)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
out = m.engine.train_step(batch, 1e-4)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("loss", float(out[0]), "launches/step", m.engine.launches // (warm + 1))
