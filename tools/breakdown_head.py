"""Per-call time breakdown of the cached-feature head step at P voxels (config 3). usage: breakdown_head.py [P] [N]"""
import collections, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import lib
from phantom_vlb_b200.engine import Engine
from phantom_vlb_b200.model import FrozenWeights, ModelConfig, Trainables

P = int(sys.argv[1]) if len(sys.argv) > 1 else 250000
N = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dev = torch.device("cuda", 0)
mc = ModelConfig(num_frames=16, num_target=P, use_lora=False, layers=0, vit_layers=2)
W = FrozenWeights()
W.llm = {"norm_w": torch.ones(mc.hidden, device=dev, dtype=torch.bfloat16)}
T = Trainables(mc, dev); T.init_default(1)
eng = Engine(mc, W, T, device=dev)
if os.environ.get("VLB_HEAD_FUSED") is not None:
    eng.fused_head_opt = os.environ["VLB_HEAD_FUSED"] == "1"
g = torch.Generator(device=dev).manual_seed(7)
u = torch.randn((N, mc.hidden), device=dev, generator=g); sw = torch.rand((N,), device=dev, generator=g) + 0.5
y = torch.randn((N, P), device=dev, generator=g)
for _ in range(3):
    eng.cached_loss_and_grads(u, sw, y); eng.optimizer_step(1e-4)
lib.profile_begin()
eng.cached_loss_and_grads(u, sw, y); eng.optimizer_step(1e-4)
rows = lib.profile_end()
tot = sum(ms for _, ms, _ in rows)
print("fused_head_opt", eng.fused_head_opt, "total %.3f ms" % tot)
for name, ms, a in rows:
    extra = " M=%d N=%d K=%d" % (a[6], a[7], a[8]) if name == "vlb_gemm_bf16" else ""
    print("%-26s %8.3f ms%s" % (name, ms, extra))
