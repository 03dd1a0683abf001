"""Hand-off timeline of one attention-forward CTA (blockIdx 0 = the heaviest causal query tile): clock64 stamps per key block.
usage: VLB_ATT_TS=1 python tools/attn_timeline.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import lib as _l, ops
B, S, Hq, Hkv, D = int(os.environ.get("TL_B", "4")), 2048, 32, 8, 128
qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
for _ in range(3):
    ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
torch.cuda.synchronize()
n = 32 * 8
buf = (ctypes.c_longlong * n)()
_l.call("vlb_attn_debug_ts", ctypes.cast(buf, ctypes.c_void_p).value, n)
t = [buf[i] for i in range(n)]
t0 = min(x for x in t if x > 0)
names = ["sm:S seen", "sm:ld done", "sm:exp done", "sm:P arrive", "is:P seen", "is:V seen", "is:PV issued", "is:QK(j+2) issued"]
print("j   " + "  ".join("%-9s" % s[:9] for s in names))
for j in range(32):
    print("%-3d " % j + "  ".join("%9d" % (t[j * 8 + k] - t0) for k in range(8)))

if len(sys.argv) > 1 and sys.argv[1] == "bwd":
    do = torch.randn(B * S, Hq * D, device="cuda").bfloat16()
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
    for _ in range(2):
        ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, True)
    torch.cuda.synchronize()
    n = 128 * 16
    buf = (ctypes.c_longlong * n)()
    _l.call("vlb_attn_bwd_debug_ts", ctypes.cast(buf, ctypes.c_void_p).value, n)
    t = [buf[i] for i in range(n)]
    t0 = min(x for x in t if x > 0)
    names = ["sm:qdo", "sm:S", "sm:A done", "sm:dP", "sm:arrive", "sm:mmadone", "sm:drained", "is:pds", "is:dP+1", "is:dVdK", "is:dqfree",
             "is:dQ", "is:S+2"]
    print("t   " + " ".join("%-10s" % s for s in names))
    for j in range(8, 28):
        print("%-3d " % j + " ".join("%10d" % (t[j * 16 + k] - t0) for k in range(13)))
