"""Same process, same engine, same box: the 7B step with lora_dropout / dropout_rate 0.1 and 0.0 in alternating blocks, so that the
cost of dropout is not lost in box-to-box and warm-up drift (the two bench.py runs of a pair differ by 1-2 % on their own)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import host_batches, lit_config
from phantom_vlb_b200.litmodule import VLBLitModule

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 4
block = int(sys.argv[3]) if len(sys.argv) > 3 else 4
m = VLBLitModule(lit_config(lora_dropout=0.1, dropout_rate=0.1))
m.configure_model(device=torch.device("cuda", 0))
eng = m.engine
batches = [{k: v.cuda() for k, v in b.items()} for b in host_batches(m.model_config, B, 2, 1234)]
for i in range(4):
    eng.train_step(batches[i & 1], 1e-4)
res = {"drop": [], "nodrop": []}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for r in range(rounds):
    for name, p in (("drop", 0.1), ("nodrop", 0.0)) if r % 2 == 0 else (("nodrop", 0.0), ("drop", 0.1)):
        eng.lora_dropout = p
        eng.dropout_rate = p
        eng.train_step(batches[0], 1e-4)            # one untimed step after the switch
        torch.cuda.synchronize()
        e0.record()
        for i in range(block):
            eng.train_step(batches[i & 1], 1e-4)
        e1.record()
        torch.cuda.synchronize()
        res[name].append(e0.elapsed_time(e1) / block)
d, n = sum(res["drop"]) / rounds, sum(res["nodrop"]) / rounds
print(json.dumps(dict(B=B, rounds=rounds, block=block, ms_dropout=round(d, 2), ms_no_dropout=round(n, 2), overhead_pct=round(100 * (d / n - 1), 2),
                      per_round_dropout=[round(x, 1) for x in res["drop"]], per_round_no_dropout=[round(x, 1) for x in res["nodrop"]])))
