#!/usr/bin/env python
"""Headline benchmark: LoRA train clips/sec for VideoLLaMA2-7B (CLIP ViT-L/14-336 + STC + Mistral-7B, r=16, bf16),
8-frame 336^2 clips, 1000-parcel head -- BASELINE.json configs[1] at N=1, configs[2] (data-parallel) at N>1.

    python bench.py --gpus N --steps K --warmup W            # ours (torchrun launches N ranks when N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # reference CPU path (oracle port) on the host cores

One "step" = one full training step (ViT + STC forward, decoder forward/backward with fused LoRA, brain head,
gradient all-reduce when N > 1, global-norm clip + AdamW) on `--batch` synthetic clips per GPU.
Prints ONE JSON line (rank 0). See the keys' definitions in DESIGN.md section "Measurement".
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# algorithmic FLOPs per 8-frame clip of one LoRA train step (SURVEY 8d / BASELINE.md section 3)
TF_DECODER_LAYER = 61.54 / 32
TF_VIT_LAYER = 2.93 / 23
TF_STC = 1.67
TF_FULL = 66.1


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def collate(items):
    import numpy as np
    import torch
    out = {}
    for k in items[0]:
        v = [it[k] for it in items]
        out[k] = torch.stack([x if isinstance(x, torch.Tensor) else torch.from_numpy(np.asarray(x)) for x in v], 0)
    return out


def host_batches(mc, B, n, seed):
    from phantom_vlb_b200.datamodule import SyntheticVLBDataset
    ds = SyntheticVLBDataset(mc, length=B * n, seed=seed)
    res = []
    for i in range(n):
        b = collate([ds[i * B + j] for j in range(B)])
        res.append({k: v.contiguous().pin_memory() for k, v in b.items()})
    return res


def lit_config(P=1000, lora_dropout=0.0, dropout_rate=0.0):
    from phantom_vlb_b200.litmodule import VLBLitModuleConfig
    # config/experiment/VLB_vllama2_friends_lora.yaml:19-40. Default dropouts 0 (deterministic, parity-exact); the yaml's
    # lora_dropout 0.1 / dropout_rate 0.1 are reachable with --lora-dropout / --head-dropout (masked-adapter path).
    return VLBLitModuleConfig(model_path="random:7b", freeze_backbone=False, use_lora=True, lora_r=16, lora_alpha=32,
                              lora_dropout=lora_dropout, dropout_rate=dropout_rate, num_target=P, l2_lambda=0.001, lr=1e-4,
                              betas=[0.9, 0.999], eps=1e-8, weight_decay=1e-2, lr_scheduler_name="CosineAnnealingLR",
                              last_epoch=-1, t_max=50000)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from phantom_vlb_b200.engine import cosine_lr
    from phantom_vlb_b200.litmodule import VLBLitModule

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, K, Wm = args.batch, args.steps, max(args.warmup, 3)
    cfg = lit_config(lora_dropout=args.lora_dropout, dropout_rate=args.head_dropout)
    module = VLBLitModule(cfg)
    module.configure_model(device=dev, seed=1234)            # identical weights on every rank
    eng, mc = module.engine, module.model_config
    hosts = host_batches(mc, B, 2, seed=1234 + rank)         # rank-private clips
    devs = [{k: v.to(dev) for k, v in h.items()} for h in hosts]
    allreduce = (lambda t: dist.all_reduce(t)) if world > 1 else None
    h2d = sum(v.numel() * v.element_size() for v in hosts[0].values())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident arm (value)
    def step_dev(i):
        eng.train_step(devs[i % 2], cosine_lr(cfg.lr, eng.T.step_count, cfg.t_max), world_size=world, allreduce=allreduce)

    for i in range(Wm):
        step_dev(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.launches
    ms = timed(step_dev, K)
    launches = eng.launches - l0
    value = world * B * K / (ms / 1e3)

    # ---- end-to-end arm through the public API with pinned host buffers (e2e)
    (opt,), (sch,) = module.configure_optimizers()
    opt.world_size, opt.allreduce = world, allreduce
    sched = sch["scheduler"]
    losses = []

    def step_e2e(i):
        loss = module.training_step(hosts[i % 2])            # H2D copies of the batch happen inside
        loss.backward()
        opt.step()
        sched.step()
        opt.zero_grad(set_to_none=True)
        losses.append(float(loss.item()))                     # D2H read of the step's loss

    step_e2e(0)
    ms_e2e = timed(step_e2e, K)
    e2e_value = world * B * K / (ms_e2e / 1e3)
    clocks = sampler.stop() if rank == 0 else None

    # ---- roofline of the dominant kernel (gemm_tcgen05_kernel, LoRA fused): one instrumented step, CUDA events per launch
    eng.profile = []
    step_dev(0)
    torch.cuda.synchronize()
    prof, eng.profile = eng.profile, None
    g_ms = sum(e0.elapsed_time(e1) for e0, e1, _, _ in prof)
    g_fl = sum(f for _, _, f, _ in prof)
    big = [(e0.elapsed_time(e1), f) for e0, e1, f, s in prof if s[1] >= 1024 and s[0] >= 1024]
    peak_tf, peak_bw, peak_src = peaks()
    achieved = sum(f for _, f in big) / (sum(t for t, _ in big) / 1e3) / 1e12
    step_ms = ms / K
    roofline = {"bound": "tensor", "kernel": "gemm2_tcgen05_kernel (cta_group::2 pair tiles 256x256) / gemm_tcgen05_kernel: decoder, ViT and STC GEMMs with the fused LoRA k-block",
                "achieved": round(achieved, 1), "peak": peak_tf, "unit": "TFLOP/s", "frac": round(achieved / peak_tf, 4),
                "traffic": 3.17e9 if B == 8 else None,
                "traffic_note": "DRAM read+write bytes of ONE gate/up forward launch with SwiGLU in its epilogue (M=16384, N=28672, "
                                "K=4096+64) from profiles/r01d_ncu_full_gemm2_swiglu.json (ncu --set full, B=8): 1.79 GB read + 1.38 GB "
                                "written vs 1.78 GB algorithmic 2(MK+NK+MN)+2MI (gu and act are both outputs); the weight panel is "
                                "re-read once per 16-block M group",
                "peak_source": peak_src, "launches_timed": len(big),
                "how": "one extra instrumented step: CUDA events around every GEMM launch on the launching stream; achieved = sum of algorithmic 2*M*N*(K+K2) over launches with M,N >= 1024 / sum of their times. The host-side event overhead leaves short idle gaps between launches, so the power-capped SM clock can sit slightly higher than inside the plain timed step",
                "gemm_share_of_step": round(g_ms / step_ms, 4), "all_gemm_tflops": round(g_fl / (g_ms / 1e3) / 1e12, 1),
                "step_tflops": round(B * TF_FULL / (step_ms / 1e3), 1)}
    res = None
    if rank == 0:
        res = {"metric": "lora_train_clips_per_sec", "value": round(value, 4), "unit": "clips/s", "n_gpus": world,
               "steps": K, "warmup": Wm, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
               "config": {"workload": "VideoLLaMA2-7B (CLIP ViT-L/14-336 23L + STC + Mistral-7B 32L) LoRA r=16 bf16, "
                                      "8-frame 336^2 clips, 1000-parcel head, S=2048 (BASELINE configs[%d])" % (1 if world == 1 else 2),
                          "clips_per_gpu": B, "global_batch": B * world, "lora_dropout": args.lora_dropout, "head_dropout": args.head_dropout, "parallelism": "dp%d" % world,
                          "weights": "random-init, frozen bf16; LoRA A kaiming / B zero; fp32 master + Adam state",
                          "l2_policy": "per-step working set (>%d GB activations) far exceeds the 126 MB L2; two alternating input batches" % (B * 9 // 4)},
               "e2e": {"value": round(e2e_value, 4), "unit": "clips/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                       "ms_per_step": round(ms_e2e / K, 3), "last_loss": losses[-1]},
               "gpu_launches": launches, "clocks": clocks, "roofline": roofline}
        if world == 1 and not args.no_cpu_baseline:
            res["cpu_baseline"] = cpu_reference(steps=1, warmup=0, budget_s=60.0)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return res


def cpu_reference(steps, warmup, budget_s):
    """Oracle port (oracle/ref_model.py: HF CLIP/Mistral fp32 + restated STC/LoRA/head) timed on the host cores on a
    bounded sample: BASELINE configs[0] (2 ViT + 2 LLM layers, full widths, one 8-frame 336^2 clip, fwd+bwd), scaled
    to the 7B step by the algorithmic FLOP ratio."""
    import torch
    from oracle.ref_model import RefVLB, config1, enable_training, synthetic_batch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = config1()
    t0 = time.time()
    ref = enable_training(RefVLB(cfg, seed=1234, lora_b_std=0.0))
    params = [p for _, p in ref.trainable()]
    opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-2)
    batch = synthetic_batch(cfg, 1, seed=1234)
    build_s = time.time() - t0

    def one():
        opt.zero_grad(set_to_none=True)
        out = ref.step(batch)
        out["loss"].backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        return float(out["loss"].detach())

    times = []
    t_first = time.time()
    for _ in range(warmup):
        one()
    done = 0
    while done < steps:
        t1 = time.time()
        one()
        times.append(time.time() - t1)
        done += 1
        if time.time() - t_first > budget_s:
            break
    t_step = sum(times) / len(times)
    tf_sample = 2 * TF_DECODER_LAYER + 2 * TF_VIT_LAYER + TF_STC
    clips_per_s = 1.0 / (t_step * TF_FULL / tf_sample)
    return {"value": round(clips_per_s, 6), "unit": "clips/s", "cores": cores, "kind": "port",
            "sample": "configs[0]: 2 ViT + 2 LLM layers at full width, 1 clip fwd+bwd+AdamW fp32, %d timed step(s) of %.2f s "
                      "(%.2f TF algorithmic); scaled to the 66.1 TF 7B step by FLOP ratio (EXTRAPOLATED)" % (len(times), t_step, tf_sample),
            "sample_seconds_per_step": round(t_step, 3), "steps_timed": len(times), "build_seconds": round(build_s, 1)}


def gpu_torch_reference(B, steps, warmup):
    """OPTIONAL context number (`--impl reference --ref-device cuda`): the oracle module itself -- Hugging Face CLIP / Mistral in
    bf16 with sdpa attention, the restated peft-style LoRA (two extra GEMMs + add per linear), torch clip_grad_norm_ + AdamW on
    fp32 adapter/head parameters under bf16 autocast -- stepped on THIS GPU on the same synthetic clips. It stands in for "the
    reference PyTorch/HF path on a B200" (SURVEY 8d-i); it is not the reference-arm number the driver compares (that is the CPU
    run) and nothing of it is on the product path."""
    import torch
    from oracle.ref_model import RefConfig, RefVLB, enable_training, synthetic_batch
    torch.cuda.set_device(0)
    torch.set_default_device("cuda")
    torch.set_default_dtype(torch.bfloat16)
    cfg = RefConfig(attn_impl="sdpa")
    ref = enable_training(RefVLB(cfg, seed=1234, lora_b_std=0.0))
    torch.set_default_dtype(torch.float32)
    params = []
    for _, p in ref.trainable():
        p.data = p.data.float()                    # fp32 master adapters / head, bf16 frozen backbone (bf16-mixed)
        params.append(p)
    opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-2)
    torch.set_default_device("cpu")
    batch = {k: v.cuda() for k, v in synthetic_batch(cfg, B, seed=1234).items()}
    torch.set_default_device("cuda")              # the module's glue (mask rows, splice padding) allocates with the default device

    def one():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = ref.step(batch)
        out["loss"].backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        return out["loss"]

    for _ in range(max(warmup, 1)):
        one()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = one()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"impl": "reference-torch-gpu", "metric": "lora_train_clips_per_sec", "value": round(B / ms * 1e3, 4), "unit": "clips/s",
            "n_gpus": 1, "steps": steps, "ms_per_step": round(ms, 2), "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "VideoLLaMA2-7B LoRA train step via the oracle module on the GPU: HF transformers %s CLIP/Mistral "
                                   "bf16 + sdpa attention + unfused peft-style LoRA + torch AdamW, %d clips/step" % (
                                       __import__("transformers").__version__, B)},
            "last_loss": float(loss), "max_mem_GB": round(torch.cuda.max_memory_allocated() / 1e9, 1)}


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return None
    if args.ref_device == "cuda":
        return gpu_torch_reference(min(args.batch, 6), args.steps, args.warmup)      # 19 GB per clip of autograd state: 6 clips fit 180 GB
    cb = cpu_reference(steps=args.steps, warmup=min(args.warmup, 1), budget_s=180.0)
    return {"impl": "reference", "metric": "lora_train_clips_per_sec", "value": cb["value"], "unit": "clips/s",
            "n_gpus": args.gpus, "steps": cb["steps_timed"], "warmup": min(args.warmup, 1),
            "ms_per_step": round(cb["sample_seconds_per_step"] * 1e3, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "VideoLLaMA2-7B LoRA train step, 8-frame 336^2 clips, 1000-parcel head (reference CPU path)"},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=8, help="clips per GPU per step (8 x 2048 tokens: ~95 GB of the 180 GB HBM)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-device", default="cpu", choices=["cpu", "cuda"], help="--impl reference only: cuda = time the oracle module on this GPU (context number)")
    ap.add_argument("--lora-dropout", type=float, default=0.0, help="peft lora_dropout (reference yaml: 0.1)")
    ap.add_argument("--head-dropout", type=float, default=0.0, help="dropout_rate of the brain head (reference yaml: 0.1)")
    args = ap.parse_args()
    res = run_reference(args) if args.impl == "reference" else run_ours(args)
    if res is not None:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
