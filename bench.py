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


def lit_config(P=1000, lora_dropout=0.1, dropout_rate=0.1):
    from phantom_vlb_b200.litmodule import VLBLitModuleConfig
    # config/experiment/VLB_vllama2_friends_lora.yaml:19-40 (lora_dropout 0.1, dropout_rate 0.1 are the reference's values)
    return VLBLitModuleConfig(model_path="random:7b", freeze_backbone=False, use_lora=True, lora_r=16, lora_alpha=32,
                              lora_dropout=lora_dropout, dropout_rate=dropout_rate, num_target=P, l2_lambda=0.001, lr=1e-4,
                              betas=[0.9, 0.999], eps=1e-8, weight_decay=1e-2, lr_scheduler_name="CosineAnnealingLR",
                              last_epoch=-1, t_max=50000)


NCU_SHAPE_ORDER = [("qkv", 16384, 6144, 4096), ("gu", 16384, 28672, 4096), ("d", 16384, 4096, 14336),
                   ("d_gu", 16384, 4096, 28672), ("d_d", 16384, 14336, 4096)]     # tools/ncu_gemm_shapes.py: two launches per shape


def _to_bytes(txt):
    v, unit = str(txt).split()[:2]
    return float(v) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


def worst_ncu_traffic():
    """DRAM read+write bytes per launch of the dominant kernel from the newest committed `ncu --set full` capture of the decoder
    GEMM shapes (profiles/*gemm2_shapes*.json, written by tools/ncu_summary.py from tools/ncu_gemm_shapes.py): the WORST shape,
    i.e. the largest measured / algorithmic ratio."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*gemm2_shapes*.json")))
    if not files:
        return None
    path = files[-1]
    try:
        rows = json.load(open(path))["launches"]
    except Exception:
        return None
    best = None
    for i, r in enumerate(rows):
        try:
            name, M, N, K = NCU_SHAPE_ORDER[i // 2]
            meas = _to_bytes(r["dram__bytes_read.sum"]) + _to_bytes(r["dram__bytes_write.sum"])
        except Exception:
            continue
        alg = 2.0 * (M * K + N * K + M * N)
        cand = dict(file=os.path.basename(path), shape=name, M=M, N=N, K=K, measured_bytes=meas, algorithmic_bytes=alg,
                    ratio=round(meas / alg, 2))
        if best is None or cand["ratio"] > best["ratio"]:
            best = cand
    return best


def run_ours(args):
    import torch
    import torch.distributed as dist
    from phantom_vlb_b200 import dp
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
    h2d = sum(v.numel() * v.element_size() for v in hosts[0].values())
    # data parallel: the flat gradient is all-reduced as dp.OverlappedGradReducer schedules it (VLB_DP_REDUCE: "deferred" = one
    # all-reduce after the backward, "inline" = per finished chunk on the compute stream, "overlap" = per chunk on a side stream)
    reducer = dp.OverlappedGradReducer(device=dev, timing=True) if world > 1 else None
    eng.grad_ready = reducer
    if world > 1:
        eng.row_gather, eng.bcast0 = dp.gather_rows, (lambda t: dist.broadcast(t, 0))

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
    marks = None

    def step_dev(i):
        if marks is not None:
            marks.append([torch.cuda.Event(enable_timing=True) for _ in range(3)])
            marks[-1][0].record()
        eng.loss_and_grads(devs[i % 2])
        if marks is not None:
            marks[-1][1].record()
        if reducer is not None:
            reducer.wait()
        if marks is not None:
            marks[-1][2].record()
        eng.optimizer_step(cosine_lr(cfg.lr, eng.T.step_count, cfg.t_max), world_size=world)

    for i in range(Wm):
        step_dev(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if reducer is not None:
        torch.cuda.synchronize(); reducer.pop_ms()
    l0 = eng.launches
    ms = timed(step_dev, K)
    launches = eng.launches - l0
    value = world * B * K / (ms / 1e3)
    step_ms = ms / K
    dp_info = None
    if world > 1:
        allreduce_ms = reducer.pop_ms() / K
        marks = []
        barrier()
        for i in range(3):
            step_dev(i)
        torch.cuda.synchronize()
        compute = sum(m[0].elapsed_time(m[1]) for m in marks) / len(marks)
        exposed = sum(m[1].elapsed_time(m[2]) for m in marks) / len(marks)
        marks = None
        reducer.pop_ms()
        t = torch.tensor([compute, exposed, allreduce_ms], device=dev)
        allv = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        allv = torch.stack(allv).cpu()
        dp_info = {"allreduce_ms": round(float(allv[:, 2].max()), 3), "allreduce_exposed_ms": round(float(allv[:, 1].max()), 3),
                   "rank_skew_ms": round(float(allv[:, 0].max() - allv[:, 0].min()), 3),
                   "rank_compute_ms": [round(float(x), 2) for x in allv[:, 0]], "allreduce_bytes": 4 * eng.T.numel,
                   "schedule": reducer.mode,
                   "how": "allreduce_ms = sum of the ncclAllReduce durations per step (CUDA events on the stream they run on, max "
                          "over ranks); exposed = backward end -> gradients reduced, on the compute stream; rank_skew = max - "
                          "min over ranks of the step-start -> backward-end device time (3 extra steps)"}

    # ---- end-to-end arm through the public API with pinned host buffers (e2e)
    (opt,), (sch,) = module.configure_optimizers()
    opt.world_size, opt.prereduced = world, world > 1
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

    # ---- roofline of the dominant kernel (the CTA-pair tcgen05 GEMM with the fused LoRA term): one instrumented step with CUDA
    # events around every GEMM launch. Per-launch events leave idle gaps in which the power-capped clock recovers, so the sum of
    # the per-launch times is NOT what those launches cost inside the plain step; the kernel's SHARE of the instrumented step is
    # applied to the plain step time instead.
    eng.profile = []
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ea.record()
    step_dev(0)
    eb.record()
    torch.cuda.synchronize()
    prof, eng.profile = eng.profile, None
    t_instr = ea.elapsed_time(eb)
    g_ms = sum(e0.elapsed_time(e1) for e0, e1, _, _ in prof)
    g_fl = sum(f for _, _, f, _ in prof)
    big = [(e0.elapsed_time(e1), f) for e0, e1, f, s in prof if s[1] >= 1024 and s[0] >= 1024]
    big_ms, big_fl = sum(t for t, _ in big), sum(f for _, f in big)
    peak_tf, peak_bw, peak_src = peaks()
    share = big_ms / t_instr
    achieved = big_fl / (share * step_ms / 1e3) / 1e12
    traffic = worst_ncu_traffic()
    roofline = {"bound": "tensor", "kernel": "gemm2_tcgen05_kernel (cta_group::2 pair tiles 256x256; fused LoRA k-block / masked adapter epilogue / SwiGLU epilogues): decoder, ViT and STC GEMMs",
                "achieved": round(achieved, 1), "peak": peak_tf, "unit": "TFLOP/s", "frac": round(achieved / peak_tf, 4),
                "traffic": traffic["measured_bytes"] if traffic else None, "traffic_note": traffic,
                "peak_source": peak_src, "launches_timed": len(big), "share_of_step": round(share, 4),
                "achieved_per_launch_events": round(big_fl / (big_ms / 1e3) / 1e12, 1),
                "how": "algorithmic 2*M*N*(K+K2) over the launches with M,N >= 1024 / (their share of one event-instrumented step x the "
                       "plain step time); achieved_per_launch_events is the raw sum over per-launch event times (clock recovers in the gaps)",
                "all_gemm_tflops": round(g_fl / (g_ms / 1e3) / 1e12, 1), "step_tflops": round(B * TF_FULL / (step_ms / 1e3), 1)}
    res = None
    if rank == 0:
        res = {"metric": "lora_train_clips_per_sec", "value": round(value, 4), "unit": "clips/s", "n_gpus": world,
               "steps": K, "warmup": Wm, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
               "config": {"workload": "VideoLLaMA2-7B (CLIP ViT-L/14-336 23L + STC + Mistral-7B 32L) LoRA r=16 bf16, "
                                      "8-frame 336^2 clips, 1000-parcel head, S=2048 (BASELINE configs[%d]); lora_dropout %g / dropout_rate %g "
                                      "as in the reference's VLB_vllama2_friends_lora.yaml" % (1 if world == 1 else 2, args.lora_dropout, args.head_dropout),
                          "clips_per_gpu": B, "global_batch": B * world, "lora_dropout": args.lora_dropout, "head_dropout": args.head_dropout, "parallelism": "dp%d" % world,
                          "weights": "random-init, frozen bf16; LoRA A kaiming / B zero; fp32 master + Adam state",
                          "l2_policy": "per-step working set (>%d GB activations) far exceeds the 126 MB L2; two alternating input batches" % (B * 9 // 4)},
               "e2e": {"value": round(e2e_value, 4), "unit": "clips/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                       "ms_per_step": round(ms_e2e / K, 3), "last_loss": losses[-1]},
               "gpu_launches": launches, "clocks": clocks, "roofline": roofline}
        if dp_info is not None:
            res["data_parallel"] = dp_info
    # ---- BASELINE configs[3] / configs[4] on the same N GPUs (extra keys; the driver only runs this file)
    if world > 1 and not args.no_extra_configs:
        import gc
        W = eng.W
        module.engine = None
        eng._bufs.clear()
        del eng, module, devs
        gc.collect(); torch.cuda.empty_cache()
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import run_dp_configs as rdc
        extra = []
        collect = lambda s, flush=True: extra.append(json.loads(s))
        try:
            rdc.measure_c5(rank, world, dev, 22528, 250000, out=collect)
            gc.collect(); torch.cuda.empty_cache()
            rdc.measure_c4(rank, world, dev, 4, 250000, W=W, out=collect)
        except Exception as e:                      # never lose the headline line over an extra measurement
            extra.append({"error": repr(e)[:300]})
        if rank == 0:
            res["configs_3_4"] = extra
    elif world == 1 and rank == 0:
        if not args.no_cpu_baseline:
            res["cpu_baseline"] = cpu_reference(steps=1, warmup=0, budget_s=60.0)
        if not args.no_gpu_reference:
            import gc
            module.engine = None
            eng._bufs.clear()
            del eng, module, devs
            gc.collect(); torch.cuda.empty_cache()
            res["gpu_reference"] = gpu_reference_subprocess()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return res


def gpu_reference_subprocess():
    """The 'real bar' of SURVEY 8d(i): the reference's PyTorch / Hugging Face path on THIS GPU (oracle module on CUDA: HF CLIP +
    Mistral bf16 with flash_attention_2 as litmodule.py:71-72 configures, unfused peft-style LoRA, torch clip + AdamW under bf16
    autocast), at the yaml's batch_size 3 and at 6 clips. Separate processes so its allocator state never meets ours."""
    out = {}
    for b in (3, 6):
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--ref-device", "cuda", "--batch", str(b),
                                "--steps", "3", "--warmup", "1"], capture_output=True, text=True, timeout=420)
            line = [l for l in r.stdout.splitlines() if l.startswith("{")]
            out["B%d" % b] = json.loads(line[-1]) if line else {"error": (r.stderr or r.stdout)[-300:]}
        except Exception as e:
            out["B%d" % b] = {"error": repr(e)[:300]}
    return out


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
    return {"value": round(clips_per_s, 6), "unit": "clips/s", "cores": cores, "kind": "port", "extrapolated": True,
            "sample": "configs[0]: 2 ViT + 2 LLM layers at full width, 1 clip fwd+bwd+AdamW fp32, %d timed step(s) of %.2f s "
                      "(%.2f TF algorithmic); scaled to the 66.1 TF 7B step by FLOP ratio (EXTRAPOLATED)" % (len(times), t_step, tf_sample),
            "sample_seconds_per_step": round(t_step, 3), "steps_timed": len(times), "build_seconds": round(build_s, 1)}


def gpu_torch_reference(B, steps, warmup):
    """Context number (`--impl reference --ref-device cuda`, and the `gpu_reference` key of the N=1 line): the oracle module itself
    -- Hugging Face CLIP / Mistral in bf16 with flash_attention_2 (what the reference configures: litmodule.py:71-72, encoder.py:24;
    sdpa if flash-attn cannot run here), the restated peft-style LoRA (two extra GEMMs + dropout + add per linear), torch
    clip_grad_norm_ + AdamW on fp32 adapter/head parameters under bf16 autocast -- stepped on THIS GPU on the same synthetic clips.
    It stands in for "the reference PyTorch/HF path on a B200" (SURVEY 8d-i); it is not the reference-arm number the driver
    compares (that is the CPU run) and nothing of it is on the product path."""
    import torch
    from oracle.ref_model import RefConfig, RefVLB, enable_training, synthetic_batch
    torch.cuda.set_device(0)

    def build(attn):
        torch.set_default_device("cuda")
        torch.set_default_dtype(torch.bfloat16)
        cfg = RefConfig(attn_impl=attn)
        ref = enable_training(RefVLB(cfg, seed=1234, lora_b_std=0.0, lora_dropout=0.1, head_dropout=0.1))
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
        return one

    attn = "flash_attention_2"
    try:
        import flash_attn  # noqa: F401
        one = build(attn)
        one()
        torch.cuda.synchronize()
    except Exception as e:
        sys.stderr.write("flash_attention_2 unavailable for the reference module (%s); using sdpa\n" % (repr(e)[:200],))
        attn = "sdpa"
        import gc
        one = None
        gc.collect(); torch.cuda.empty_cache()
        one = build(attn)
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
            "n_gpus": 1, "steps": steps, "ms_per_step": round(ms, 2), "dtype": "bf16", "data": "synthetic", "attention": attn,
            "config": {"workload": "VideoLLaMA2-7B LoRA train step via the oracle module on the GPU: HF transformers %s CLIP/Mistral "
                                   "bf16 + %s attention + unfused peft-style LoRA (lora_dropout 0.1, dropout_rate 0.1) + torch "
                                   "AdamW, %d clips/step" % (__import__("transformers").__version__, attn, B)},
            "last_loss": float(loss), "max_mem_GB": round(torch.cuda.max_memory_allocated() / 1e9, 1)}


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return None
    if args.ref_device == "cuda":
        return gpu_torch_reference(min(args.batch, 6), args.steps, args.warmup)      # 19 GB per clip of autograd state: 6 clips fit 180 GB
    cb = cpu_reference(steps=args.steps, warmup=min(args.warmup, 1), budget_s=180.0)
    return {"impl": "reference", "extrapolated": True, "metric": "lora_train_clips_per_sec", "value": cb["value"], "unit": "clips/s",
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
    ap.add_argument("--lora-dropout", type=float, default=0.1, help="peft lora_dropout (reference yaml: 0.1)")
    ap.add_argument("--head-dropout", type=float, default=0.1, help="dropout_rate of the brain head (reference yaml: 0.1)")
    ap.add_argument("--no-gpu-reference", action="store_true", help="N=1: skip timing the oracle module on this GPU (gpu_reference key)")
    ap.add_argument("--no-extra-configs", action="store_true", help="N>1: skip the configs[3]/configs[4] measurements (configs_3_4 key)")
    args = ap.parse_args()
    res = run_reference(args) if args.impl == "reference" else run_ours(args)
    if res is not None:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
