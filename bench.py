#!/usr/bin/env python
"""bench.py -- the judged measurement of the hot path (see the contract in DESIGN.md §Measurement).

Workload ("cfg3-hotpath"): BASELINE config 3 restricted to SURVEY.md §8's hot path.  One *step* =
one denoising step's worth of hot-path work for one 16-frame 1024^2 video: two UNet evaluations
(uncond + cond, as inference_animatediff.py:110-121 runs CFG), each = 70 spatial self-attention +
70 text cross-attention processor calls (UnZipLoRA rank 64 content+style, fused projections) and
15 motion modules (GroupNorm5D, rearranges, temporal-LoRA rank 32 projections, frame attention,
GEGLU FF).  Synthetic latents, random-init weights (1.2 B parameters, distinct per module).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  N>1: launched by torchrun, one rank per GPU; ranks are data-parallel over videos (weak scaling,
  no data-path collective -- DESIGN.md §Multi-GPU).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "16f 1024^2 SDXL-AnimateDiff denoise steps/s (hot path: UnZipLoRA attention + motion modules, x2 CFG)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--latent", type=int, default=128)
    ap.add_argument("--depth-scale", type=float, default=1.0, help="<1 shrinks the stack (debug only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true",
                    help="launch every kernel from Python instead of replaying the captured CUDA graph")
    ap.add_argument("--workload", default="cfg3", choices=["cfg3", "cfg2", "cfg4", "cfg5"],
                    help="cfg3 = the judged headline line; the others are the secondary BASELINE "
                         "configs (bench_extra.py)")
    ap.add_argument("--cfg-mode", default="batched", choices=["batched", "two-pass"],
                    help="batched: uncond|cond as one B=2 pass (default); two-pass: the reference's two B=1 calls")
    ap.add_argument("--no-sp", action="store_true", help="skip the sequence-parallel probe at N>1")
    ap.add_argument("--with-spatial-ff", action="store_true",
                    help="secondary line: every spatial transformer block in full (LayerNorms + GEGLU feed-forward "
                         "around the two attentions, SURVEY.md §8f-1) instead of the attention processors alone")
    ap.add_argument("--ncu-step", action="store_true",
                    help="warm up, then run ONE step between cudaProfilerStart/Stop and exit "
                         "(for `ncu --profile-from-start off`; prints no bench line)")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# --------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# --------------------------------------------------------------------------------------------------
# workload description shared by both arms (the driver compares the two `config` objects)
# --------------------------------------------------------------------------------------------------
def common_config(frames, latent, world):
    return {"workload": "cfg3-hotpath", "frames": frames, "latent": latent, "cfg_passes": 2,
            "modules_per_unet_call": {"self_attn": 70, "cross_attn": 70, "motion": 15},
            "unziplora_rank": 64, "temporal_lora_rank": 32, "parallelism": f"dp{world} (videos)",
            "l2": "per-step working set (2.4 GB of weights + >0.4 GB activations) exceeds the 126 MB L2"}


# --------------------------------------------------------------------------------------------------
# CPU arm: the reference's own modules (oracle/_ref) on the host cores, bounded sample
# --------------------------------------------------------------------------------------------------
def cpu_arm_line(arm, walls, med, best):
    sec = arm.step_seconds(med)
    return {"value": 1.0 / sec, "unit": "steps/s", "cores": arm.cores, "kind": arm.kind,
            "sample": arm.sample_text,
            "extrapolation": {"kind": "module_sum", "cfg_passes": 2,
                              "calls_x_downbatch_factor": {k: round(v, 3) for k, v in arm.factors().items()},
                              "unit_seconds_median": {k: round(v, 4) for k, v in med.items()},
                              "unit_seconds_best": {k: round(v, 4) for k, v in best.items()},
                              "step_seconds_median": round(sec, 2),
                              "step_seconds_best": round(arm.step_seconds(best), 2)},
            "timed_passes": len(walls), "seconds_per_pass": round(sum(walls) / len(walls), 3)}


def run_reference(args):
    """`--impl reference`: K timed passes (after W untimed ones) over the bounded sample of the cfg3
    hot-path step through the reference's own modules on all host threads.  `ms_per_step` is what one
    timed pass took (steps x ms_per_step is the timed region); `value` is the full step's rate by the
    module-sum extrapolation spelled out under cpu_baseline.extrapolation."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.ref_cpu import CpuArm
    arm = CpuArm(args.frames, args.latent)
    walls, med, best = arm.run(max(1, args.steps), max(0, args.warmup))
    cb = cpu_arm_line(arm, walls, med, best)
    val = cb["value"]
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sum(walls) / len(walls), "ms_per_full_step_extrapolated": 1e3 / val,
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": common_config(args.frames, args.latent, args.gpus),
            "cpu_baseline": cb,
            "e2e": {"value": val, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# per-kernel event timing of one eager step (roofline of the dominant kernel)
# --------------------------------------------------------------------------------------------------
class KernelTimer:
    """Wraps ops.lora_gemm_raw / ops.attn_fwd_raw with CUDA events on the launching stream and groups
    the launches by the kernel instantiation they run on (vst_lora_gemm_plan)."""

    def __init__(self, ops, lib):
        self.ops, self.lib, self.rows = ops, lib, []
        self._gemm, self._attn = ops.lora_gemm_raw, ops.attn_fwd_raw

    def __enter__(self):
        ops, lib = self.ops, self.lib

        def gemm(a1, b1, a2=None, b2=None, bias=None, residual=None, out_scale=1.0, act="none",
                 k2_group_n=0, out=None, out_dtype=torch.bfloat16):
            m, k1 = a1.shape
            n = b1.shape[0]
            k2 = 0 if b2 is None else b2.shape[1]
            nn = n // 2 if act == "geglu" else n
            code = lib.vst_lora_gemm_plan(m, nn, k2, k2_group_n if a2 is not None else 0,
                                          {"none": 0, "gelu": 1, "geglu": 2}[act])
            name = ("lora_gemm2_kernel" if code >= 1000 else "lora_gemm_kernel") + \
                f"<{code % 1000},{1 if (residual is not None and out_dtype == torch.bfloat16) else 0}>" + \
                (" geglu" if act == "geglu" else "")
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = self._gemm(a1, b1, a2, b2, bias, residual, out_scale, act, k2_group_n, out, out_dtype)
            e.record()
            byts = 2.0 * (m * k1 + n * k1 + m * nn * (2 if residual is not None else 1) + (m + n) * k2)
            self.rows.append((name, s, e, 2.0 * m * n * (k1 + k2), byts))
            return r

        def attn(q, k, v, heads, scale=None, return_lse=False, mask=None):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = self._attn(q, k, v, heads, scale, return_lse, mask)
            e.record()
            name = "attn_fwd_pers_kernel " + ("self" if k.shape[1] == q.shape[1] else "text-cross")
            fl = 4.0 * q.shape[0] * q.shape[1] * k.shape[1] * q.shape[2]
            byts = 2.0 * (2 * q.numel() + k.shape[0] * k.shape[1] * q.shape[2] * 2)
            self.rows.append((name, s, e, fl, byts))
            return r

        ops.lora_gemm_raw, ops.attn_fwd_raw = gemm, attn
        return self

    def __exit__(self, *a):
        self.ops.lora_gemm_raw, self.ops.attn_fwd_raw = self._gemm, self._attn

    def table(self):
        torch.cuda.synchronize()
        out = {}
        for name, s, e, fl, by in self.rows:
            t = out.setdefault(name, {"ms": 0.0, "flop": 0.0, "bytes": 0.0, "launches": 0})
            t["ms"] += s.elapsed_time(e)
            t["flop"] += fl
            t["bytes"] += by
            t["launches"] += 1
        return out


# --------------------------------------------------------------------------------------------------
# sequence-parallel probe (WORLD_SIZE > 1): BASELINE config 5 on the driver's clock
# --------------------------------------------------------------------------------------------------
def sp_probe(args, world, rank, dev, steps):
    """One 32-frame video sharded over the ranks (frames for the spatial processors, positions for
    the motion-module bodies; the two all-to-alls per motion module are the layout kernels writing
    to / reading from peer memory over NVLink).  Returns the `sp` object of the bench line."""
    import torch.distributed as dist
    from video_style_transfer_b200 import sdxl_hotpath as hpth, sequence_parallel as spm
    from video_style_transfer_b200 import fused as fusedmod
    bf = torch.bfloat16
    F, S = 32, args.latent
    fl_ = F // world
    torch.manual_seed(4321)                      # every rank holds the same weights: ONE video
    with torch.device(dev):
        model = hpth.SDXLHotPath(temporal_rank=32, depth_scale=args.depth_scale)
    model = model.to(bf).eval()
    plan = model.forward_plan()
    nb = 2                                       # batched CFG: uncond | cond stacked along the batch
    full_shapes = hpth.level_shapes(F, S, batch=nb)
    g = torch.Generator().manual_seed(777)       # the same full video on every rank; each keeps its frames
    full4 = [torch.randn(s, generator=g).to(bf) for s in full_shapes]
    full3 = [None] + [torch.randn(s[0], s[2] * s[3], s[1], generator=g).to(bf) for s in full_shapes[1:]]
    texts = [torch.randn(nb, 77, 2048, generator=g).to(bf).to(dev)]
    sl = slice(rank * fl_, (rank + 1) * fl_)

    def mine(t):                                 # (nb * F, ...) -> this rank's frames of both CFG halves
        return t.view(nb, F, *t.shape[1:])[:, sl].reshape(nb * fl_, *t.shape[1:]).contiguous()

    xs4 = [mine(t).to(dev) for t in full4]
    xs3 = [None] + [mine(t).to(dev) for t in full3[1:]]
    nbytes = max(spm.SPMotionStep.buffer_bytes(nb, s[1], s[2] * s[3], F, world) for s in full_shapes)
    comm = spm.SPComm.from_process_group(nbytes)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() / n

    # ---- parity: this rank's slice of the unsharded modules' outputs (same kernels, one GPU) ------
    with torch.no_grad():
        want = hpth.run_unet_hotpath(plan, [t.to(dev) for t in full4], [None] + [t.to(dev) for t in full3[1:]],
                                     texts[0], F)
        got = hpth.run_unet_hotpath_sp(plan, xs4, xs3, texts[0], F, comm)
    errs = []
    for w_, g_ in zip(want, got):
        w_ = mine(w_).float()
        errs.append(((g_.float() - w_).abs().max() / w_.abs().max().clamp_min(1e-30)).item())
    perr = torch.tensor([max(errs)], device=dev)
    dist.all_reduce(perr, op=dist.ReduceOp.MAX)
    del want, got

    # ---- the sharded step, graph replay, max over ranks -------------------------------------------
    graphed = hpth.GraphedSPDenoiseStep(plan, xs4, xs3, texts, F, comm)
    for _ in range(3):
        graphed.replay()
    with ClockSampler(dev.index) as clk_sp:               # rank 0's SM clock under the sharded step ...
        sp_steps = max(steps, 20 * world)                 # ~5 s of timed region at every N: nvidia-smi answers in
                                                          # 0.3-1 s with N ranks polling, 20 steps at N = 8 gave ONE sample
        ms_sp = timed(graphed.replay, sp_steps)

    # ---- per-phase device time of one eager step (events on the launching stream) -----------------
    phases = {k: [] for k in ("gn_partials_share", "push", "normalize", "compute", "pull", "barrier_wait", "spatial")}

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def motion_sp_timed(module, x_local):
        st = spm.SPMotionStep(module, comm, x_local, F)
        t = [ev()]
        st.share_partials(); t.append(ev())
        st.push_raw(); t.append(ev())
        comm.barrier(); t.append(ev())
        st.normalize(); t.append(ev())
        st.compute(); t.append(ev())
        comm.barrier(); t.append(ev())
        out = st.pull(); t.append(ev())
        for name, i in (("gn_partials_share", 0), ("push", 1), ("normalize", 3), ("compute", 4), ("pull", 6)):
            phases[name].append((t[i], t[i + 1]))
        for i in (2, 5):
            phases["barrier_wait"].append((t[i], t[i + 1]))
        return out

    barrier()
    with torch.no_grad():
        for text in texts:
            t_prev = ev()
            for kind, lvl, mod in plan:
                if kind == "motion":
                    phases["spatial"].append((t_prev, ev()))
                    motion_sp_timed(mod, xs4[lvl])
                    t_prev = ev()
                elif kind == "self":
                    mod(xs3[lvl])
                else:
                    mod(xs3[lvl], encoder_hidden_states=text)
            phases["spatial"].append((t_prev, ev()))
    torch.cuda.synchronize()
    ph = torch.tensor([sum(a.elapsed_time(b) for a, b in phases[k]) for k in phases], device=dev)
    dist.all_reduce(ph, op=dist.ReduceOp.MAX)
    phase_ms = {k: round(v, 3) for k, v in zip(phases, ph.tolist())}

    # ---- the same video on ONE GPU (rank 0 alone, same process, same clocks window) --------------
    ms_sp1 = torch.zeros(1, device=dev)
    barrier()
    if rank == 0:
        g1 = hpth.GraphedDenoiseStep(plan, [t.to(dev) for t in full4], [None] + [t.to(dev) for t in full3[1:]],
                                     texts, F)
        for _ in range(3):
            g1.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(dev.index) as clk_sp1:          # ... and under the same video on this GPU alone
            e0.record()
            for _ in range(steps):
                g1.replay()
            e1.record()
            torch.cuda.synchronize()
        ms_sp1[0] = e0.elapsed_time(e1) / steps
        del g1
    dist.all_reduce(ms_sp1, op=dist.ReduceOp.MAX)
    ms_sp1 = ms_sp1.item()
    egress = 0
    for kind, lvl, _ in plan:
        if kind == "motion":
            s = full_shapes[lvl]
            egress += 2 * (s[0] * s[1] * s[2] * s[3] * 2 // world) * (world - 1) // world   # push + pull
    # (full_shapes already carry both CFG halves)
    barrier()
    del graphed
    comm.close()
    return {"workload": "cfg5-sequence-parallel: ONE 32-frame 1024^2 video, hot path, x2 CFG (batched B=2)",
            "n_gpus": world, "frames": F, "frames_per_rank": fl_, "steps": sp_steps, "steps_sp1": steps,
            "ms_per_step": ms_sp, "ms_sp1": ms_sp1, "speedup": ms_sp1 / ms_sp,
            "efficiency": ms_sp1 / (world * ms_sp), "scaling": "strong",
            "clocks": {"sp": clk_sp.summary(), "sp1": clk_sp1.summary() if rank == 0 else None,
                       "note": "rank 0's SM clock sampled during each timed region; efficiency_iso_clock rescales "
                               "ms_sp1 to the sharded run's median clock (compute-bound step)"},
            "efficiency_iso_clock": (None if rank != 0 or not clk_sp.summary()["sm_mhz"] or not clk_sp1.summary()["sm_mhz"]
                                     else (ms_sp1 * clk_sp1.summary()["sm_mhz"] / clk_sp.summary()["sm_mhz"]) / (world * ms_sp)),
            "nvlink_bytes_per_rank": egress, "parity_rel_err": perr.item(),
            "parity_note": "rank-sliced output of the sharded step vs the unsharded modules on one GPU, "
                           "max |a-b| / max |b| over the per-level outputs, max over ranks",
            "phase_ms": phase_ms,
            "phase_note": "one eager step, CUDA events on the launching stream, summed over the 30 motion-module "
                          "calls, max over ranks; barrier_wait = 2 flag barriers per motion module (includes peer "
                          "skew and, in eager mode, launch skew); the timed ms_per_step is the graph replay",
            "launch": "CUDA graph replay (one graph per rank, no NCCL and no host sync on the data path)",
            "adapters": "folded" if fusedmod.MERGE_LORA else "low-rank"}


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist
    from video_style_transfer_b200 import _lib, ops, sdxl_hotpath as hpth
    from video_style_transfer_b200 import fused as fusedmod

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    bf = torch.bfloat16
    F, S = args.frames, args.latent
    batched = args.cfg_mode == "batched"
    nb = 2 if batched else 1                                   # CFG halves stacked along the batch

    torch.manual_seed(1234 + rank)
    with torch.device(dev):
        model = hpth.SDXLHotPath(temporal_rank=32, depth_scale=args.depth_scale, spatial_ff=args.with_spatial_ff)
    model = model.to(bf).eval()
    plan = model.forward_plan()
    shapes = hpth.level_shapes(F, S, batch=nb)

    # host (pinned) copies of one step's inputs, and device-resident copies for `value`
    g = torch.Generator().manual_seed(rank)
    h_x4 = [torch.randn(s, generator=g).to(bf).pin_memory() for s in shapes]
    h_x3 = [None] + [torch.randn(s[0], s[2] * s[3], s[1], generator=g).to(bf).pin_memory() for s in shapes[1:]]
    h_text = [torch.randn(1, 77, 2048, generator=g).to(bf) for _ in range(2)]     # uncond, cond: per PROMPT
    d_x4 = [t.to(dev) for t in h_x4]
    d_x3 = [None] + [t.to(dev) for t in h_x3[1:]]
    d_text = [torch.cat(h_text).to(dev)] if batched else [t.to(dev) for t in h_text]
    h_out = [torch.empty(s, dtype=bf).pin_memory() for s in shapes]
    h2d = sum(t.numel() * 2 for t in h_x4 + h_x3[1:])
    d2h = sum(t.numel() * 2 for t in h_out)

    def step_eager():
        outs = None
        for text in d_text:           # two-pass: uncond, cond B=1 calls; batched: one B=2 call
            outs = hpth.run_unet_hotpath(plan, d_x4, d_x3, text, F)
        return outs

    # the product's step: the same launches captured once in a CUDA graph and replayed
    graphed = None if (args.no_graph or args.ncu_step) else hpth.GraphedDenoiseStep(plan, d_x4, d_x3, d_text, F)
    step_device = step_eager if graphed is None else graphed.replay
    streamed = None if graphed is None else hpth.StreamedDenoiseStep(graphed)
    used_graph = graphed is not None

    def step_e2e():
        if graphed is None:
            x4 = [t.to(dev, non_blocking=True) for t in h_x4]
            x3 = [None] + [t.to(dev, non_blocking=True) for t in h_x3[1:]]
            outs = None
            for text in d_text:
                outs = hpth.run_unet_hotpath(plan, x4, x3, text, F)
        else:
            # the product's host-fed path: copies of step n+1 / n-1 overlap the replay of step n
            streamed.submit(h_x4 + h_x3[1:], h_out)
            return
        for lvl in range(3):                                   # the step's result back on the host
            o = outs[lvl]
            if o.dim() == 3:                                   # last call at L1/L2 may be a token tensor
                h_out[lvl].view(o.shape).copy_(o, non_blocking=True)
            else:
                h_out[lvl].copy_(o, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, after=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        if after is not None:
            after()                                            # e.g. wait for the last D2H copy
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    for _ in range(max(args.warmup, 3)):
        step_device()
    if os.environ.get("VST_KERNEL_TABLE"):
        # per-kernel device time of ONE step through torch.profiler (CUPTI): cheap next to an ncu pass; tab-separated
        from torch.profiler import ProfilerActivity, profile
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step_device()
            torch.cuda.synchronize()
        rows = sorted(((e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0),
                      key=lambda r: -r[2])
        tot = sum(r[2] for r in rows)
        with open(os.environ["VST_KERNEL_TABLE"], "w") as f:
            f.write(f"# one hot-path denoise step, kernel device time (us), total {tot:.1f} us over {sum(r[1] for r in rows)} launches\n")
            f.write("kernel\tlaunches\ttotal_us\tshare\n")
            for k, n, t in rows:
                f.write(f"{k[:120]}\t{n}\t{t:.1f}\t{t / tot:.4f}\n")
    if args.ncu_step:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step_device()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    launches0 = _lib.launch_count
    with ClockSampler(local) as clk:
        ms_dev = timed(step_device, args.steps)
    launches = (_lib.launch_count - launches0) if graphed is None else graphed.launches * args.steps
    step_e2e()
    ms_e2e = timed(step_e2e, args.steps, after=None if streamed is None else streamed.drain)

    # ---- rooflines: one extra eager step with CUDA events around every GEMM / attention launch ----
    pk, pk_src = peaks()
    step_eager()                      # allocator / text K/V warm-up on this stream (untimed)
    torch.cuda.synchronize()
    if graphed is not None:           # keep the GPU busy and power-capped while the host runs ahead:
        graphed.replay()              # the eager launches below queue up behind two replays, so every
        graphed.replay()              # event pair brackets a kernel, not a host launch gap
    with KernelTimer(ops, _lib.lib) as kt:
        step_eager()
    table = kt.table()
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    traffic_db = {}
    try:
        traffic_db = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        pass
    ms_step = ms_dev / args.steps

    def roof(name):
        t = table[name]
        ach = t["flop"] / (t["ms"] * 1e-3) / 1e12
        tr = traffic_db.get(name.split(" ")[0])
        return {"kernel": name, "bound": "tensor", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": ach / peak_tf,
                "traffic": None if tr is None else tr["dram_bytes_per_launch"],
                "traffic_note": None if tr is None else tr["note"],
                "algorithmic_bytes_per_launch": t["bytes"] / t["launches"],
                "algorithmic_flop_per_launch": t["flop"] / t["launches"],
                "us_per_launch": 1e3 * t["ms"] / t["launches"],
                "peak_source": f"bf16_tflops_sustained, {pk_src}",
                "launches_timed": t["launches"], "share_of_step": t["ms"] / ms_step,
                "timing": "CUDA events around every launch of one eager step queued behind two graph replays (GPU never "
                          "waits for the host; adds ~2 us of event overhead per launch)"}

    gemm_names = [n for n in table if n.startswith("lora_gemm")]
    dominant = max(gemm_names, key=lambda n: table[n]["ms"])
    by_kernel = {n: {"ms": round(t["ms"], 3), "launches": t["launches"],
                     "tflops": round(t["flop"] / (t["ms"] * 1e-3) / 1e12, 1),
                     "frac": round(t["flop"] / (t["ms"] * 1e-3) / 1e12 / peak_tf, 3)}
                 for n, t in sorted(table.items(), key=lambda kv: -kv[1]["ms"])}
    fl = hpth.hotpath_flops(plan, F, S, merged=fusedmod.MERGE_LORA, text_kv_cached=True)   # FLOPs executed
    step_flops = 2.0 * sum(fl.values())
    step_flops_lowrank = 2.0 * sum(hpth.hotpath_flops(plan, F, S).values())  # SURVEY §8d low-rank form

    sp = None
    if world > 1 and not args.no_sp and 32 % world == 0:
        if graphed is not None:
            del streamed, graphed
            streamed = graphed = None
        torch.cuda.empty_cache()
        sp = sp_probe(args, world, rank, dev, max(3, min(args.steps, 10)))

    if rank != 0:
        return
    cfg = common_config(F, S, world)
    if args.with_spatial_ff:
        cfg["workload"] = "cfg3-hotpath+spatial-ff"
        cfg["modules_per_unet_call"] = {"spatial_transformer_block (LN, self-attn, LN, cross-attn, LN, GEGLU FF)": 70,
                                        "motion": 15}
    line = {
        "metric": METRIC, "value": world * 1e3 / ms_step, "unit": "steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic", "config": cfg,
        "execution": {"cfg": ("batched: uncond | cond stacked along the batch, ONE B=2 pass per step "
                              "(pipeline_animatediff_xl.py:353,373-375)" if batched else
                              "two B=1 passes per step (inference_animatediff.py:110-121)"),
                      "launch": "CUDA graph replay" if used_graph else "eager (one Python call per kernel)",
                      "text_kv": "to_k | to_v of the text computed once per prompt (constant over the denoise loop), "
                                 "outside the step",
                      "adapters": ("folded into the weights once per parameter version (W + B'A), one GEMM per "
                                   "projection" if fusedmod.MERGE_LORA else
                                   "low-rank form: u = x A^T pre-pass + extra K blocks in the fused GEMM"),
                      "step_tflop": step_flops / 1e12, "step_tflop_lowrank_form": step_flops_lowrank / 1e12},
        "e2e": {"value": world * 1e3 / (ms_e2e / args.steps), "unit": "steps/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        "clocks": clk.summary(),
        "tflops_hotpath": step_flops / (ms_step * 1e-3) / 1e12,
        "frac_step": step_flops / (ms_step * 1e-3) / 1e12 / peak_tf,
        "roofline": roof(dominant),
        "roofline_attn": roof("attn_fwd_pers_kernel self"),
        "by_kernel": by_kernel,
    }
    if sp is not None:
        line["sp"] = sp
    if world == 1 and not args.no_cpu_baseline:
        from oracle.ref_cpu import CpuArm
        arm = CpuArm(F, S)
        walls, med, best = arm.run(3, 1)
        line["cpu_baseline"] = cpu_arm_line(arm, walls, med, best)
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload != "cfg3":
        import bench_extra
        bench_extra.run(args, {"ClockSampler": ClockSampler, "peaks": peaks})
    else:
        run_ours(args)
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
