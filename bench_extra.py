"""Secondary workloads of bench.py (`--workload cfg2|cfg4|cfg5`): the BASELINE.json configs that are
not the headline line.  Same timing rules as bench.py (CUDA events on the launching stream, >= 3
warm-up steps, barrier + synchronize on both sides, max over ranks); each prints one JSON line.

  cfg2  UnZipLoRA attention processor fwd+bwd at the SDXL mid-block geometry (C=1280, 20 heads,
        16 x 1024 tokens, text 1 x 77 x 2048, rank-64 content+style), self and cross, Stage-2
        (mergers + dX) and Stage-1 (A, B, mergers + dX) trainable sets.
  cfg4  one Stage-2 training step of the hot-path stack (fwd + bwd + clip + AdamW), 16 frames; every
        module output receives a synthetic upstream gradient (the rest of the UNet is out of scope).
  cfg5  one denoising step of ONE 32-frame video, sequence-parallel over --gpus ranks
        (frames <-> positions all-to-alls fused into the layout kernels, peer memory over NVLink).
"""
import json
import os
import time
import types

import torch


def _timed(fn, steps, world, dev):
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item() / steps


# --------------------------------------------------------------------------------------------------
# cfg2
# --------------------------------------------------------------------------------------------------
def _midblock_attn(kind, stage, dev):
    from video_style_transfer_b200.unziplora_unet.lora_linear import LoRACompatibleLinear
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    C, heads, ctx, r = 1280, 20, 2048, 64
    kv = C if kind == "self" else ctx

    def mk(i, o, bias):
        lin = LoRACompatibleLinear(i, o, bias=bias)
        lay = UnZipLoRALinearLayer(i, o, r, ["content", "style"], lora_matrix_num=2)
        with torch.no_grad():
            lay.merge_content.uniform_(0, 1)
            lay.merge_style.uniform_(0, 1)
        lin.set_lora_layer(lay)
        lin = lin.to(dev)          # fp32 parameters, bf16 activations (the reference under autocast)
        for n, p in lin.named_parameters():
            merger = "merge_" in n
            lora = "lora_matrix_dic" in n
            p.requires_grad_(merger or (lora and stage == 1))
        return lin

    attn = types.SimpleNamespace(
        spatial_norm=None, group_norm=None, norm_cross=False, heads=heads,
        to_q=mk(C, C, False), to_k=mk(kv, C, False), to_v=mk(kv, C, False),
        to_out=torch.nn.ModuleList([mk(C, C, True), torch.nn.Dropout(0.0)]),
        residual_connection=False, rescale_output_factor=1.0)
    return attn


def _cfg2_cpu(kind, frames_sample):
    """Oracle fwd+bwd (fp32 torch CPU, reference's dense form) on `frames_sample` of the 16 frames."""
    from oracle import processor as op
    g = torch.Generator().manual_seed(0)
    rn = lambda *s: torch.randn(*s, generator=g)
    C, ctx, r = 1280, 2048, 64
    kv = C if kind == "self" else ctx

    def proj(i, o, bias):
        rq = lambda t: t.requires_grad_(True)
        return op.UnzipProj(rn(o, i) / i ** 0.5, rn(o) if bias else None, rq(rn(r, i) / r), rq(rn(o, r) / r),
                            rq(rn(r, i) / r), rq(rn(o, r) / r), rq(torch.rand(o, generator=g)),
                            rq(torch.rand(o, generator=g)), infer=False)
    P = op.SpatialAttnParams(20, proj(C, C, False), proj(kv, C, False), proj(kv, C, False), proj(C, C, True))
    x = rn(frames_sample, 1024, C).requires_grad_(True)
    text = None if kind == "self" else rn(1, 77, ctx).requires_grad_(True)
    t0 = time.perf_counter()
    y = op.animatediff_attn_processor(P, x, text, scale=1.0)
    y.backward(torch.ones_like(y))
    return time.perf_counter() - t0


def run_cfg2(args, common):
    from video_style_transfer_b200 import _lib
    from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    bf = torch.bfloat16
    proc = AnimateDiffAttnProcessor2_0()
    F, S, C = 16, 1024, 1280
    res = {}
    launches = {}
    for kind in ("self", "cross"):
        for stage in (2, 1):
            attn = _midblock_attn(kind, stage, dev)
            x = torch.randn(F, S, C, device=dev).to(bf).requires_grad_(True)
            text = None if kind == "self" else torch.randn(1, 77, 2048, device=dev).to(bf).requires_grad_(True)
            dy = torch.randn(F, S, C, device=dev).to(bf)

            def fwd():
                with torch.no_grad():
                    return proc(attn, x, text)

            def fwd_bwd():
                x.grad = None
                y = proc(attn, x, text)
                y.backward(dy)

            for _ in range(max(3, args.warmup)):
                fwd_bwd()
            l0 = _lib.launch_count
            fwd_bwd()
            launches[f"{kind}_stage{stage}"] = _lib.launch_count - l0
            res[f"{kind}_stage{stage}_fwd_bwd_ms"] = _timed(fwd_bwd, args.steps, 1, dev)
            if stage == 2:
                for _ in range(3):
                    fwd()
                res[f"{kind}_fwd_ms"] = _timed(fwd, args.steps, 1, dev)
    # algorithmic FLOPs: attention 4*Bt*h*Sq*Sk*d fwd, x2.5 bwd; projections
    # 2*M*in*out + 2*M*r2*(in+out) fwd, dX doubles it (Stage 2), dA/dB add the low-rank products
    M = F * S
    att = {"self": 4.0 * F * S * S * C, "cross": 4.0 * F * S * 77 * C}
    total_ms = res["self_stage2_fwd_bwd_ms"] + res["cross_stage2_fwd_bwd_ms"]
    line = {"metric": "UnZipLoRA attention processor fwd+bwd, SDXL mid-block (self + cross), ms",
            "value": total_ms, "unit": "ms", "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": total_ms, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "cfg2-processor-fwd-bwd", "tokens": [F, S, C], "heads": 20,
                       "text": [1, 77, 2048], "unziplora_rank": 64,
                       "trainable": "stage2 = dX + mergers (headline); stage1 adds A, B",
                       "l2": "inputs + saved activations (~0.5 GB) exceed the 126 MB L2"},
            "detail_ms": {k: round(v, 4) for k, v in res.items()},
            "attn_core_tflops_fwd_bwd_self": 3.5 * att["self"] / 1e12,
            "gpu_launches": launches, "clocks": None}
    if not args.no_cpu_baseline:
        fs = 2
        sec = (_cfg2_cpu("self", fs) + _cfg2_cpu("cross", fs)) * (F / fs)
        line["cpu_baseline"] = {"value": sec * 1e3, "unit": "ms", "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": f"oracle fp32 fwd+bwd (Stage-1 trainable set) on {fs} of 16 frames, scaled x{F // fs}"}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# cfg4
# --------------------------------------------------------------------------------------------------
def run_cfg4(args, common):
    from video_style_transfer_b200 import _lib, sdxl_hotpath as hpth
    from video_style_transfer_b200.animatediff.temporal_lora import build_spatial_lora_index
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    bf = torch.bfloat16
    F, S = args.frames, args.latent
    torch.manual_seed(1234)
    with torch.device(dev):
        model = hpth.SDXLHotPath(temporal_rank=32, depth_scale=args.depth_scale)
    model = model.train()        # fp32 parameters, bf16 activations: train_animatediff.py under autocast
    plan = model.forward_plan()
    params = hpth.freeze_for_stage2(model, unfreeze_mergers=True)
    index = build_spatial_lora_index(model)
    opt = torch.optim.AdamW(params, lr=1e-5, weight_decay=1e-2, fused=True, capturable=not args.no_graph)
    shapes = hpth.level_shapes(F, S)
    g = torch.Generator().manual_seed(0)
    xs4 = [torch.randn(s, generator=g).to(bf).to(dev) for s in shapes]
    xs3 = [None] + [torch.randn(s[0], s[2] * s[3], s[1], generator=g).to(bf).to(dev) for s in shapes[1:]]
    text = torch.randn(1, 77, 2048, generator=g).to(bf).to(dev)
    # upstream gradients dL/dy of every module output (what the rest of the UNet would send back)
    targets = {"motion": [(1e-3 * torch.randn(s, generator=g)).to(bf).to(dev) for s in shapes],
               "self": [None] + [1e-3 * torch.randn_like(t) for t in xs3[1:]],
               "cross": [None] + [1e-3 * torch.randn_like(t) for t in xs3[1:]]}

    def step_eager():
        return hpth.train_step_hotpath(model, plan, xs4, xs3, text, targets, F, opt, index)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(max(3, args.warmup)):
            loss = step_eager()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    if args.ncu_step:
        torch.cuda.profiler.start()
        step_eager()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    graph = None
    if not args.no_graph:
        # whole training step (forward, backward, clip, AdamW) captured once: ~4,300 kernel launches
        # plus the autograd bookkeeping between them are replayed without the Python interpreter
        graph = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        l0 = _lib.launch_count
        with torch.cuda.graph(graph):
            loss = step_eager()
        captured = _lib.launch_count - l0
        step = graph.replay
        for _ in range(2):
            step()
    else:
        step = step_eager
    torch.cuda.synchronize()
    prof_out = os.environ.get("VST_KERNEL_TABLE")
    if prof_out:
        # per-kernel device time of ONE step (CUPTI through torch.profiler: kernels of a graph replay are recorded
        # like any others; far cheaper than an ncu pass over ~4000 launches) -> tab-separated file
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        rows = sorted(((e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0),
                      key=lambda r: -r[2])
        tot = sum(r[2] for r in rows)
        with open(prof_out, "w") as f:
            f.write(f"# one cfg4 training step, kernel device time (us), total {tot:.1f} us over {sum(r[1] for r in rows)} launches\n")
            f.write("kernel\tlaunches\ttotal_us\tshare\n")
            for k, n, t in rows:
                f.write(f"{k[:120]}\t{n}\t{t:.1f}\t{t / tot:.4f}\n")
    l0 = _lib.launch_count
    with common["ClockSampler"](0) as clk:
        ms = _timed(step, args.steps, 1, dev)
    launches = captured if graph is not None else (_lib.launch_count - l0) // args.steps
    fl = hpth.hotpath_flops(plan, F, S)
    fwd = sum(fl.values())
    # backward: attention cores 2.5x forward; GEMMs 2x forward for dX + dW of the trainable
    # matrices (motion FF / proj_in / proj_out full dW; LoRA A/B low-rank) -- reported as a
    # lower bound "fwd + dX" figure: 3x forward GEMMs + 3.5x attention
    att = fl["attn_self"] + fl["attn_cross"] + fl["temporal_attn"]
    step_fl = 3.5 * att + 3.0 * (fwd - att)
    line = {"metric": "Stage-2 training step (temporal LoRA + UnZipLoRA mergers) steps/s, hot-path stack, 1x16 frames",
            "value": 1e3 / ms, "unit": "steps/s", "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "cfg4-train-step", "frames": F, "latent": S, "temporal_lora_rank": 32,
                       "unziplora_rank": 64, "lambda_orth": 1e-4, "unfreeze_mergers": True,
                       "optimizer": "torch AdamW (fused), fp32 parameters",
                       "launch": "eager" if graph is None else "CUDA graph replay (fwd + bwd + clip + AdamW)",
                       "trainable_params": sum(p.numel() for p in params),
                       "gradient_checkpointing": "off (activations of the whole stack fit in HBM)",
                       "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1),
                       "step_tflop_fwd_plus_dx": step_fl / 1e12},
            "tflops_fwd_plus_dx": step_fl / (ms * 1e-3) / 1e12,
            "orth_loss": float(loss), "gpu_launches": launches, "clocks": clk.summary()}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# cfg5
# --------------------------------------------------------------------------------------------------
def run_cfg5(args, common):
    import torch.distributed as dist
    from video_style_transfer_b200 import _lib, sdxl_hotpath as hpth
    from video_style_transfer_b200.sequence_parallel import SPComm, SPMotionStep
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    bf = torch.bfloat16
    F = args.frames if args.frames != 16 else 32
    S = args.latent
    fl_ = F // world
    torch.manual_seed(1234)                      # every rank holds the same weights (one video)
    with torch.device(dev):
        model = hpth.SDXLHotPath(temporal_rank=32, depth_scale=args.depth_scale)
    model = model.to(bf).eval()
    plan = model.forward_plan()
    shapes = hpth.level_shapes(fl_, S)
    g = torch.Generator().manual_seed(100 + rank)
    xs4 = [torch.randn(s, generator=g).to(bf).to(dev) for s in shapes]
    xs3 = [None] + [torch.randn(s[0], s[2] * s[3], s[1], generator=g).to(bf).to(dev) for s in shapes[1:]]
    texts = [torch.randn(1, 77, 2048, generator=g).to(bf).to(dev) for _ in range(2)]
    comm = None
    if world > 1:
        nbytes = max(SPMotionStep.buffer_bytes(1, s[1], s[2] * s[3], F, world) for s in shapes)
        comm = SPComm.from_process_group(nbytes)

    def step_eager():
        outs = None
        for text in texts:
            if world > 1:
                outs = hpth.run_unet_hotpath_sp(plan, xs4, xs3, text, F, comm)
            else:
                outs = hpth.run_unet_hotpath(plan, xs4, xs3, text, F)
        return outs

    graphed = None
    if not args.no_graph:       # every rank captures the same launch sequence, replays in lockstep
        graphed = (hpth.GraphedSPDenoiseStep(plan, xs4, xs3, texts, F, comm) if world > 1
                   else hpth.GraphedDenoiseStep(plan, xs4, xs3, texts, F))
    step = step_eager if graphed is None else graphed.replay

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    l0 = _lib.launch_count
    with common["ClockSampler"](local) as clk:
        ms = _timed(step, args.steps, world, dev)
    launches = (_lib.launch_count - l0) // args.steps if graphed is None else graphed.launches
    from video_style_transfer_b200 import fused as fusedmod
    fl = hpth.hotpath_flops(plan, F, S, merged=fusedmod.MERGE_LORA)
    step_fl = 2.0 * sum(fl.values())
    if comm is not None:
        torch.cuda.synchronize()
        dist.barrier()
        comm.close()
    if rank == 0:
        egress = 0
        if world > 1:     # bytes this rank pushes to / pulls from peers per step (both all-to-alls)
            for kind, lvl, _ in plan:
                if kind == "motion":
                    s = shapes[lvl]
                    egress += 2 * 2 * s[0] * s[1] * s[2] * s[3] * 2 * (world - 1) // world
        line = {"metric": "32f 1024^2 single-video denoise steps/s, sequence-parallel hot path (x2 CFG)",
                "value": 1e3 / ms, "unit": "steps/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": "cfg5-sequence-parallel", "frames": F, "latent": S,
                           "frames_per_rank": fl_, "parallelism": f"sp{world} (frames <-> positions)",
                           "launch": "eager" if graphed is None else "CUDA graph replay (one graph per rank)",
                           "nvlink_bytes_per_rank_per_step": egress, "step_tflop": step_fl / 1e12},
                "tflops_hotpath": step_fl / (ms * 1e-3) / 1e12, "gpu_launches": launches,
                "clocks": clk.summary()}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run(args, common):
    {"cfg2": run_cfg2, "cfg4": run_cfg4, "cfg5": run_cfg5}[args.workload](args, common)
