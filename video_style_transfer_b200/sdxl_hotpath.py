"""The SDXL-AnimateDiff hot-path stack: every spatial attention (UnZipLoRA, rank 64 content+style)
and every motion module (temporal LoRA, rank 32) of one ``UNetMotionModel`` forward, under the
module names diffusers gives them, random-initialised.  It is the workload of BASELINE config 3
restricted to SURVEY.md §8's hot path -- the ResNet / conv / spatial feed-forward layers between
these modules belong to the un-vendored diffusers UNet and are out of scope.

SDXL geometry (unziplora_unet/utils.py:35-47,88-104): latent side S = 128 at 1024^2;
  L0  C=320   S^2        no spatial attention          motion modules: down 2 + up 3
  L1  C=640   (S/2)^2    2 blocks/attn, 10 heads       attn: down 2 + up 3   motion: 2 + 3
  L2  C=1280  (S/4)^2    10 blocks/attn, 20 heads      attn: down 2 + mid 1 + up 3   motion: 2 + 3
=> 70 self + 70 cross attention calls and 15 motion modules per UNet forward.
"""
from typing import List, Tuple

import torch
import torch.nn as nn

from .animatediff.attention_processor import AnimateDiffAttnProcessor2_0
from .animatediff.temporal_lora import inject_temporal_lora
from .motion_module import MotionModule
from .unziplora_unet.lora_linear import LoRACompatibleLinear
from .unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayerInfer

TEXT_DIM = 2048
TEXT_LEN = 77


class SpatialAttention(nn.Module):
    """The slice of diffusers' ``Attention`` that the processor protocol reads
    (animatediff/attention_processor.py:30-94; SURVEY.md §8b duck-type)."""

    def __init__(self, query_dim: int, heads: int, cross_dim: int = None, rank: int = 64):
        super().__init__()
        kv = cross_dim or query_dim
        self.heads = heads
        self.spatial_norm = None
        self.group_norm = None
        self.norm_cross = None
        self.residual_connection = False
        self.rescale_output_factor = 1.0
        self.to_q = LoRACompatibleLinear(query_dim, query_dim, bias=False)
        self.to_k = LoRACompatibleLinear(kv, query_dim, bias=False)
        self.to_v = LoRACompatibleLinear(kv, query_dim, bias=False)
        self.to_out = nn.ModuleList([LoRACompatibleLinear(query_dim, query_dim, bias=True),
                                     nn.Dropout(0.0)])
        for lin in (self.to_q, self.to_k, self.to_v, self.to_out[0]):
            lay = UnZipLoRALinearLayerInfer(lin.in_features, lin.out_features, rank,
                                            ["content", "style"])
            with torch.no_grad():
                lay.merge_content.uniform_(0, 1)
                lay.merge_style.uniform_(0, 1)
            lin.set_lora_layer(lay)
        self.processor = None       # created on first use (needs a CUDA device)

    def set_processor(self, processor):
        self.processor = processor

    def forward(self, hidden_states, encoder_hidden_states=None, **kw):
        if self.processor is None:
            self.processor = AnimateDiffAttnProcessor2_0()
        return self.processor(self, hidden_states, encoder_hidden_states=encoder_hidden_states, **kw)


class _Block(nn.Module):
    def __init__(self, dim, heads):
        super().__init__()
        self.attn1 = SpatialAttention(dim, heads)
        self.attn2 = SpatialAttention(dim, heads, TEXT_DIM)


def _full_block(dim, heads):
    """The whole spatial BasicTransformerBlock (LayerNorms + GEGLU feed-forward around the two
    attentions; unziplora_unet/unzip_attention.py:113-239)."""
    from .unziplora_unet.unzip_attention import BasicTransformerBlock
    return BasicTransformerBlock(dim, heads, 64, cross_attention_dim=TEXT_DIM,
                                 attn1=SpatialAttention(dim, heads), attn2=SpatialAttention(dim, heads, TEXT_DIM))


class _Transformer2D(nn.Module):
    def __init__(self, dim, heads, depth, spatial_ff=False):
        super().__init__()
        mk = _full_block if spatial_ff else _Block
        self.transformer_blocks = nn.ModuleList([mk(dim, heads) for _ in range(depth)])


class _UNetBlock(nn.Module):
    def __init__(self, dim, n_attn, depth, n_motion, spatial_ff=False):
        super().__init__()
        self.dim = dim
        if n_attn:
            self.attentions = nn.ModuleList([_Transformer2D(dim, dim // 64, depth, spatial_ff)
                                             for _ in range(n_attn)])
        if n_motion:
            self.motion_modules = nn.ModuleList([MotionModule(dim, 8) for _ in range(n_motion)])


class SDXLHotPath(nn.Module):
    """levels: (C, latent-side divisor).  ``forward_plan()`` lists the modules in UNet order."""

    def __init__(self, temporal_rank: int = 32, depth_scale: float = 1.0, spatial_ff: bool = False):
        """spatial_ff: build every spatial transformer block in full (norm1/2/3 + GEGLU ff around the two
        attentions); the plan then lists ("block", level, module) entries."""
        super().__init__()
        self.spatial_ff = spatial_ff
        d2, d10 = max(1, round(2 * depth_scale)), max(1, round(10 * depth_scale))
        U = lambda *a: _UNetBlock(*a, spatial_ff=spatial_ff)
        self.down_blocks = nn.ModuleList([U(320, 0, 0, 2), U(640, 2, d2, 2), U(1280, 2, d10, 2)])
        self.mid_block = U(1280, 1, d10, 0)
        self.up_blocks = nn.ModuleList([U(1280, 3, d10, 3), U(640, 3, d2, 3), U(320, 0, 0, 3)])
        self.n_temporal_lora = inject_temporal_lora(self, rank=temporal_rank, alpha=1.0)
        with torch.no_grad():      # B = 0 at init would make the low-rank update a no-op
            for n, p in self.named_parameters():
                if n.endswith("lora_B"):
                    p.normal_(0, 1.0 / temporal_rank)

    @staticmethod
    def level_of(dim: int) -> int:
        return {320: 0, 640: 1, 1280: 2}[dim]

    def forward_plan(self) -> List[Tuple[str, int, nn.Module]]:
        """[(kind, level, module)] with kind in {self, cross, motion}, in UNet execution order
        (per UNet block: attention i then motion module i, as UNetMotionModel interleaves them)."""
        plan = []
        for blk in list(self.down_blocks) + [self.mid_block] + list(self.up_blocks):
            lvl = self.level_of(blk.dim)
            attns = list(getattr(blk, "attentions", []))
            motions = list(getattr(blk, "motion_modules", []))
            for i in range(max(len(attns), len(motions))):
                if i < len(attns):
                    for tb in attns[i].transformer_blocks:
                        if self.spatial_ff:
                            plan.append(("block", lvl, tb))
                        else:
                            plan.append(("self", lvl, tb.attn1))
                            plan.append(("cross", lvl, tb.attn2))
                if i < len(motions):
                    plan.append(("motion", lvl, motions[i]))
        return plan


def level_shapes(frames: int, latent: int = 128, batch: int = 1):
    """Per level: (B*F, C, H, W) hidden-state shape."""
    return [(batch * frames, 320, latent, latent), (batch * frames, 640, latent // 2, latent // 2),
            (batch * frames, 1280, latent // 4, latent // 4)]


@torch.no_grad()
def run_unet_hotpath(plan, xs4, xs3, text, frames: int):
    """One UNet forward's worth of hot-path calls.  xs4[l]: (B*F,C,H,W) motion-module input,
    xs3[l]: (B*F,HW,C) spatial token input, text: (B,77,2048).  Every module reads its level's
    input (the conv/FF layers that would sit between them are out of scope) and the last output
    per level is returned."""
    outs = [None, None, None]
    for kind, lvl, mod in plan:
        if kind == "self":
            outs[lvl] = mod(xs3[lvl])
        elif kind in ("cross", "block"):
            outs[lvl] = mod(xs3[lvl], encoder_hidden_states=text)
        else:
            outs[lvl] = mod(xs4[lvl], num_frames=frames)
    return outs


class GraphedDenoiseStep:
    """One denoising step's worth of hot-path work captured ONCE into a CUDA graph and replayed: the
    step is ~1,200-2,300 kernel launches of 10-800 us each, and the Python + launch overhead between
    them (~10 % of the step at 16 frames) disappears on replay.

    `texts` lists the text embeddings of the step, one run_unet_hotpath per entry:
      * two-pass CFG (inference_animatediff.py:110-121): texts = [uncond (1,77,2048), cond], xs hold
        the F frames of the video and both passes read them;
      * batched CFG (pipeline_animatediff_xl.py:353,373-375): texts = [cat(uncond, cond) (2,77,2048)],
        xs hold 2F frames (uncond half, cond half): every weight is read once per step and every
        GEMM runs at twice the M.
    Inputs live in static device buffers (`x4`, `x3`); copy new data into them before ``replay()``.
    The text is per PROMPT, not per step: ``set_text`` copies it and refreshes the cached text K / V
    of every cross-attention (attention_processor.text_kv) outside the graph.
    Everything the C ABI needs is graph-safe: no host synchronisation, workspaces come from the
    caching allocator's graph pool, tensor maps travel as kernel parameters."""

    def __init__(self, plan, xs4, xs3, texts, frames: int, warmup: int = 2):
        from . import _lib
        self.plan, self.frames = plan, frames
        self.x4 = [t.clone() for t in xs4]
        self.x3 = [None] + [t.clone() for t in xs3[1:]]
        self.texts = [t.clone() for t in texts]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                   # bf16 caches, text K/V, smem attributes, allocator
            for _ in range(warmup):
                self._run()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count
        with torch.cuda.graph(self.graph):
            self.outs = self._run()
        self.launches = _lib.launch_count - n0           # C-ABI calls captured per replay

    def _run(self):
        outs = None
        for text in self.texts:
            outs = run_unet_hotpath(self.plan, self.x4, self.x3, text, self.frames)
        return outs

    @torch.no_grad()
    def set_text(self, texts):
        """New prompt: copy the embeddings into the static buffers and recompute the text K / V the
        captured graph reads (in place, eagerly, on the current stream)."""
        from .animatediff.attention_processor import text_kv
        for dst, src in zip(self.texts, texts):
            dst.copy_(src, non_blocking=True)
        for kind, _, mod in self.plan:
            if kind in ("cross", "block"):
                for t in self.texts:
                    text_kv(mod.attn2 if kind == "block" else mod, t, 1.0)

    def replay(self):
        self.graph.replay()
        return self.outs


class StreamedDenoiseStep:
    """Host-fed pipeline around a GraphedDenoiseStep: every step's inputs come from (pinned) host
    tensors and its outputs go back to host tensors, with the PCIe traffic of step n+1 / n-1 hidden
    behind the graph replay of step n (the prompt's text embedding is not a per-step input: it is
    set once with GraphedDenoiseStep.set_text).  Three streams: `h2d` fills one of two device staging sets,
    the compute stream moves it into the graph's static inputs (device-to-device, ~0.3 ms for the
    420 MB of a 16-frame step), replays, and parks the outputs in one of two output staging sets,
    `d2h` drains those to the host.  submit() never blocks the host; drain() makes the calling
    stream wait for the last device-to-host copy."""

    def __init__(self, graphed: "GraphedDenoiseStep"):
        self.g = graphed
        self.dst = graphed.x4 + graphed.x3[1:]        # the text is per prompt: GraphedDenoiseStep.set_text
        self.h2d, self.d2h = torch.cuda.Stream(), torch.cuda.Stream()
        self.stage_in = [[torch.empty_like(t) for t in self.dst] for _ in range(2)]
        self.stage_out = [[torch.empty_like(t) for t in graphed.outs] for _ in range(2)]
        ev = lambda: [torch.cuda.Event() for _ in range(2)]
        self.in_ready, self.in_consumed, self.out_ready, self.out_free = ev(), ev(), ev(), ev()
        self.n = 0

    def submit(self, host_inputs, host_outs):
        """host_inputs: tensors matching x4 + x3[1:] of the wrapped step; host_outs: one
        host tensor per level output (same numel).  Enqueues one step and returns at once."""
        k, first = self.n & 1, self.n < 2
        cur = torch.cuda.current_stream()
        with torch.cuda.stream(self.h2d):
            if not first:
                self.h2d.wait_event(self.in_consumed[k])      # step n-2 has left this staging set
            for d, h in zip(self.stage_in[k], host_inputs):
                d.copy_(h, non_blocking=True)
            self.in_ready[k].record(self.h2d)
        cur.wait_event(self.in_ready[k])
        torch._foreach_copy_(self.dst, self.stage_in[k])
        self.in_consumed[k].record(cur)
        outs = self.g.replay()
        if not first:
            cur.wait_event(self.out_free[k])                  # step n-2's outputs are on the host
        torch._foreach_copy_(self.stage_out[k], list(outs))
        self.out_ready[k].record(cur)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(self.out_ready[k])
            for h, d in zip(host_outs, self.stage_out[k]):
                h.view(d.shape).copy_(d, non_blocking=True)
            self.out_free[k].record(self.d2h)
        self.n += 1

    def drain(self):
        """The calling stream waits until every submitted step's outputs are on the host."""
        cur = torch.cuda.current_stream()
        for k in range(min(self.n, 2)):
            cur.wait_event(self.out_free[k])


class GraphedSPDenoiseStep(GraphedDenoiseStep):
    """GraphedDenoiseStep for ONE video sharded over the ranks of an SPComm: every rank captures its
    own graph of the same launch sequence (spatial processors on its frames, motion modules through
    the fused NVLink all-to-alls and flag barriers) and all ranks replay in lockstep.  `frames` is
    the video's total frame count; xs4 / xs3 hold this rank's frames."""

    def __init__(self, plan, xs4_local, xs3_local, texts, frames: int, comm, warmup: int = 2):
        self.comm = comm
        super().__init__(plan, xs4_local, xs3_local, texts, frames, warmup)

    def _run(self):
        outs = None
        for text in self.texts:
            outs = run_unet_hotpath_sp(self.plan, self.x4, self.x3, text, self.frames, self.comm)
        return outs



@torch.no_grad()
def run_unet_hotpath_sp(plan, xs4_local, xs3_local, text, f_total: int, comm, group=None):
    """run_unet_hotpath for ONE video whose frames are sharded over the ranks of `comm`
    (BASELINE config 5): the spatial processors run on this rank's frames with no
    exchange; every motion module goes through sequence_parallel.motion_module_sp (two fused
    NVLink all-to-alls + the 256-byte GroupNorm reduction)."""
    from .sequence_parallel import motion_module_sp
    outs = [None, None, None]
    for kind, lvl, mod in plan:
        if kind == "self":
            outs[lvl] = mod(xs3_local[lvl])
        elif kind in ("cross", "block"):
            outs[lvl] = mod(xs3_local[lvl], encoder_hidden_states=text)
        else:
            outs[lvl] = motion_module_sp(mod, comm, xs4_local[lvl], f_total, group)
    return outs


def freeze_for_stage2(model: nn.Module, unfreeze_mergers: bool = True):
    """The Stage-2 trainable set (reference animatediff/utils.py:66-95, which needs diffusers to
    import): motion-module parameters except the TemporalLoRALinear bases, plus the UnZipLoRA
    mergers when `unfreeze_mergers`; everything else frozen.  Returns the trainable parameters."""
    out = []
    for name, p in model.named_parameters():
        if "motion_modules" in name:
            p.requires_grad_(not (".base.weight" in name or ".base.bias" in name))
        elif unfreeze_mergers and ("merge_content" in name or "merge_style" in name):
            p.requires_grad_(True)
        else:
            p.requires_grad_(False)
        if p.requires_grad:
            out.append(p)
    return out


def train_step_hotpath(model, plan, xs4, xs3, text, upstream, frames: int, optimizer, spatial_index,
                       lambda_orth: float = 1e-4, max_grad_norm: float = 1.0):
    """One Stage-2 training step over the hot-path stack (train_animatediff.py:212-317 restricted
    to SURVEY.md §8): forward of every module with autograd, ONE backward pass in which every
    module output receives its upstream gradient `upstream[kind][level]` (inside the UNet that
    gradient comes from the layers behind the module; the reference's MSE sits on the final UNet
    output, :298-300) together with the temporal / spatial orthogonality loss (:307-308), gradient
    clipping (:316), optimizer step (:317).  The inputs require grad (every module has to return dX
    to the layer before it)."""
    from .animatediff.temporal_lora import compute_orth_loss
    optimizer.zero_grad(set_to_none=True)
    x4 = [t.detach().requires_grad_(True) for t in xs4]
    x3 = [None] + [t.detach().requires_grad_(True) for t in xs3[1:]]
    outs, grads = [], []
    for kind, lvl, mod in plan:
        if kind == "self":
            y = mod(x3[lvl])
        elif kind in ("cross", "block"):
            y = mod(x3[lvl], encoder_hidden_states=text)
        else:
            y = mod(x4[lvl], num_frames=frames)
        outs.append(y)
        grads.append(upstream[kind][lvl])
    orth = compute_orth_loss(model, spatial_index, lambda_orth)
    if orth.requires_grad:
        outs.append(orth)
        grads.append(torch.ones_like(orth))
    torch.autograd.backward(outs, grads)
    params = [p for g in optimizer.param_groups for p in g["params"]]
    torch.nn.utils.clip_grad_norm_(params, max_grad_norm)
    optimizer.step()
    return orth.detach()


def hotpath_flops(plan, frames: int, latent: int = 128, batch: int = 1, merged: bool = False,
                  text_kv_cached: bool = False):
    """Algorithmic FLOPs of one UNet forward of the plan (SURVEY.md §8d formulas), by kernel family.
    merged=True: the FLOPs actually executed when the frozen adapters are folded into the weights
    (fused.merged_weight): the low-rank terms cost nothing per call."""
    fl = {"attn_self": 0.0, "attn_cross": 0.0, "proj": 0.0, "temporal_attn": 0.0, "motion_gemm": 0.0}
    side = [latent, latent // 2, latent // 4]
    dims = [320, 640, 1280]
    for kind, lvl, _ in plan:
        c, hw = dims[lvl], side[lvl] ** 2
        m = batch * frames * hw
        r2 = 0 if merged else 128
        rt = 0 if merged else 32
        if kind == "block":                  # self + cross + LN/FF of one spatial transformer block
            fl.setdefault("spatial_ff", 0.0)
            fl["spatial_ff"] += 2.0 * m * c * 8 * c + 2.0 * m * 4 * c * c
        if kind in ("self", "block"):
            fl["attn_self"] += 4.0 * batch * frames * hw * hw * c
            fl["proj"] += 4 * (2.0 * m * c * c + 2.0 * m * r2 * (c + c))
        if kind in ("cross", "block"):
            fl["attn_cross"] += 4.0 * batch * frames * hw * TEXT_LEN * c
            mt = batch * TEXT_LEN
            fl["proj"] += 2 * (2.0 * m * c * c + 2.0 * m * r2 * 2 * c)
            if not text_kv_cached:       # once per prompt, not per step (attention_processor.text_kv)
                fl["proj"] += 2 * (2.0 * mt * TEXT_DIM * c + 2.0 * mt * r2 * (TEXT_DIM + c))
        if kind == "motion":
            n = batch * hw
            fl["temporal_attn"] += 2 * 4.0 * n * frames * frames * c
            fl["motion_gemm"] += 2.0 * m * c * c * 2                       # proj_in, proj_out
            fl["motion_gemm"] += 8 * (2.0 * m * c * c + 2.0 * m * rt * 2 * c)   # 2 x (q,k,v,out) + LoRA
            fl["motion_gemm"] += 2.0 * m * c * 8 * c + 2.0 * m * 4 * c * c   # GEGLU FF
    return fl
