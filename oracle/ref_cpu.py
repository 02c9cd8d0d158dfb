"""The CPU arm of the measurement (test infrastructure): the reference's own modules of the hot path,
assembled the way the live path assembles them, timed on the host cores by bench.py.

When oracle/_ref is present (oracle/build_ref.py placed the reference's six pure-torch files there)
the arm runs the REFERENCE code: `AnimateDiffAttnProcessor2_0` (animatediff/attention_processor.py
:18-96, which calls F.scaled_dot_product_attention :78-80) on a duck-typed `attn`
whose projections are `lora_linear.LoRACompatibleLinear` + `UnZipLoRALinearLayerInfer` (the dense
3-GEMM form of unziplora_linear_layer.py:298-346), and a motion block whose eight attention
projections are the reference's `TemporalLoRALinear` (temporal_lora.py:10-41).  The rest of the
motion block lives in diffusers (`TransformerTemporalModel`, un-vendored): it is restated here with
the torch library calls diffusers makes (nn.GroupNorm, nn.LayerNorm, F.scaled_dot_product_attention,
F.gelu), i.e. the same restatement as oracle/processor.py:motion_module but on library kernels.
Without oracle/_ref the arm falls back to the oracle port (kind "port").

Only bench.py (`--impl reference`, `cpu_baseline`) and tests/ import this module.
"""
import os
import sys
import time
import types
from typing import Callable, Dict, List, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def load_reference():
    """Namespace of the reference's classes (imported from oracle/_ref), or None."""
    if not os.path.exists(os.path.join(REF_DIR, "animatediff", "attention_processor.py")):
        return None
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    from animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    from animatediff.temporal_lora import TemporalLoRALinear
    from animatediff.temporal_transformer import PositionalEncoding, TemporalTransformer
    from unziplora_unet.lora_linear import LoRACompatibleLinear
    from unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer, UnZipLoRALinearLayerInfer
    return types.SimpleNamespace(
        AnimateDiffAttnProcessor2_0=AnimateDiffAttnProcessor2_0, TemporalLoRALinear=TemporalLoRALinear,
        PositionalEncoding=PositionalEncoding, TemporalTransformer=TemporalTransformer,
        LoRACompatibleLinear=LoRACompatibleLinear, UnZipLoRALinearLayer=UnZipLoRALinearLayer,
        UnZipLoRALinearLayerInfer=UnZipLoRALinearLayerInfer)


# ---- reference-module assembly --------------------------------------------------------------------
def ref_spatial_attn(ns, c: int, heads: int, kv_dim: int, rank: int = 64):
    """(attn duck-type, processor): the live spatial path (unziplora_unet/utils.py:435-483 wiring)."""
    def mk(i, o, bias):
        lin = ns.LoRACompatibleLinear(i, o, bias=bias)
        lay = ns.UnZipLoRALinearLayerInfer(i, o, rank, ["content", "style"])
        with torch.no_grad():
            lay.merge_content.uniform_(0, 1)
            lay.merge_style.uniform_(0, 1)
        lin.set_lora_layer(lay)
        return lin

    attn = types.SimpleNamespace(
        spatial_norm=None, group_norm=None, norm_cross=False, heads=heads,
        to_q=mk(c, c, False), to_k=mk(kv_dim, c, False), to_v=mk(kv_dim, c, False),
        to_out=nn.ModuleList([mk(c, c, True), nn.Dropout(0.0)]),
        residual_connection=False, rescale_output_factor=1.0)
    return attn, ns.AnimateDiffAttnProcessor2_0()


class RefMotionModule(nn.Module):
    """diffusers TransformerTemporalModel (un-vendored; structure as oracle/processor.py:motion_module)
    on torch library calls, attention projections = the reference's TemporalLoRALinear."""

    def __init__(self, ns, c: int, heads: int = 8, rank: int = 32, alpha: float = 1.0):
        super().__init__()
        self.heads = heads
        self.norm = nn.GroupNorm(32, c, eps=1e-6)
        self.proj_in, self.proj_out = nn.Linear(c, c), nn.Linear(c, c)
        self.norms = nn.ModuleList([nn.LayerNorm(c) for _ in range(3)])
        wrap = lambda bias: ns.TemporalLoRALinear(nn.Linear(c, c, bias=bias), rank, alpha)
        self.attns = nn.ModuleList([nn.ModuleList([wrap(False), wrap(False), wrap(False), wrap(True)])
                                    for _ in range(2)])
        for m in self.modules():
            if isinstance(m, ns.TemporalLoRALinear):
                nn.init.normal_(m.lora_B, 0.0, 1.0 / rank)
        self.ff_in, self.ff_out = nn.Linear(c, 8 * c), nn.Linear(4 * c, c)
        self.register_buffer("pe", ns.PositionalEncoding(c, 32).pe.clone())

    def forward(self, x: torch.Tensor, frames: int) -> torch.Tensor:
        bf, c, h, w = x.shape
        b = bf // frames
        x5 = self.norm(x.reshape(b, frames, c, h, w).permute(0, 2, 1, 3, 4))
        y = self.proj_in(x5.permute(0, 3, 4, 2, 1).reshape(b * h * w, frames, c))
        n, d = y.shape[0], c // self.heads
        sp = lambda t: t.view(n, frames, self.heads, d).transpose(1, 2)
        for norm, (tq, tk, tv, to) in zip(self.norms[:2], self.attns):
            yn = norm(y) + self.pe[:, :frames]
            o = F.scaled_dot_product_attention(sp(tq(yn)), sp(tk(yn)), sp(tv(yn)))
            y = y + to(o.transpose(1, 2).reshape(n, frames, c))
        hid, gate = self.ff_in(self.norms[2](y)).chunk(2, dim=-1)
        y = y + self.ff_out(hid * F.gelu(gate))
        y = self.proj_out(y)
        y = y.view(b, h, w, frames, c).permute(0, 3, 4, 1, 2).reshape(bf, c, h, w)
        return y + x


# ---- oracle-port assembly (fallback when oracle/_ref is absent) ------------------------------------
def _port_spatial(c, heads, kv, g):
    from . import processor as op
    rn = lambda *s: torch.randn(*s, generator=g)
    r = 64

    def proj(i, o, bias):
        return op.UnzipProj(rn(o, i) / i ** 0.5, rn(o) if bias else None, rn(r, i) / r, rn(o, r) / r,
                            rn(r, i) / r, rn(o, r) / r, torch.rand(o, generator=g), torch.rand(o, generator=g))
    p = op.SpatialAttnParams(heads, proj(c, c, False), proj(kv, c, False), proj(kv, c, False), proj(c, c, True))
    return lambda x, text=None: op.animatediff_attn_processor(p, x, text)


def _port_motion(c, g):
    from . import hotpath as hp
    from . import processor as op
    rn = lambda *s: torch.randn(*s, generator=g)
    w = {"norm.weight": torch.ones(c), "norm.bias": torch.zeros(c), "pos_embed.pe": hp.positional_table(c)}
    for n in ("proj_in", "proj_out"):
        w[n + ".weight"], w[n + ".bias"] = rn(c, c) / c ** 0.5, rn(c)
    tb = "transformer_blocks.0."
    for n in ("norm1", "norm2", "norm3"):
        w[tb + n + ".weight"], w[tb + n + ".bias"] = torch.ones(c), torch.zeros(c)
    for a in ("attn1", "attn2"):
        for p in ("to_q", "to_k", "to_v", "to_out.0"):
            w[tb + f"{a}.{p}.base.weight"] = rn(c, c) / c ** 0.5
            w[tb + f"{a}.{p}.lora_A"], w[tb + f"{a}.{p}.lora_B"] = rn(32, c) * 0.01, rn(c, 32) / 32
        w[tb + f"{a}.to_out.0.base.bias"] = rn(c)
    w[tb + "ff.net.0.proj.weight"], w[tb + "ff.net.0.proj.bias"] = rn(8 * c, c) / c ** 0.5, rn(8 * c)
    w[tb + "ff.net.2.weight"], w[tb + "ff.net.2.bias"] = rn(c, 4 * c) / (4 * c) ** 0.5, rn(c)
    return lambda x, frames: op.motion_module(x, w, frames, lora={"scale": 1 / 32})


# ---- the bounded sample of one cfg3 hot-path step ----------------------------------------------------
class CpuArm:
    """One call per (kind, level) of the cfg3 hot-path workload at a bounded shape, with the factor
    that scales it to the full step (L2 shapes in full, L1 / L0 down-batched and
    scaled linearly; every op on the path is linear in frames (spatial) or positions (temporal)).

        step seconds = cfg_passes x sum_k  calls_per_unet[k] x scale[k] x seconds[k]     (module_sum)
    """

    CALLS = {"L2_self": 60, "L2_cross": 60, "L1_self": 10, "L1_cross": 10,
             "L2_motion": 5, "L1_motion": 5, "L0_motion": 5}

    def __init__(self, frames: int = 16, latent: int = 128, threads: int = 0):
        torch.set_num_threads(threads or os.cpu_count() or 1)
        self.cores = torch.get_num_threads()
        self.frames, self.latent = frames, latent
        ns = load_reference()
        self.kind = "reference" if ns is not None else "port"
        g = torch.Generator().manual_seed(0)
        torch.manual_seed(0)
        rn = lambda *s: torch.randn(*s, generator=g)
        s1, s2 = latent // 2, latent // 4
        fs = max(1, frames // 8)                       # L1 spatial: 1/8 of the frames
        text = rn(1, 77, 2048)
        self.units: Dict[str, Tuple[Callable[[], torch.Tensor], float]] = {}

        def spatial(name, c, heads, kv, x, txt, scale):
            if ns is not None:
                attn, proc = ref_spatial_attn(ns, c, heads, kv)
                fn = (lambda: proc(attn, x, encoder_hidden_states=txt, scale=1.0))
            else:
                f = _port_spatial(c, heads, kv, g)
                fn = (lambda: f(x, txt))
            self.units[name] = (fn, scale)

        def motion(name, c, x, scale):
            if ns is not None:
                mm = RefMotionModule(ns, c).eval()
                fn = (lambda: mm(x, frames))
            else:
                f = _port_motion(c, g)
                fn = (lambda: f(x, frames))
            self.units[name] = (fn, scale)

        spatial("L2_self", 1280, 20, 1280, rn(frames, s2 * s2, 1280), None, 1.0)
        spatial("L2_cross", 1280, 20, 2048, rn(frames, s2 * s2, 1280), text, 1.0)
        spatial("L1_self", 640, 10, 640, rn(fs, s1 * s1, 640), None, frames / fs)
        spatial("L1_cross", 640, 10, 2048, rn(fs, s1 * s1, 640), text, frames / fs)
        motion("L2_motion", 1280, rn(frames, 1280, s2, s2), 1.0)
        w1, w0 = max(1, s1 // 8), max(1, latent // 8)   # L1 / L0 temporal: 1/8 of the positions
        motion("L1_motion", 640, rn(frames, 640, s1, w1), s1 / w1)
        motion("L0_motion", 320, rn(frames, 320, latent, w0), latent / w0)

    @property
    def sample_text(self) -> str:
        what = ("reference modules from oracle/_ref (AnimateDiffAttnProcessor2_0 + lora_linear.LoRACompatibleLinear + "
                "UnZipLoRALinearLayerInfer dense 3-GEMM form + F.scaled_dot_product_attention; motion block = "
                "TemporalLoRALinear projections inside the diffusers-shaped block on torch library calls)"
                if self.kind == "reference" else
                "oracle port (torch CPU fp32 restatement; oracle/_ref absent)")
        return (f"{what}, torch CPU fp32, {self.cores} threads; one call per (kind, level): L2 self/cross/motion at "
                "full shape, L1 self/cross on frames/8, L1/L0 motion on 1/8 of the positions; "
                "module_sum: step = 2 CFG x (60 L2 + 10 L1 self+cross, 5 motion per level) with the "
                "down-batch factors applied")

    def factors(self) -> Dict[str, float]:
        return {k: self.CALLS[k] * s for k, (_, s) in self.units.items()}

    @torch.no_grad()
    def sample_pass(self) -> Dict[str, float]:
        """One pass over the bounded sample: seconds per unit call."""
        out = {}
        for k, (fn, _) in self.units.items():
            t0 = time.perf_counter()
            fn()
            out[k] = time.perf_counter() - t0
        return out

    def step_seconds(self, unit_seconds: Dict[str, float], cfg_passes: int = 2) -> float:
        return cfg_passes * sum(self.CALLS[k] * self.units[k][1] * unit_seconds[k] for k in self.units)

    def run(self, passes: int, warmup: int = 2):
        """`warmup` untimed passes, then `passes` timed ones.  Returns (per-pass wall seconds,
        per-unit median seconds, per-unit best seconds)."""
        for _ in range(warmup):
            self.sample_pass()
        walls: List[float] = []
        rows: List[Dict[str, float]] = []
        for _ in range(passes):
            t0 = time.perf_counter()
            rows.append(self.sample_pass())
            walls.append(time.perf_counter() - t0)
        med = {k: sorted(r[k] for r in rows)[len(rows) // 2] for k in self.units}
        best = {k: min(r[k] for r in rows) for k in self.units}
        return walls, med, best
