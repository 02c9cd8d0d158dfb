"""Oracle restatement (test infrastructure, torch CPU + gloo) of the sequence-parallel motion module
of SURVEY.md §8e: rank p holds frames [p F/P, (p+1) F/P); GroupNorm statistics are all-reduced
(sum, sum of squares per (batch, group)); the frames->positions and positions->frames all-to-alls
are written with all_gather + slicing (gloo has no all_to_all).  It must agree with the unsharded
oracle/processor.py:motion_module, which is what the CUDA path's SP kernels are checked against."""
from typing import Dict

import torch
import torch.distributed as dist

from . import hotpath as hp
from .processor import motion_module_tokens


def sp_plan(f_total: int, hw: int, world: int, rank: int):
    """(frame range, position range) owned by `rank`."""
    assert f_total % world == 0 and hw % world == 0
    fl, hwl = f_total // world, hw // world
    return (rank * fl, (rank + 1) * fl), (rank * hwl, (rank + 1) * hwl)


def motion_module_sp(x_local, w: Dict[str, torch.Tensor], f_total: int, heads: int = 8, groups: int = 32,
                     lora=None, group=None):
    """x_local: this rank's frames (B*F/P, C, H, W) of a (B*F, C, H, W) video -> same shape."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    bfl, c, h, wd = x_local.shape
    fl = f_total // world
    b = bfl // fl
    (f0, f1), (p0, p1) = sp_plan(f_total, h * wd, world, rank)
    x5 = x_local.reshape(b, fl, c, h, wd).permute(0, 2, 1, 3, 4)            # (B,C,Fl,H,W)
    # GroupNorm over (C/groups, F, H, W): global sums from the local ones
    xg = x5.reshape(b, groups, -1).double()
    stats = torch.stack([xg.sum(-1), (xg * xg).sum(-1)])
    dist.all_reduce(stats, group=group)
    n = xg.shape[-1] * world
    mean = stats[0] / n
    var = stats[1] / n - mean * mean
    rstd = (var + 1e-6).rsqrt()
    cg = c // groups
    xn = (x5.reshape(b, groups, cg, fl, h, wd) - mean.float().view(b, groups, 1, 1, 1, 1)) \
        * rstd.float().view(b, groups, 1, 1, 1, 1)
    xn = xn.reshape(b, c, fl, h, wd) * w["norm.weight"].view(1, c, 1, 1, 1) + w["norm.bias"].view(1, c, 1, 1, 1)
    tok_local_frames = hp.rearrange_fwd(xn).reshape(b, h * wd, fl, c)        # all positions, my frames
    # all-to-all #1: everyone's frames for my positions
    mine = [torch.empty_like(tok_local_frames) for _ in range(world)]
    dist.all_gather(mine, tok_local_frames.contiguous(), group=group)
    tok = torch.cat([t[:, p0:p1] for t in mine], dim=2)                       # (B, HWl, F, C)
    y = motion_module_tokens(tok.reshape(b * (p1 - p0), f_total, c), w, f_total, heads, lora)
    y = y.reshape(b, p1 - p0, f_total, c)
    # all-to-all #2: everyone's positions for my frames
    parts = [torch.empty_like(y) for _ in range(world)]
    dist.all_gather(parts, y.contiguous(), group=group)
    full = torch.cat([t[:, :, f0:f1] for t in parts], dim=1)                  # (B, HW, Fl, C)
    y5 = hp.rearrange_bwd(full.reshape(b * h * wd, fl, c), b, h, wd)          # (B,C,Fl,H,W)
    return y5.permute(0, 2, 1, 3, 4).reshape(bfl, c, h, wd) + x_local
