"""-m gpu tests of the sequence-parallel motion module: the frames<->positions
all-to-alls fused into the layout kernels.  One GPU: all ranks emulated in one process (plain
pointers instead of IPC mappings, stream order instead of the flag barrier).  Two or more GPUs: the
real thing, one process per GPU through torchrun (CUDA IPC + NVLink flag barrier)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _module(c, seed=0):
    from video_style_transfer_b200.motion_module import MotionModule
    from video_style_transfer_b200.animatediff.temporal_lora import inject_temporal_lora, TemporalLoRALinear
    torch.manual_seed(seed)
    holder = torch.nn.Module()
    holder.motion_modules = torch.nn.ModuleList([MotionModule(c, 8)])
    with torch.no_grad():
        for n, p in holder.named_parameters():
            if "norm" in n or n.endswith("bias"):
                p.add_(0.1 * torch.randn_like(p))
    inject_temporal_lora(holder, rank=32, alpha=1.0)
    with torch.no_grad():
        for m in holder.modules():
            if isinstance(m, TemporalLoRALinear):
                m.lora_B.normal_(0, 0.2)
    return holder.motion_modules[0].cuda()


@pytest.mark.parametrize("world,c,hw,f", [(2, 320, 16, 16), (4, 640, 16, 32), (8, 64, 32, 16)])
def test_sp_motion_module_emulated_ranks_match_unsharded(world, c, hw, f):
    from video_style_transfer_b200.sequence_parallel import SPComm, SPMotionStep
    mm = _module(c)
    x = torch.randn(f, c, hw, hw, device="cuda").to(torch.bfloat16)
    with torch.no_grad():
        want = mm(x, num_frames=f)
        comms = SPComm.local_emulation(world, SPMotionStep.buffer_bytes(1, c, hw * hw, f, world))
        fl = f // world
        steps = [SPMotionStep(mm, comms[r], x[r * fl:(r + 1) * fl].contiguous(), f) for r in range(world)]
        parts = torch.cat([s.local_partials() for s in steps], dim=1)
        for s in steps:
            s.share_partials()                 # peer stores into every rank's buffer
        for s in steps:
            # the gathered block is the unsharded partial order: statistics are bit-identical
            assert torch.equal(s.comm.view(s.part_off, s._part_shape, torch.float32), parts)
            s.push()                           # round-1 form: GroupNorm applied on the way out
        toks_fused = [s.comm.view(0, (s.n_local * f, c)).clone() for s in steps]
        # the shipped form: raw push + partials under ONE barrier, GroupNorm applied by the owner
        for s in steps:
            s.push_raw()
        for s, t in zip(steps, toks_fused):
            s.normalize()
            assert torch.equal(s.comm.view(0, (s.n_local * f, c)), t)     # bit-identical tokens
        for s in steps:
            s.compute()
        got = torch.cat([s.pull() for s in steps], dim=0)
    torch.cuda.synchronize()
    for cm in comms:
        cm.close()
    err = ((got.float() - want.float()).abs().max() / want.float().abs().max()).item()
    assert err < 1e-2, err
    assert torch.nn.functional.cosine_similarity(got.float().flatten(), want.float().flatten(), dim=0) > 0.9999


def test_a2a_layout_is_bit_exact_emulated():
    """Without the norm the two all-to-alls are pure data movement: frames -> positions -> frames
    returns the input bit for bit, and the token rows equal the unsharded rearrange."""
    from video_style_transfer_b200 import ops
    from video_style_transfer_b200.sequence_parallel import SPComm
    world, c, hw, f = 4, 320, 256, 16
    fl, hwl = f // world, hw // world
    x = torch.randn(f, c, 16, 16, device="cuda").to(torch.bfloat16)
    tok_ref = ops.rearrange_fwd(x.view(1, f, c, 16, 16).permute(0, 2, 1, 3, 4))       # (HW, F, C)
    nbytes = hwl * f * c * 2
    comms = SPComm.local_emulation(world, nbytes)
    for r in range(world):
        x5 = x[r * fl:(r + 1) * fl].view(1, fl, c, 16, 16).permute(0, 2, 1, 3, 4)
        comms[r].frames_to_hw(x5, f, 0)
    for r in range(world):
        assert torch.equal(comms[r].view(0, (hwl, f, c)), tok_ref[r * hwl:(r + 1) * hwl])
    outs = []
    for r in range(world):
        out = torch.empty(fl, c, 16, 16, device="cuda", dtype=torch.bfloat16)
        comms[r].hw_to_frames(0, out.view(1, fl, c, 16, 16).permute(0, 2, 1, 3, 4), f)
        outs.append(out)
    assert torch.equal(torch.cat(outs), x)
    torch.cuda.synchronize()
    for cm in comms:
        cm.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs (gpurun --gpus 2)")
def test_sp_motion_module_real_ranks_over_nvlink():
    n = min(torch.cuda.device_count(), 8)
    n = 1 << (n.bit_length() - 1)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "sp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "sp_worker ok" in r.stdout
