"""Sequence parallelism for ONE video over the GPUs of an NVSwitch box (BASELINE
config 5).  Every spatial op (the UnZipLoRA attention processors) is independent per frame, every
temporal op (the motion module body) per spatial position, so rank p keeps frames
[p F/P, (p+1) F/P) for the spatial work and positions [p HW/P, (p+1) HW/P) for the temporal work.
The two transposes per motion module are done by the module's own layout kernels writing to /
reading from peer memory over NVLink (csrc/layout.cu ShardArgs, csrc/comm.cu):

    frames-sharded x --GroupNorm partials stored into every peer (vst_comm_allgather_partials)
                     --vst_a2a_frames_to_hw (push of the RAW tokens to the position owners)
        --barrier--> GroupNorm on the received tokens (vst_groupnorm_tokens_apply, same rounding as the
                     fused apply), proj_in .. proj_out on N/P sequences
        --barrier--> vst_a2a_hw_to_frames (pull, + residual) --> frames-sharded output
    (two flag barriers per motion module; round 1 used four)

torch.distributed (NCCL) is plumbing only: it carries the IPC handles at start-up.  No NCCL call
and no host synchronisation sits on the data path, so a whole sequence-parallel denoising step is
capturable in one CUDA graph per rank (the barrier's epoch lives in device memory).  The reference
has no multi-GPU path for a single video at all.
"""
import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import ops
from ._lib import check, lib


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class SPComm:
    """One rank's communicator: a symmetric buffer (data | flags) mapped on every peer."""

    def __init__(self, rank: int, world: int, data_bytes: int, storage: torch.Tensor,
                 peer_ptrs: Sequence[int], opened: Sequence[int] = ()):
        self.rank, self.world, self.data_bytes = rank, world, data_bytes
        self.storage = storage                      # own uint8 buffer: [data_bytes | world * 4 flag bytes]
        self._opened = list(opened)                 # IPC mappings to close
        bufs = (C.c_void_p * world)(*[p for p in peer_ptrs])
        flags = (C.c_void_p * world)(*[p + data_bytes for p in peer_ptrs])
        handle = C.c_void_p()
        check(lib.vst_comm_init(rank, world, bufs, flags, C.byref(handle)), "vst_comm_init")
        self.handle = handle

    # ---- construction ------------------------------------------------------------------------------
    @staticmethod
    def _alloc(data_bytes: int, world: int, device) -> torch.Tensor:
        """[data | flags] as a dedicated cudaMalloc allocation (IPC maps whole allocations), zeroed,
        wrapped as a uint8 tensor through __cuda_array_interface__."""
        nbytes = _align(data_bytes) + 256
        ptr = C.c_void_p()
        with torch.cuda.device(device):
            check(lib.vst_comm_alloc(nbytes, C.byref(ptr)), "vst_comm_alloc")

        class _Holder:
            __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr.value, False),
                                        "version": 3, "strides": None}
        t = torch.as_tensor(_Holder(), device=device)
        assert t.data_ptr() == ptr.value and ptr.value % 256 == 0
        t._vst_owned_ptr = ptr.value
        return t

    @classmethod
    def from_process_group(cls, data_bytes: int, group=None) -> "SPComm":
        """One process per GPU: exchange CUDA-IPC handles of the symmetric buffers over `group`."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        data_bytes = _align(data_bytes)
        dev = torch.device("cuda", torch.cuda.current_device())
        storage = cls._alloc(data_bytes, world, dev)
        h = (C.c_ubyte * 64)()
        check(lib.vst_comm_ipc_get(storage.data_ptr(), h), "vst_comm_ipc_get")
        gathered: List[Optional[bytes]] = [None] * world
        dist.all_gather_object(gathered, bytes(h), group=group)
        ptrs, opened = [], []
        for r in range(world):
            if r == rank:
                ptrs.append(storage.data_ptr())
                continue
            out = C.c_void_p()
            buf = (C.c_ubyte * 64).from_buffer_copy(gathered[r])
            check(lib.vst_comm_ipc_open(buf, C.byref(out)), "vst_comm_ipc_open")
            ptrs.append(out.value)
            opened.append(out.value)
        dist.barrier(group=group)
        return cls(rank, world, data_bytes, storage, ptrs, opened)

    @classmethod
    def local_emulation(cls, world: int, data_bytes: int, device="cuda") -> List["SPComm"]:
        """All ranks in ONE process on one GPU (tests): plain pointers instead of IPC mappings.  The
        flag barrier must not be used (the ranks' kernels would wait on one another on one GPU);
        stream order provides the synchronisation."""
        data_bytes = _align(data_bytes)
        stores = [cls._alloc(data_bytes, world, device) for _ in range(world)]
        ptrs = [s.data_ptr() for s in stores]
        return [cls(r, world, data_bytes, stores[r], ptrs) for r in range(world)]

    def close(self):
        if self.handle:
            lib.vst_comm_fini(self.handle)
            self.handle = None
        for p in self._opened:
            lib.vst_comm_ipc_close(p)
        self._opened = []
        owned = getattr(self.storage, "_vst_owned_ptr", None)
        if owned:
            torch.cuda.synchronize()
            self.storage._vst_owned_ptr = None
            lib.vst_comm_free(owned)

    # ---- operations --------------------------------------------------------------------------------
    def barrier(self):
        check(lib.vst_comm_barrier(self.handle, torch.cuda.current_stream().cuda_stream), "vst_comm_barrier")

    def allgather_partials(self, partials: torch.Tensor, dst_offset: int):
        """partials (rows, nparts, 3) fp32 -> (rows, world*nparts, 3) at dst_offset of every rank."""
        rows, nparts, three = partials.shape
        assert three == 3 and partials.dtype == torch.float32 and partials.is_contiguous()
        assert dst_offset + rows * nparts * self.world * 12 <= self.data_bytes, "symmetric buffer too small"
        check(lib.vst_comm_allgather_partials(self.handle, partials.data_ptr(), rows, nparts, dst_offset,
                                              torch.cuda.current_stream().cuda_stream),
              "vst_comm_allgather_partials")

    def view(self, offset_bytes: int, shape, dtype=torch.bfloat16) -> torch.Tensor:
        """A tensor over this rank's own symmetric buffer."""
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        assert offset_bytes % 16 == 0 and offset_bytes + nbytes <= self.data_bytes, "symmetric buffer too small"
        return self.storage[offset_bytes:offset_bytes + nbytes].view(dtype).view(shape)

    def frames_to_hw(self, x5: torch.Tensor, f_total: int, dst_offset: int, norm=None):
        """x5: this rank's (B, C, F_local, H, W) view.  norm = (mean, rstd, gamma, beta, groups)."""
        b, c, fl, h, w = x5.shape
        assert x5.dtype == torch.bfloat16
        sB, sC, sF = ops._planar_strides(x5)
        mean = rstd = gamma = beta = None
        groups, pdt = 1, 0
        if norm is not None:
            mean, rstd, gamma, beta, groups = norm
            pdt = ops._dt(gamma)
        check(lib.vst_a2a_frames_to_hw(self.handle, x5.data_ptr(), dst_offset, ops._ptr(mean), ops._ptr(rstd),
                                       ops._ptr(gamma), ops._ptr(beta), b, c, fl, f_total, h * w, sB, sC, sF,
                                       groups, pdt, torch.cuda.current_stream().cuda_stream),
              "vst_a2a_frames_to_hw")

    def hw_to_frames(self, src_offset: int, out5: torch.Tensor, f_total: int,
                     residual5: Optional[torch.Tensor] = None):
        b, c, fl, h, w = out5.shape
        sB, sC, sF = ops._planar_strides(out5)
        if residual5 is not None:
            assert residual5.shape == out5.shape and residual5.stride() == out5.stride()
        check(lib.vst_a2a_hw_to_frames(self.handle, src_offset, out5.data_ptr(), ops._ptr(residual5), b, c, fl,
                                       f_total, h * w, sB, sC, sF, torch.cuda.current_stream().cuda_stream),
              "vst_a2a_hw_to_frames")


class SPMotionStep:
    """One sequence-parallel motion-module call on one rank, split into the phases between which the
    ranks have to meet (so a test can drive all ranks of an emulated group from one process)."""

    def __init__(self, module, comm: SPComm, x_local: torch.Tensor, f_total: int):
        tt = module.temporal_transformer if hasattr(module, "temporal_transformer") else module
        self.tt, self.comm, self.f_total = tt, comm, f_total
        bfl, c, h, w = x_local.shape
        fl = f_total // comm.world
        self.b = bfl // fl
        assert self.b * fl == bfl and fl * comm.world == f_total
        assert x_local.dtype == torch.bfloat16 and x_local.is_contiguous()
        self.x = x_local
        self.x5 = x_local.view(self.b, fl, c, h, w).permute(0, 2, 1, 3, 4)
        self.geom = (c, h, w, fl)
        hwl = h * w // comm.world
        self.n_local = self.b * hwl
        self.tok_bytes = _align(self.n_local * f_total * c * 2)
        self.part_off = 2 * self.tok_bytes
        self._part_shape = None

    def local_partials(self) -> torch.Tensor:
        return ops.groupnorm5d_partials(self.x5, self.tt.groups)

    def share_partials(self):
        """Store this rank's GroupNorm partials into every rank's buffer (peer stores)."""
        part = self.local_partials()
        rows, nparts, _ = part.shape
        self._part_shape = (rows, nparts * self.comm.world, 3)
        self.comm.allgather_partials(part, self.part_off)

    def push(self, partials_all: Optional[torch.Tensor] = None):
        """Fused form (GroupNorm applied on the way out): needs the ranks' partials first.
        partials_all: the ranks' partials concatenated along dim 1 (rank order); None = the block
        `share_partials` of all ranks filled in this rank's buffer."""
        if partials_all is None:
            partials_all = self.comm.view(self.part_off, self._part_shape, torch.float32)
        mean, rstd = ops.groupnorm5d_finalize(partials_all, self.tt.norm.eps)
        n = self.tt.norm
        self.comm.frames_to_hw(self.x5, self.f_total, 0, (mean, rstd, n.weight.detach(), n.bias.detach(), self.tt.groups))

    def push_raw(self):
        """This rank's frames, un-normalised, into the token buffers of the position owners (peer
        stores).  Together with share_partials it needs ONE barrier before normalize()."""
        self.comm.frames_to_hw(self.x5, self.f_total, 0, None)

    def normalize(self, partials_all: Optional[torch.Tensor] = None):
        """GroupNorm on the received tokens, in place (same coefficients / rounding as the fused
        apply + rearrange: bit-identical tokens)."""
        if partials_all is None:
            partials_all = self.comm.view(self.part_off, self._part_shape, torch.float32)
        mean, rstd = ops.groupnorm5d_finalize(partials_all, self.tt.norm.eps)
        n = self.tt.norm
        c = self.geom[0]
        tok = self.comm.view(0, (self.n_local * self.f_total, c))
        gamma, beta = n.weight.detach(), n.bias.detach()
        check(lib.vst_groupnorm_tokens_apply(tok.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                                             beta.data_ptr(), self.b, (self.n_local // self.b) * self.f_total, c,
                                             self.tt.groups, ops._dt(gamma), torch.cuda.current_stream().cuda_stream),
              "vst_groupnorm_tokens_apply")

    def compute(self):
        from ._host import bf16_cached
        c, h, w, fl = self.geom
        tt, f = self.tt, self.f_total
        tok = self.comm.view(0, (self.n_local * f, c))
        x2 = ops.lora_gemm(tok, bf16_cached(tt.proj_in.weight), bias=tt.proj_in.bias.detach())
        for block in tt.transformer_blocks:
            x2 = block.forward_tokens(x2, self.n_local, f)
        send = self.comm.view(self.tok_bytes, (self.n_local * f, c))
        ops.lora_gemm_raw(x2, bf16_cached(tt.proj_out.weight), bias=tt.proj_out.bias.detach(), out=send)

    def pull(self) -> torch.Tensor:
        c, h, w, fl = self.geom
        out = torch.empty_like(self.x)
        out5 = out.view(self.b, fl, c, h, w).permute(0, 2, 1, 3, 4)
        self.comm.hw_to_frames(self.tok_bytes, out5, self.f_total, self.x5)
        return out

    @staticmethod
    def buffer_bytes(b: int, c: int, hw: int, f_total: int, world: int, groups: int = 32) -> int:
        nparts = lib.vst_groupnorm5d_partials(c, f_total // world, hw, groups)
        return 2 * _align(b * (hw // world) * f_total * c * 2) + _align(b * groups * nparts * world * 12)


@torch.no_grad()
def motion_module_sp(module, comm: SPComm, x_local: torch.Tensor, f_total: int, group=None) -> torch.Tensor:
    """Inference forward of one motion module with the video's frames sharded over the ranks of
    `comm`.  Graph-capturable: only kernels of this library, on the current stream."""
    step = SPMotionStep(module, comm, x_local, f_total)
    step.share_partials()          # (B*groups, nparts, 3) fp32 -> every peer
    step.push_raw()                # raw tokens -> position owners (peer stores over NVLink)
    comm.barrier()                 # 1: partials and tokens of every rank have landed here
    step.normalize()
    step.compute()
    comm.barrier()                 # 2: every rank's send region is complete
    return step.pull()
    # No trailing barrier: a rank next writes its send region (or a peer's token / partial region) only
    # after barrier 1 of the NEXT module, which every rank reaches only after this pull has finished on
    # its own stream; and no rank passes barrier 2 before all have finished compute(), so the token
    # buffer is never overwritten while its owner still reads it.
