"""Drop-in for unziplora_unet/unziplora_linear_layer.py (reference :7-346).

Same constructor arguments, sub-module / parameter names (state-dict compatible:
``lora_matrix_dic.{content,style}_{down,up}.weight``, ``merge_content``, ``merge_style``),
attributes and bookkeeping methods.  ``forward`` no longer materialises the two (in x out)
products and three dense GEMMs of the reference (:303-323); it runs
u = x [A_c; A_s]^T and delta = u [m_c (.) B_c | m_s (.) B_s]^T on the tcgen05 GEMM, and when the
layer sits inside a ``LoRACompatibleLinear`` the second product is folded into that linear's own
GEMM (``fused_operands``).  The mask / cone / merger bookkeeping stays host-side torch on the
same tensors, as SURVEY.md §8b specifies.
"""
from typing import List, Optional, Tuple, Union

import torch
from torch import nn

from .. import ops
from .._host import as_rows_bf16, bf16_cached, cached_build, pad8, packed_up, pad_branch_cols

_KEYS = ("content", "style")


class _UnZipBase(nn.Module):
    """State and the fused forward shared by the training and the inference layer."""

    hard_mask_multiplies = False   # training layer: bool column mask multiplies (:236-237)
                                   # inference layer: a block-masked branch is zeros (:308,:316)

    def _build(self, in_features, out_features, rank, lora_matrix_key, device, dtype):
        self.lora_matrix_dic = nn.ModuleDict()
        self.masked_matrix = {}
        for key in lora_matrix_key:
            down = nn.Linear(in_features, rank, bias=False, device=device, dtype=dtype)
            up = nn.Linear(rank, out_features, bias=False, device=device, dtype=dtype)
            nn.init.normal_(down.weight, std=1 / rank)
            nn.init.normal_(up.weight, std=1 / rank)
            self.lora_matrix_dic[f"{key}_down"] = down
            self.lora_matrix_dic[f"{key}_up"] = up
            self.masked_matrix[key] = False
        self.merge_content = nn.Parameter(torch.ones(out_features, device=device, dtype=dtype))
        self.merge_style = nn.Parameter(torch.ones(out_features, device=device, dtype=dtype))
        self.lora_matrix_key = lora_matrix_key
        self.in_features = in_features
        self.out_features = out_features
        self.rank = rank
        self.dtype = dtype
        self.device = device
        self.forward_type = "both"

    # ---- reference API ---------------------------------------------------------------------------
    def set_layer_mask(self, key, value=True):
        self.masked_matrix[key] = value

    def set_forward(self, type: str = "both"):
        assert type in ["both", "content", "style"]
        self.forward_type = type

    # ---- fused operands -----------------------------------------------------------------------------
    def _active(self):
        """[(key, use_merger, hard_mask or None)] for the current forward type / block masks."""
        keys = _KEYS if self.forward_type == "both" else (self.forward_type,)
        use_merger = self.forward_type == "both"          # :331,:343 / :249-261 skip the merger
        out = []
        for key in keys:
            masked = self.masked_matrix.get(key, False) is True
            if masked and not self.hard_mask_multiplies:
                continue                                   # branch contributes zeros
            mask = getattr(self, f"mask_{key}") if masked else None
            out.append((key, use_merger, mask))
        return out

    def site_downs(self):
        """Down-projection weights of the active branches, in branch order."""
        return [self.lora_matrix_dic[f"{k}_down"].weight for k, _, _ in self._active()]

    def site_up(self, scale: float = 1.0) -> Optional[torch.Tensor]:
        """B' (out, nact*pad8(r)) = scale * [m_c (.) mask_c (.) B_c | m_s (.) mask_s (.) B_s], bf16,
        cached until a parameter, mask, the scale or the forward type changes."""
        act = self._active()
        if not act:
            return None
        d = self.lora_matrix_dic
        ups, mrg, msk = {}, {}, {}
        for k, use_merger, mask in act:
            ups[k] = d[f"{k}_up"].weight
            mrg[k] = getattr(self, f"merge_{k}") if use_merger else None
            msk[k] = mask
        return packed_up(ups.get("content"), mrg.get("content"), msk.get("content"),
                         ups.get("style"), mrg.get("style"), msk.get("style"), float(scale), self.rank)

    def fused_operands(self, x_content: torch.Tensor, x_style: Optional[torch.Tensor],
                       scale: float = 1.0) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
        """(u (M,R), B' (out,R)) such that scale * self(x_content, x_style) == u B'^T, with
        x_* already 2-D bf16.  None when every branch is block-masked."""
        act = self._active()
        if not act:
            return None
        if x_style is None:
            x_style = x_content
        xs = {"content": x_content, "style": x_style}
        downs = self.site_downs()
        if len(act) == 2 and x_style is x_content:
            u = ops.lora_gemm(x_content, bf16_cached(*downs, pad_rows=True))   # one GEMM, N = 2r
        elif torch.is_grad_enabled() and any(t.requires_grad for t in (x_content, x_style, *downs)):
            u = torch.cat([ops.lora_gemm(xs[k], bf16_cached(downs[i], pad_rows=True))
                           for i, (k, _, _) in enumerate(act)], dim=1)
        else:
            m, r8 = x_content.shape[0], pad8(self.rank)
            u = torch.empty((m, len(act) * r8), dtype=torch.bfloat16, device=x_content.device)
            for i, (k, _, _) in enumerate(act):
                ops.lora_gemm(xs[k], bf16_cached(downs[i], pad_rows=True), out=u[:, i * r8:(i + 1) * r8])
        return u, self.site_up(scale)

    def forward(self, hidden_states_content: torch.Tensor,
                hidden_states_style: torch.Tensor = None) -> torch.Tensor:
        ref = hidden_states_content if self.forward_type != "style" or hidden_states_style is None \
            else hidden_states_style
        if self.forward_type == "style" and hidden_states_style is None:
            if self.hard_mask_multiplies:
                raise AttributeError("forward_type 'style' needs hidden_states_style")  # :256
            hidden_states_style = hidden_states_content                               # :335-336
        xc = as_rows_bf16(hidden_states_content)
        xs = None
        if hidden_states_style is not None and hidden_states_style is not hidden_states_content:
            xs = as_rows_bf16(hidden_states_style)
        lead = ref.shape[:-1]
        ops_ = self.fused_operands(xc, xs, 1.0)
        if ops_ is None:
            return torch.zeros(lead + (self.out_features,), dtype=ref.dtype, device=ref.device)
        u, bp = ops_
        delta = ops.lora_gemm(u, bp)
        return delta.reshape(lead + (self.out_features,)).to(ref.dtype)


class UnZipLoRALinearLayerInfer(_UnZipBase):
    """Reference: unziplora_linear_layer.py:265-346."""

    def __init__(self, in_features: int, out_features: int, rank: int = 64,
                 lora_matrix_key: List[str] = None,
                 device: Optional[Union[torch.device, str]] = None,
                 dtype: Optional[torch.dtype] = None):
        super().__init__()
        self._build(in_features, out_features, rank, lora_matrix_key, device, dtype)


class UnZipLoRALinearLayer(_UnZipBase):
    """Reference: unziplora_linear_layer.py:7-263 (training variant: hard column masks multiply,
    cone scoring and merger bookkeeping)."""

    hard_mask_multiplies = True

    def __init__(self, in_features: int, out_features: int, rank: int = 64,
                 lora_matrix_key: List[str] = None,
                 device: Optional[Union[torch.device, str]] = None,
                 dtype: Optional[torch.dtype] = None, **model_kwargs):
        super().__init__()
        self._build(in_features, out_features, rank, lora_matrix_key, device, dtype)
        self.fixed_matrix = {}
        self.lora_matrix_dic_norm = {}
        for key in lora_matrix_key:
            for part in ("down", "up"):
                w = self.lora_matrix_dic[f"{key}_{part}"].weight.detach()
                self.lora_matrix_dic_norm[f"{key}_norm_{part}"] = torch.norm(w, dim=0, keepdim=True)
        self.column_score_content = torch.ones(out_features)
        self.column_score_style = torch.ones(out_features)
        self.mask_content = torch.zeros(out_features, device=device, dtype=torch.bool)
        self.mask_style = torch.zeros(out_features, device=device, dtype=torch.bool)

    # ---- bookkeeping (host-side; reference :53-220) ----------------------------------------------
    def _w(self, key, part):
        return self.lora_matrix_dic[f"{key}_{part}"].weight

    def set_cone_score(self, key):
        setattr(self, f"column_score_{key}", torch.zeros(self.out_features))

    def compute_mergers_similarity(self):
        mc, ms = self.merge_content, self.merge_style
        if self.masked_matrix["content"] is not False and self.masked_matrix["style"] is not False:
            mc, ms = mc * self.mask_content, ms * self.mask_style
        return (mc * ms).abs().mean().unsqueeze(0)

    def set_merger_gradient(self, key, value=False):
        getattr(self, f"merge_{key}").requires_grad = value

    def clamp_merger(self, key):
        getattr(self, f"merge_{key}").clamp_(0, 1)

    def get_merger_mask(self, key):
        return getattr(self, f"merge_{key}")

    def _dense(self, key, with_merger):
        dense = self._w(key, "down").T @ self._w(key, "up").data.T          # (in, out)
        return dense * getattr(self, f"merge_{key}") if with_merger else dense

    def get_unziplora_norm(self, key, dim="L2", quick_log=False, multiple=True):
        dense = self._dense(key, multiple is True)
        if self.masked_matrix[key] is True:
            dense = dense * getattr(self, f"mask_{key}")
        p = {"L2": "fro", "L1": 1, "nuc": "nuc"}[dim]
        norm = torch.norm(dense, p=p)
        return torch.tensor(norm.item()).to(self.device) if quick_log else norm.unsqueeze(0)

    def get_unziplora_weight(self, key):
        col = getattr(self, f"mask_{key}") if self.masked_matrix[key] is not False \
            else getattr(self, f"merge_{key}")
        return self._w(key, "down").data, self._w(key, "up").data * col.unsqueeze(1)

    def get_unziplora_cone(self, key, accumulate=True):
        """weight x gradient "cone" per (in, out) entry from the grads of A, B and the merger."""
        m = getattr(self, f"merge_{key}")
        down, up = self._w(key, "down"), self._w(key, "up")
        dense = down.data.T @ up.data.T
        grad = down.data.T @ up.grad.T * m
        if down.grad is not None:
            grad = down.grad.T @ up.data.T * m + grad
        if m.grad is not None:
            grad = grad + dense * m.grad
        cone = dense * grad
        if accumulate:
            prev = getattr(self, f"column_score_{key}").to(self.device)
            setattr(self, f"column_score_{key}", prev + cone.to(self.device))
        else:
            sparsity = torch.sum(torch.abs(cone) > 1e-5, dim=0).float() / cone.shape[0]
            setattr(self, f"column_score_{key}", sparsity)

    def set_gradient_mask(self, finetune_mask):
        mc, ms = self.merge_content, self.merge_style
        if mc.grad is None or ms.grad is None:
            return
        if finetune_mask:
            mc.grad *= self.mask_content
            ms.grad *= self.mask_style
        else:
            both = self.mask_content & self.mask_style
            mc.grad *= both
            ms.grad *= both

    @staticmethod
    def _top_mask(score, count):
        top, _ = torch.topk(score, count)
        thr = top.min() if top.numel() > 0 else float("inf")
        return score > thr

    def mask_updated_elements(self, key=None, step_ratio=0.1, avoid=True):
        count = int(self.out_features * step_ratio)
        if key is None:
            self.mask_content = self._top_mask(self.column_score_content, count) | self.mask_content
            style_score = self.column_score_style.clone()
            if avoid:
                style_score[self.mask_content] = float("-inf")
            self.mask_style = self._top_mask(style_score, count) | self.mask_style
        else:
            score = getattr(self, f"column_score_{key}")
            setattr(self, f"mask_{key}", self._top_mask(score, count) | getattr(self, f"mask_{key}"))
            other = "content" if key == "style" else "style"
            setattr(self, f"mask_{other}",
                    torch.ones(self.out_features, device=self.device, dtype=torch.bool))

    def log_selected_mask(self, key):
        return getattr(self, f"column_score_{key}") * getattr(self, f"mask_{key}")
