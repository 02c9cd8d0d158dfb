"""TEST INFRASTRUCTURE.  Generate tests/golden/formats.npz by running the REFERENCE's own on-disk
format helpers (imported from /root/reference in the build container) on the toy UNet of
oracle/toy_unet.py:

  unziplora_unet/utils.py   insert_unziplora_to_unet :388-483 (reads the LoRA safetensors + merger
                            .pth files), unet_inverse_ziplora_state_dict :289-310 (Stage-1 export)
  animatediff/utils.py      freeze_spatial_layers :66-96, save_checkpoint :103-146 with
                            _extract_merger_state_dicts :149-163 (motion_modules.pth,
                            merger_*_stage2.pth)
  animatediff/temporal_lora.py  inject_temporal_lora / get_merged_motion_state_dict :86-192

Both utils files import diffusers / matplotlib-backed modules at module scope that are not in this
image; only those IMPORT LINES are satisfied by empty stub modules -- every function body that runs
is the reference's.  The files the reference writes are read back and stored (key names and
tensors), so tests/test_cpu.py can require this repository's savers / loaders to produce and accept
exactly the same files.

Run here only:  python -m oracle.make_golden_formats
"""
import logging
import os
import sys
import tempfile
import types

import numpy as np
import torch

REF = os.environ.get("VST_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
RANK, TRANK = 4, 2


def _stub_imports():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _Any:
        def __init__(self, *a, **kw):
            pass

    log = types.SimpleNamespace(get_logger=lambda name=None: logging.getLogger(name or "ref"))
    stub("diffusers", DDPMScheduler=_Any, AutoencoderKL=_Any, MotionAdapter=_Any,
         UNet2DConditionModel=_Any, UNetMotionModel=_Any)
    stub("diffusers.utils", logging=log)
    stub("unziplora_unet.pipeline_stable_diffusion_xl", StableDiffusionXLUnZipLoRAPipeline=_Any)
    stub("unziplora_unet.unet_2d_condition", UNet2DConditionModel=_Any)
    stub("record_utils", )
    stub("record_utils.cone", cone_matrix=None, cone_column_sparsity=None, draw_concatenated_heatmap=None)
    if "huggingface_hub" not in sys.modules:
        try:
            import huggingface_hub  # noqa: F401
        except Exception:
            stub("huggingface_hub", hf_hub_download=None)


def make_inputs(unet, seed=3):
    """The LoRA files and Stage-1 merger files of both branches for every spatial attention."""
    g = torch.Generator().manual_seed(seed)
    names = [".".join(k.split(".")[:-1]) for k in unet.attn_processors if "motion_modules" not in k]
    files, mergers = {}, {}
    for branch in ("content", "style"):
        sd, md = {}, {}
        for n in names:
            attn = unet.get_submodule(n)
            for part in ("to_q", "to_k", "to_v", "to_out.0"):
                lin = attn.get_submodule(part)
                sd[f"unet.unet.{n}.{part}.lora.down.weight"] = torch.randn(RANK, lin.in_features, generator=g)
                sd[f"unet.unet.{n}.{part}.lora.up.weight"] = torch.randn(lin.out_features, RANK, generator=g)
                md[f"unet.{n}.{part}.lora.merge_{branch}"] = torch.rand(lin.out_features, generator=g)
        files[branch], mergers[branch] = sd, md
    return files, mergers


def main():
    sys.path.insert(0, REF)
    _stub_imports()
    from safetensors.torch import save_file
    import unziplora_unet.utils as RU
    import animatediff.utils as RAU
    from animatediff.temporal_lora import inject_temporal_lora, TemporalLoRALinear
    from unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    from oracle.toy_unet import seeded_toy

    out = {}
    unet = seeded_toy()
    for k, v in unet.state_dict().items():
        out[f"toy/{k}"] = v.clone()
    files, mergers = make_inputs(unet)
    for b in files:
        for k, v in files[b].items():
            out[f"in/lora_{b}/{k}"] = v
        for k, v in mergers[b].items():
            out[f"in/merger_{b}/{k}"] = v

    with tempfile.TemporaryDirectory() as tmp:
        for b in files:
            os.makedirs(os.path.join(tmp, b))
            save_file(files[b], os.path.join(tmp, b, RU.LORA_WEIGHT_NAME_SAFE))
            torch.save(mergers[b], os.path.join(tmp, f"merger_{b}.pth"))
        out["lora_file_name"] = np.array(RU.LORA_WEIGHT_NAME_SAFE)
        # ---- the reference reads the files and wires its inference layers ------------------------------
        RU.insert_unziplora_to_unet(unet, os.path.join(tmp, "content"), os.path.join(tmp, "style"),
                                    os.path.join(tmp, "merger_content.pth"), os.path.join(tmp, "merger_style.pth"),
                                    rank=RANK)
        for k, v in unet.state_dict().items():
            out[f"inserted/{k}"] = v.clone()
        # ---- Stage 2: temporal LoRA, freezing, a trained step's worth of change, the checkpoint ---------
        n_wrapped = inject_temporal_lora(unet, rank=TRANK, alpha=1.0)
        out["n_wrapped"] = np.array(n_wrapped)
        gen = torch.Generator().manual_seed(5)
        with torch.no_grad():
            for n, m in unet.named_modules():
                if isinstance(m, TemporalLoRALinear):
                    m.lora_B.copy_(torch.randn(m.lora_B.shape, generator=gen) * 0.3)
                    out[f"lora_B/{n}"] = m.lora_B.clone()
                    out[f"lora_A/{n}"] = m.lora_A.clone()
            for n, m in unet.named_modules():
                ll = getattr(m, "lora_layer", None)
                if ll is not None:
                    ll.merge_content.mul_(0.5)
                    ll.merge_style.add_(0.25)
        RAU.freeze_spatial_layers(unet, unfreeze_mergers=True)
        for n, p in unet.named_parameters():
            out[f"requires_grad/{n}"] = np.array(p.requires_grad)
        RAU.save_checkpoint(unet, tmp, 7, save_mergers=True)
        ck = os.path.join(tmp, "checkpoint-7")
        out["checkpoint_files"] = np.array(sorted(os.listdir(ck)))
        for fn in sorted(os.listdir(ck)):
            for k, v in torch.load(os.path.join(ck, fn)).items():
                out[f"ckpt/{fn}/{k}"] = v

    # ---- Stage-1 export of trained layers (unet_inverse_ziplora_state_dict) --------------------------------
    unet1 = seeded_toy()
    gen = torch.Generator().manual_seed(9)
    for n in [".".join(k.split(".")[:-1]) for k in unet1.attn_processors if "motion_modules" not in k]:
        attn = unet1.get_submodule(n)
        for parent, leaf in ((attn, "to_q"), (attn, "to_k"), (attn, "to_v"), (attn.to_out, "0")):
            RU._make_lora_compatible(parent, leaf)
            lin = getattr(parent, leaf)
            lay = UnZipLoRALinearLayer(lin.in_features, lin.out_features, RANK, ["content", "style"],
                                       lora_matrix_num=2)
            with torch.no_grad():
                lay.merge_content.copy_(torch.rand(lay.merge_content.shape, generator=gen))
                lay.merge_style.copy_(torch.rand(lay.merge_style.shape, generator=gen))
            lin.set_lora_layer(lay)
            for k, v in lay.state_dict().items():
                out[f"stage1/layer/{n}.{leaf if parent is attn else 'to_out.0'}/{k}"] = v.clone()
    for key in ("content", "style"):
        lora_sd, mask_sd = RU.unet_inverse_ziplora_state_dict(unet1, key)
        for k, v in lora_sd.items():
            out[f"stage1/export_{key}/lora/{k}"] = v.detach().clone()
        for k, v in mask_sd.items():
            out[f"stage1/export_{key}/mask/{k}"] = v.detach().clone()

    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "formats.npz"),
                        **{k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
                           for k, v in out.items()})
    print(f"formats.npz: {len(out)} arrays")


if __name__ == "__main__":
    main()
