"""Recipe for oracle/_ref: the reference's own pure-torch modules of the hot path, taken from where
they lie under /root/reference (test infrastructure; git-ignored output, nothing is committed).

    python -m oracle.build_ref            # also run by __graft_entry__.build() when the reference is present

The reference is Python, so "building" it is placing the six files that import without diffusers
 in a package tree the GPU box can import:

    oracle/_ref/animatediff/{temporal_transformer,temporal_lora,attention_processor}.py
    oracle/_ref/unziplora_unet/{unziplora_linear_layer,lora_linear,lora_unzip}.py

oracle/_ref/ is listed in .gitignore (never in history) and not in .gpurunignore (it travels to the
GPU box with the snapshot, like the built .so).  Only tests/, __graft_entry__.smoke() and bench.py's
CPU legs import it (through oracle/ref_cpu.py); the product package never does.
"""
import hashlib
import json
import os
import shutil
import sys

REF = os.environ.get("VST_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

FILES = [
    "animatediff/temporal_transformer.py",
    "animatediff/temporal_lora.py",
    "animatediff/attention_processor.py",
    "unziplora_unet/unziplora_linear_layer.py",
    "unziplora_unet/lora_linear.py",
    "unziplora_unet/lora_unzip.py",
]


def build(ref: str = REF, out: str = OUT) -> bool:
    """Returns True when oracle/_ref is in place (freshly made or already there)."""
    if not os.path.isdir(ref):
        return os.path.exists(os.path.join(out, "MANIFEST.json"))
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(ref, rel), os.path.join(out, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = hashlib.sha256(open(src, "rb").read()).hexdigest()
    for pkg in ("animatediff", "unziplora_unet"):
        init = os.path.join(out, pkg, "__init__.py")
        if not os.path.exists(init):
            open(init, "w").close()
    with open(os.path.join(out, "MANIFEST.json"), "w") as f:
        json.dump({"source": ref, "sha256": manifest}, f, indent=1)
    return True


if __name__ == "__main__":
    ok = build()
    print("oracle/_ref ready" if ok else f"no reference tree at {REF} and no oracle/_ref", file=sys.stderr)
    sys.exit(0 if ok else 1)
