"""Stage the UNMODIFIED reference model package for the CPU arm  --  TEST / BENCH INFRASTRUCTURE, NOT PRODUCT.

    python oracle/make_ref.py            # build container only (needs /root/reference)

The reference (Orca0917/TransformerTTS) is not an installable package (no setup.py / pyproject.toml), but the files on
the hot path -- model/{__init__,model,layers,module}.py and loss.py -- need only torch.  This recipe copies them byte for
byte from /root/reference into oracle/_ref/ (git-ignored, NOT gpurun-ignored, so it travels to the GPU box next to the
built .so) and records their SHA-256 in oracle/_ref/MANIFEST.json.  Nothing under oracle/_ref is tracked or edited.

Users: `bench.py --impl reference` and bench.py's `cpu_baseline` leg instantiate `model.TransformerTTS(device='cpu')`
from here (kind "reference"); `oracle/make_golden.py` imports the same files straight from /root/reference.  The product
package never imports it (tests/test_abi.py checks).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
DST = os.path.join(HERE, "_ref")
FILES = ["model/__init__.py", "model/model.py", "model/layers.py", "model/module.py", "loss.py"]
# The reference's training harness, staged in its own directory (oracle/_ref/harness/) so that importing it never puts the
# staged `model` / `loss` on the path: tests/test_gpu_ref_harness.py runs this file UNCHANGED on the B200 module, with
# `model`, `loss`, `utils.util` resolved to transformertts_b200 and pytorch_lightning / loguru / utils.plot stubbed.
HARNESS_FILES = ["lightning_module.py"]
HARNESS_DST = os.path.join(DST, "harness")


def sha256(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def stage(src=REF_SRC, dst=DST):
    """Copy the hot-path files; returns the manifest dict.  Raises if the reference tree is absent."""
    if not os.path.isdir(src):
        raise FileNotFoundError(f"{src} not found: the staged copy can only be made in the build container")
    manifest = {"source": src, "files": {}}
    for rel in FILES:
        s, d = os.path.join(src, rel), os.path.join(dst, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        manifest["files"][rel] = sha256(d)
    manifest["harness_files"] = {}
    for rel in HARNESS_FILES:
        s, d = os.path.join(src, rel), os.path.join(dst, "harness", rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        manifest["harness_files"][rel] = sha256(d)
    with open(os.path.join(dst, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    return manifest


def load(dst=DST):
    """Import the staged reference: returns (TransformerTTS class, TransformerTTSLoss class) or None when oracle/_ref
    is absent or a file no longer matches its recorded hash."""
    mpath = os.path.join(dst, "MANIFEST.json")
    if not os.path.exists(mpath):
        return None
    with open(mpath) as f:
        manifest = json.load(f)
    for rel, digest in manifest["files"].items():
        p = os.path.join(dst, rel)
        if not os.path.exists(p) or sha256(p) != digest:
            return None
    if dst not in sys.path:
        sys.path.insert(0, dst)
    for name in [m for m in sys.modules if m == "model" or m.startswith("model.") or m == "loss"]:
        del sys.modules[name]                              # a different `model` package must not shadow the staged one
    from model.model import TransformerTTS                  # noqa: E402  (the reference)
    from loss import TransformerTTSLoss                     # noqa: E402
    return TransformerTTS, TransformerTTSLoss


def harness_path(name="lightning_module.py", dst=DST):
    """Path of a staged harness file if it is present and matches its recorded hash, else None."""
    mpath = os.path.join(dst, "MANIFEST.json")
    if not os.path.exists(mpath):
        return None
    with open(mpath) as f:
        digest = json.load(f).get("harness_files", {}).get(name)
    p = os.path.join(dst, "harness", name)
    if digest is None or not os.path.exists(p) or sha256(p) != digest:
        return None
    return p


if __name__ == "__main__":
    m = stage()
    print(json.dumps(m, indent=1))
