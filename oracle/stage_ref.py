"""Stage the reference's own CPU implementation of the path into oracle/_ref/ (git-ignored; travels to the GPU box
with the snapshot like a built .so).  TEST / BASELINE INFRASTRUCTURE - never imported by the product.

    python oracle/stage_ref.py            (also run by __graft_entry__.build() when /root/reference is present)

The reference is a Python package: "building" it means making its pure-torch algorithm sub-package importable
where /root/reference does not exist.  The files are copied byte for byte (sha256 recorded in MANIFEST.json), only
rvc/lib/algorithm/** - the decoder (generators/hifigan_nsf.py, hifigan.py, residuals.py, commons.py, modules.py) and
the rest of Synthesizer (synthesizers.py, encoders.py, attentions.py, normalization.py, discriminators.py and the
sibling generators), none of which import anything outside torch / numpy.  Nothing of it is committed to git.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference"
DST = os.path.join(HERE, "_ref")
PKG_DIRS = ["rvc", "rvc/lib", "rvc/lib/algorithm", "rvc/lib/algorithm/generators"]


def stage(src: str = SRC, dst: str = DST) -> bool:
    if not os.path.isdir(os.path.join(src, "rvc", "lib", "algorithm")):
        return False
    manifest = {}
    for d in PKG_DIRS:
        os.makedirs(os.path.join(dst, d), exist_ok=True)
        init = os.path.join(src, d, "__init__.py")
        if os.path.exists(init):
            shutil.copyfile(init, os.path.join(dst, d, "__init__.py"))
        else:
            open(os.path.join(dst, d, "__init__.py"), "a").close()
    for d in ("rvc/lib/algorithm", "rvc/lib/algorithm/generators"):
        for f in sorted(os.listdir(os.path.join(src, d))):
            if not f.endswith(".py"):
                continue
            a, b = os.path.join(src, d, f), os.path.join(dst, d, f)
            shutil.copyfile(a, b)
            manifest[os.path.join(d, f)] = hashlib.sha256(open(a, "rb").read()).hexdigest()
    # the real RMVPE contour of the README clip (BASELINE.md section 3 times the reference on it)
    f0 = os.path.join(src, "ios_test_data", "rmvpe_f0.npy")
    if os.path.exists(f0):
        shutil.copyfile(f0, os.path.join(dst, "rmvpe_f0.npy"))
    with open(os.path.join(dst, "MANIFEST.json"), "w") as fh:
        json.dump({"source": src, "files": manifest}, fh, indent=1)
    return True


def ref_path():
    """Directory to put on sys.path to import the reference's ``rvc.lib.algorithm``: the live checkout when present
    (build container), else the staged copy; None when neither exists."""
    if os.path.isdir(os.path.join(SRC, "rvc", "lib", "algorithm")):
        return SRC
    if os.path.isfile(os.path.join(DST, "rvc", "lib", "algorithm", "generators", "hifigan_nsf.py")):
        return DST
    return None


if __name__ == "__main__":
    ok = stage()
    print("staged" if ok else "reference checkout not found", "->", DST)
    sys.exit(0 if ok else 1)
