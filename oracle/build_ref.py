"""TEST INFRASTRUCTURE — copy recipe for the reference arm of bench.py (`--impl reference`, `cpu_baseline`).

The reference (xxxxx23124/-GAN-) is pure Python with no build system and no package metadata: there is
nothing to compile or pip-install.  `/root/reference` exists only in the build container, so this recipe copies its
`.py` files — unmodified, byte for byte — into `oracle/_ref/`, which is listed in .gitignore (it never enters the
history: no reference source is committed) but not in .gpurunignore (it travels to the GPU box like a built .so).
`oracle/reference.py` then loads the tree from there when /root/reference is absent, and bench.py times the
reference's own `train/<loop>.py: Train.discriminator_trainstep / generator_trainstep` on the box's host cores
(`cpu_baseline.kind` = "reference").  `__graft_entry__.build()` runs this when /root/reference is present.

    python oracle/build_ref.py [src=/root/reference] [dst=oracle/_ref]
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKGS = ("generators", "discriminators", "train", "units")


def build(src="/root/reference", dst=os.path.join(HERE, "_ref")):
    if not os.path.isdir(os.path.join(src, "generators")):
        return False
    manifest = {}
    for pkg in PKGS:
        for dirpath, _, files in os.walk(os.path.join(src, pkg)):
            for f in sorted(files):
                if not f.endswith(".py"):
                    continue
                a = os.path.join(dirpath, f)
                rel = os.path.relpath(a, src)
                b = os.path.join(dst, rel)
                os.makedirs(os.path.dirname(b), exist_ok=True)
                data = open(a, "rb").read()
                if not (os.path.exists(b) and open(b, "rb").read() == data):
                    shutil.copyfile(a, b)
                manifest[rel] = hashlib.sha256(data).hexdigest()
    json.dump(manifest, open(os.path.join(dst, "MANIFEST.json"), "w"), indent=0, sort_keys=True)
    return True


if __name__ == "__main__":
    ok = build(*sys.argv[1:3])
    print("oracle/_ref:", "copied" if ok else "reference tree not found, nothing done")
