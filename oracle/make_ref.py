"""TEST / BASELINE INFRASTRUCTURE — stages the reference's own hot-path files under ``oracle/_ref/``.

``/root/reference`` exists only in the build container; the GPU box sees what travels with the repo snapshot.
``oracle/_ref/`` is git-ignored (the history holds no reference source) but NOT gpurun-ignored, so the staged files
reach the GPU box, where ``bench.py --impl reference`` and ``cpu_baseline`` execute them UNMODIFIED through
``oracle/ref_shim.py`` — the reference arm is then the reference's file, not a restatement.

    python oracle/make_ref.py          (also run by ``__graft_entry__.build()`` whenever /root/reference is present)

Staged, byte for byte (a manifest with SHA-256 sums is written next to them):
  pyscf/mpcc/dfrmpcc_lowlevel.py   the low-level solver (the hot path, :72-176)
  pyscf/mpcc/mpcc.py               the solver object that takes a low-level *module* (:5-17)
  pyscf/mpcc/dfrmpcc.py            RMPCC (:5-12)
  pyscf/mpcc/__init__.py           factory (:3-17)
  pyscf/lib/diis.py                lib.diis.DIIS (:100-275)
Nothing under ``pyscf_mpcc_b200/`` reads this directory.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SOURCE_ROOT = "/root/reference"
STAGE_ROOT = os.path.join(HERE, "_ref")
FILES = (
    "pyscf/mpcc/dfrmpcc_lowlevel.py",
    "pyscf/mpcc/mpcc.py",
    "pyscf/mpcc/dfrmpcc.py",
    "pyscf/mpcc/__init__.py",
    "pyscf/lib/diis.py",
)


def stage(source_root: str = SOURCE_ROOT, stage_root: str = STAGE_ROOT) -> dict:
    """Copy FILES from the reference tree; returns {relative path: sha256}. Raises if the tree is absent."""
    manifest = {}
    for rel in FILES:
        src = os.path.join(source_root, rel)
        if not os.path.isfile(src):
            raise FileNotFoundError(src)
        dst = os.path.join(stage_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        with open(dst, "rb") as f:
            manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(stage_root, "MANIFEST.json"), "w") as f:
        json.dump({"source": source_root, "files": manifest}, f, indent=1)
    return manifest


def staged() -> bool:
    return all(os.path.isfile(os.path.join(STAGE_ROOT, rel)) for rel in FILES)


if __name__ == "__main__":
    if not os.path.isdir(SOURCE_ROOT):
        print(f"{SOURCE_ROOT} is absent; staged copy present: {staged()}")
        sys.exit(0 if staged() else 1)
    for rel, digest in stage().items():
        print(f"{digest[:16]}  {rel}")
