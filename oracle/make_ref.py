"""Recipe: stage the reference's OWN files for the hot path under oracle/_ref/ so that they travel to the GPU box.

TEST INFRASTRUCTURE (see oracle/__init__.py).  The reference is pure Python: there is nothing to compile, so "building"
`oracle/_ref` is a byte-for-byte copy of the few files the path lives in — `fly.py` (compute_fly_observations,
compute_fly_reward2), `flyLowGrav.py` (the low-gravity twin), `ppo.py` (Net, PPO.run / make_data / update) and the one
checkpoint the golden vectors use — from where they lie under /root/reference.  `oracle/_ref/` is git-ignored (no
reference source ever enters the history) but NOT gpurun-ignored, so the snapshot carries it and `bench.py --impl
reference` / `cpu_baseline` can time the reference's unmodified code on the GPU box's host cores
(`cpu_baseline.kind = "reference"`), and `tests/test_oracle_vs_reference.py` can re-derive the oracle's pin there.

    python oracle/make_ref.py            # run in the build container; __graft_entry__.build() calls it too

Nothing here is imported by the product package.  MANIFEST.json records the sha256 of every staged file.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("FLYB200_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = ["fly.py", "flyLowGrav.py", "ppo.py", "trainer.py", "saves/save7_1_23/save1_standing_18dofs_7800.pth"]


def sha256(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def stage(verbose: bool = True) -> bool:
    """Copy FILES from SRC to DST unmodified.  Returns False (and stages nothing) when SRC does not exist."""
    if not os.path.isfile(os.path.join(SRC, "fly.py")):
        if verbose:
            print(f"make_ref: {SRC} not present; keeping whatever oracle/_ref already holds")
        return False
    manifest = {"source": SRC, "files": {}}
    for rel in FILES:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not (os.path.exists(dst) and sha256(dst) == sha256(src)):
            if os.path.exists(dst):
                os.chmod(dst, 0o644)
            shutil.copyfile(src, dst)
            os.chmod(dst, 0o644)
        manifest["files"][rel] = sha256(dst)
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    if verbose:
        print(f"make_ref: staged {len(FILES)} reference files under {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
