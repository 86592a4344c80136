"""Stages the reference's own attack.py for the timed CPU arm (bench.py --impl reference / cpu_baseline, kind "reference").

The reference is pure Python with nothing to compile, so "building oracle/_ref" is copying the one unmodified source
file the arm executes from where it lies under /root/reference into oracle/_ref/ (git-ignored, not gpurun-ignored: it
travels to the GPU box with the snapshot, the repository history never holds it).  A no-op when /root/reference is
absent (on the GPU box the staged file is already there; without it bench.py falls back to the oracle port)."""
from __future__ import annotations

import hashlib
import shutil
from pathlib import Path

REF = Path("/root/reference/attack.py")
DST = Path(__file__).resolve().parent / "_ref"


def stage() -> bool:
    if not REF.is_file():
        return (DST / "attack.py").is_file()
    DST.mkdir(exist_ok=True)
    shutil.copyfile(REF, DST / "attack.py")
    (DST / "attack.py.sha256").write_text(hashlib.sha256(REF.read_bytes()).hexdigest() + "  /root/reference/attack.py\n")
    return True


if __name__ == "__main__":
    print("staged" if stage() else "reference not available")
