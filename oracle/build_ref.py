"""TEST / BASELINE INFRASTRUCTURE ONLY -- recipe for oracle/_ref/.

Byte-compiles the UNMODIFIED reference hot-path module where it lies, /root/reference/alexnet_resnet.py,
into oracle/_ref/alexnet_resnet.bin (CPython byte code; git-ignored output; no reference source enters the repository).  The
file travels to the GPU box with the snapshot, where /root/reference does not exist, so that bench.py's
`--impl reference` arm and its `cpu_baseline` leg time the reference's OWN code (`kind: "reference"`)
instead of the restatement in oracle/ref_path.py.  Run by __graft_entry__.build() when /root/reference is
present; same interpreter version on both machines (byte code is tied to it).
"""
from __future__ import annotations

import os
import py_compile

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/alexnet_resnet.py"
OUT = os.path.join(HERE, "_ref", "alexnet_resnet.bin")      # CPython byte code (.pyc content; snapshots skip *.pyc)


def build() -> str | None:
    if not os.path.exists(SRC):
        return OUT if os.path.exists(OUT) else None
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < os.path.getmtime(SRC):
        py_compile.compile(SRC, cfile=OUT, doraise=True)
    return OUT


if __name__ == "__main__":
    print(build())
