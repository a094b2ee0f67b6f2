"""Build a variant of libmxrcnn_b200.so with extra -D defines into .ab/<name>.so (travels to the GPU box, git-ignored);
run anything against it with MXRCNN_B200_LIB=.ab/<name>.so.   python tools/ab_build.py <name> [-DX=1 ...]"""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxrcnn_b200 import build as b  # noqa: E402

name, defs = sys.argv[1], sys.argv[2:]
out = os.path.join(b.ROOT, ".ab", name + ".so")
os.makedirs(os.path.dirname(out), exist_ok=True)
srcs = [os.path.join(b.CSRC, s) for s in b.SOURCES]
cmd = [b._nvcc()] + b.NVCC_FLAGS + defs + ["-o", out] + srcs
r = subprocess.run(cmd, capture_output=True, text=True)
if r.returncode != 0:
    raise SystemExit(r.stderr[-4000:])
print(out)
