#!/usr/bin/env python
"""Builds variants of libdeepore_b200.so that differ only in compile-time switches of the inflate kernel (csrc/bgzf.cu,
csrc/inflate_core.h) into variants/ (travels to the GPU box, git-ignored), for scripts/bgzf_bench.py --lib.
    python scripts/inflate_variants.py name=-DFLAG,-DFLAG2 ..."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepore_b200 import build          # noqa: E402

build.build()
out = os.path.join(ROOT, "variants")
os.makedirs(out, exist_ok=True)
bdir = os.path.join(build.HERE, "build")
others = [os.path.join(bdir, os.path.basename(s)[:-3] + ".o") for s in build.sources() if not s.endswith("bgzf.cu")]
for spec in sys.argv[1:]:
    name, _, flags = spec.partition("=")
    obj = os.path.join(out, name + ".o")
    cmd = ["nvcc"] + build.NVCC_FLAGS + [f for f in flags.split(",") if f] + ["-c", os.path.join(build.CSRC, "bgzf.cu"), "-o", obj]
    log = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, check=True).stdout
    regs = [l for l in log.splitlines() if "bgzf_inflate" in l or ("Used" in l)]
    k = [i for i, l in enumerate(log.splitlines()) if "bgzf_inflate_kernel" in l and "Compiling" in l]
    info = log.splitlines()[k[0] + 2:k[0] + 4] if k else []
    subprocess.check_call(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", os.path.join(out, "lib_%s.so" % name), obj]
                          + others + ["-lcudart", "-lcuda"])
    os.remove(obj)
    print(name, flags, " ".join(x.strip() for x in info))
