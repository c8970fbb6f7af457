"""Builds libpt_<name>.so next to libpt.so with extra -D macros on the two kernel translation units (A/B
measurements of kernel variants in ONE gpurun call: PT_LIBPT=<path> selects the build at run time).

    python tools/build_variant.py oldfold -DPT_FOLD_TAGGED=0
"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_ray_tracing_b200 import build as B

name, defs = sys.argv[1], sys.argv[2:]
fast_only = "--fast-only" in defs              # perf A/B in PT_MODE_FAST: reuse the default build's parity object
defs = [d for d in defs if d not in ("--fast-only", "--no-ftz")]
if "--no-base" not in defs:
    B.build_pt()
defs = [d for d in defs if d != "--no-base"]
fast_flags = ["-prec-div=false", "-prec-sqrt=false"]
for d in list(defs):
    if d.startswith("--fast-flags="):         # replaces the production TU's floating-point flags, e.g. --fast-flags=-fmad=false
        fast_flags = d.split("=", 1)[1].split()
        defs.remove(d)
vdir = os.path.join(B.BUILD, "variant_" + name)
os.makedirs(vdir, exist_ok=True)
objs = []
for unit, extra in (("pt_kernels_parity", ["-fmad=false"]), ("pt_kernels_fast", fast_flags), ("pt_kernels_fast_v2", fast_flags + ([] if "--no-ftz" in sys.argv else ["-ftz=true"]))):
    if fast_only and unit == "pt_kernels_parity":
        objs.append(os.path.join(B.BUILD, unit + ".o"))
        continue
    obj = os.path.join(vdir, unit + ".o")
    B._run(["nvcc", *B.ARCH, *B.NVCC_COMMON, *extra, *defs, "-c", os.path.join(B.CSRC, unit + ".cu"), "-o", obj], log=os.path.join(vdir, unit + ".ptxas.log"))
    objs.append(obj)
objs.append(B._sass_hash_object(objs[:3], force=True) if not fast_only else os.path.join(B.BUILD, "pt_sasshash.o"))
objs += [os.path.join(B.BUILD, u + ".o") for u in ("pt_api", "pt_hdrcache", "pt_basic", "pt_probe", "pt_comm", "pt_build")]
out = os.path.join(B.HERE, "libpt_%s.so" % name)
B._run(["nvcc", *B.ARCH, "-shared", "-Xcompiler", "-fPIC", *objs, "-lgomp", "-ldl", "-o", out])
print(out)
