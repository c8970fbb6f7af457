#!/usr/bin/env python3
"""oracle/build_ref.py — TEST INFRASTRUCTURE.  Builds oracle/_ref/ from the reference where it lies.

Nothing under /root/reference is copied into the repository: the generated intermediates and the
binaries go to oracle/_ref/ only (git-ignored, NOT gpurun-ignored, so the .so files travel to the GPU
box like the product's own).  The reference's own build system (Visual Studio .vcxproj) is not run.

Outputs:
  libref_host.so                 host functions of IS/main.cpp + hdrloader.cpp  (see ref_host_wrap.cpp)
  libref_shader_<v>_<math>.so    the reference pass1.fsh of variant v in {ogl, disney, is} compiled as
                                 C++ (see glsl_compat.h / ref_shader_wrap.cpp); math in {libm, ptmath} are the
                                 parity builds (-O2 -ffp-contract=off); math = libmO3 is the TIMING build of SURVEY.md 8d
                                 (g++ -O3 -fopenmp, libm) that bench.py --impl reference runs
  libref_pass3.so                shaders/pass3.fsh (the display pass) compiled the same way
  basic_tracer                   BasicRayTracingWithC++/main.cpp, g++ -O3 -fopenmp, SAMPLE overridable
                                 through the REF_SAMPLE macro (the only edit: `const int SAMPLE = 128`)
  libref_basic_det.so            the same source with a reproducible random stream, as a library (ref_basic_wrap.cpp)
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PT_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")
INC = os.path.join(os.path.dirname(HERE), "include")
IS_DIR = "ImportanceSampling_LowDiscrepancySequence"
SHADERS = {
    "ogl": (0, "OpenglRayTracing/shaders/pass1.fsh"),
    "disney": (1, "DisneyBRDF/shaders/pass1.fsh"),
    "is": (2, IS_DIR + "/shaders/pass1.fsh"),
}
CXXFLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-w"]


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def newer(target, *sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


# ---------------------------------------------------------------- shader text -> C++ struct body
_FLOAT_LIT = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")


def shader_to_cpp(src: str, pass1: bool = True) -> str:
    src = src.replace("\r\n", "\n")
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)              # block comments
    src = re.sub(r"//[^\n]*", "", src)                            # line comments (GBK bytes, may end in '\')
    src = re.sub(r"^\s*#version[^\n]*\n", "\n", src, flags=re.M)
    src = _FLOAT_LIT.sub(lambda m: m.group(1) + "f", src)        # GLSL float literals are 32-bit
    for sw in ("xyz", "rgb", "rg", "xy"):                         # multi-component swizzles -> methods
        src = re.sub(r"\.%s\b(?!\()" % sw, ".%s()" % sw, src)
    src = re.sub(r"\binout\s+(\w+)\s+", r"\1& ", src)            # inout T x -> T& x
    if not pass1:
        return src
    src = src.replace("const uint V[8*32] = {", "static constexpr uint V[8*32] = {")
    # hitBVH's `int stack[256]` (pass1.fsh:340) is overrun by the reference's own SAH trees on large meshes (chains
    # thousands of nodes deep); on a GPU that is undefined behaviour, on the CPU it would smash the C++ stack.
    # The CPU build gets room for the deepest tree, so those scenes stay defined; nothing else changes.
    assert src.count("int stack[256];") == 1
    src = src.replace("int stack[256];", "int stack[16384];")
    # GLSL evaluates argument lists left to right; C++ leaves it open.  Two call sites have
    # side-effecting arguments (rand() twice): brace-initialisation is left-to-right in C++.
    aa_old = "vec2((rand()-0.5f)/float(width), (rand()-0.5f)/float(height))"
    assert src.count(aa_old) == 1, "AA jitter call site not found"
    src = src.replace(aa_old, "vec2{(rand()-0.5f)/float(width), (rand()-0.5f)/float(height)}")
    src = src.replace("SampleHdr(rand(), rand())", "SampleHdr(vec2{rand(), rand()})")
    return src


def build_shaders():
    wrap = os.path.join(HERE, "ref_shader_wrap.cpp")
    compat = os.path.join(HERE, "glsl_compat.h")
    for name, (variant, rel) in SHADERS.items():
        path = os.path.join(REF, rel)
        inc = os.path.join(OUT, "shader_%s.inc" % name)
        with open(path, "r", encoding="latin-1") as f:
            text = shader_to_cpp(f.read())
        with open(inc, "w", encoding="latin-1") as f:
            f.write(text)
        for math in ("libm", "ptmath", "libmO3"):
            so = os.path.join(OUT, "libref_shader_%s_%s.so" % (name, math))
            if newer(so, path, wrap, compat, os.path.join(INC, "ptmath.h"), __file__):
                continue
            flags = ["-std=c++17", "-O3", "-fopenmp", "-fPIC", "-w"] if math == "libmO3" else CXXFLAGS
            cmd = ["g++", *flags, "-shared", "-I", HERE, "-I", INC,
                   '-DREF_SHADER_INC="%s"' % inc, "-DREF_VARIANT=%d" % variant]
            if math == "ptmath":
                cmd.append("-DREF_MATH_PTMATH")
            run(cmd + [wrap, "-o", so])


def build_pass3():
    """The display pass (toneMapping, pass3.fsh:14-24) compiled from the reference's text: oracle of pt_readback_display."""
    path = os.path.join(REF, IS_DIR, "shaders", "pass3.fsh")
    wrap = os.path.join(HERE, "ref_pass3_wrap.cpp")
    inc = os.path.join(OUT, "pass3.inc")
    so = os.path.join(OUT, "libref_pass3.so")
    if newer(so, path, wrap, os.path.join(HERE, "glsl_compat.h"), __file__):
        return
    with open(path, "r", encoding="latin-1") as f:
        text = shader_to_cpp(f.read(), pass1=False)
    with open(inc, "w", encoding="latin-1") as f:
        f.write(text)
    run(["g++", *CXXFLAGS, "-shared", "-I", HERE, "-I", INC, '-DREF_PASS3_INC="%s"' % inc, wrap, "-o", so])


# ---------------------------------------------------------------- host functions of IS/main.cpp
HOST_RANGES = [  # (first, last, token that must appear on the first line)
    (19, 21, "#define INF"),
    (26, 72, "struct Material"),
    (153, 167, "bool cmpx"),
    (243, 259, "mat4 getTransformMatrix"),
    (262, 373, "void readObj"),
    (377, 552, "int buildBVH"),
    (555, 652, "float* calculateHdrCache"),
]


def build_host():
    main_cpp = os.path.join(REF, IS_DIR, "main.cpp")
    with open(main_cpp, "r", encoding="latin-1") as f:
        lines = f.read().replace("\r\n", "\n").split("\n")
    parts = []
    for a, b, tok in HOST_RANGES:
        assert tok in lines[a - 1], "reference layout changed: line %d lacks %r" % (a, tok)
        chunk = [re.sub(r"//[^\n]*", "", ln) for ln in lines[a - 1:b]]
        chunk = [ln for ln in chunk if not ln.lstrip().startswith("#include")]
        parts.append("\n".join(chunk))
    inc = os.path.join(OUT, "ref_host_extract.inc")
    with open(inc, "w", encoding="latin-1") as f:
        f.write("\n\n".join(parts) + "\n")
    hdr_src = os.path.join(REF, IS_DIR, "hdrloader.cpp")
    with open(hdr_src, "r", encoding="latin-1") as f:
        h = f.read().replace("\r\n", "\n")
    assert h.count('"-Y %ld +X %ld"') == 1
    h = h.replace('"-Y %ld +X %ld"', '"-Y %d +X %d"')            # int targets: the LP64 fix
    hdr_patched = os.path.join(OUT, "hdrloader_patched.cpp")
    with open(hdr_patched, "w", encoding="latin-1") as f:
        f.write(h)
    so = os.path.join(OUT, "libref_host.so")
    wrap = os.path.join(HERE, "ref_host_wrap.cpp")
    if newer(so, main_cpp, hdr_src, wrap, __file__):
        return
    run(["g++", *CXXFLAGS, "-shared", "-I", OUT, "-I", os.path.join(REF, "include"),
         "-I", os.path.join(REF, IS_DIR), wrap, hdr_patched, "-o", so])


# ---------------------------------------------------------------- BasicRayTracingWithC++ CPU tracer
def build_basic():
    src = os.path.join(REF, "BasicRayTracingWithC++", "main.cpp")
    exe = os.path.join(OUT, "basic_tracer")
    if newer(exe, src, __file__):
        return
    with open(src, "r", encoding="latin-1") as f:
        t = f.read().replace("\r\n", "\n")
    old = "const int SAMPLE = 128;"
    assert t.count(old) == 1
    t = t.replace(old, "#ifndef REF_SAMPLE\n#define REF_SAMPLE 128\n#endif\nconst int SAMPLE = REF_SAMPLE;")
    gen = os.path.join(OUT, "basic_main.cpp")
    with open(gen, "w", encoding="latin-1") as f:
        f.write(t)
    shim = ["-include", "cstring", "-include", "cstdio",
            "-Dfopen_s(pf,n,m)=((*(pf)=fopen(n,m))==0)"]
    for spp in (128, 200):
        run(["g++", "-std=c++17", "-O3", "-fopenmp", "-w", "-DREF_SAMPLE=%d" % spp, *shim,
             "-I", os.path.join(REF, "include"), gen, "-o", exe + ("_%dspp" % spp)])
    run(["cp", exe + "_200spp", exe])


def build_basic_det():
    """libref_basic_det.so: the same source with a reproducible random stream (oracle/ref_basic_wrap.cpp says what is edited)."""
    src = os.path.join(REF, "BasicRayTracingWithC++", "main.cpp")
    wrap = os.path.join(HERE, "ref_basic_wrap.cpp")
    so = os.path.join(OUT, "libref_basic_det.so")
    if newer(so, src, wrap, __file__):
        return
    with open(src, "r", encoding="latin-1") as f:
        t = f.read().replace("\r\n", "\n")

    def sub1(old, new, start=0):
        nonlocal t
        i = t.find(old, start)
        assert i >= 0 and t.find(old, i + 1) < 0 or start > 0, old
        t = t[:i] + new + t[i + len(old):]

    sub1("const int SAMPLE = 128;", "static int SAMPLE = 128;")
    sub1("const double BRIGHTNESS = (2.0f * 3.1415926f) * (1.0f / double(SAMPLE));",
         "#define BRIGHTNESS ((2.0f * 3.1415926f) * (1.0f / double(SAMPLE)))")
    sub1("const int WIDTH = 256;", "static int WIDTH = 256;")
    sub1("const int HEIGHT = 256;", "static int HEIGHT = 256;")
    sub1("return dis(gen);", "return det_randf();")
    sub1('#include "svpng.inc"', "static void svpng(FILE*, unsigned, unsigned, const unsigned char*, int) {}")
    m = t.find("int main()")
    assert m >= 0
    sub1("int main()", "static int ref_basic_main()")
    sub1("omp_set_num_threads(50);", "omp_set_num_threads(1);")
    loop = "for (int j = 0; j < WIDTH; j++)\n            {"
    j = t.find(loop, m)
    assert j >= 0
    t = t[:j + len(loop)] + "\n                det_seed(k, i, j);" + t[j + len(loop):]
    sub1('imshow(image, "basic_ray_tracing_with_c++.png");', "det_export(image);")
    with open(os.path.join(OUT, "basic_det.inc"), "w", encoding="latin-1") as f:
        f.write(t)
    run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-w", "-shared",
         "-include", "cstring", "-include", "cstdio", "-Dfopen_s(pf,n,m)=((*(pf)=fopen(n,m))==0)",
         "-I", OUT, "-I", os.path.join(REF, "include"), wrap, "-o", so])


def main():
    if not os.path.isdir(REF):
        print("build_ref: %s not present (GPU box?) — using prebuilt oracle/_ref as is" % REF)
        return 0
    os.makedirs(OUT, exist_ok=True)
    build_host()
    build_shaders()
    build_pass3()
    if "--no-basic" not in sys.argv:
        build_basic()
        build_basic_det()
    return 0


if __name__ == "__main__":
    sys.exit(main())
