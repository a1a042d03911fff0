"""Builds the CUDA library of this package in-tree with nvcc for sm_100a (no torch extension machinery:
the library is a plain C-ABI shared object, include/strus_pattern_b200.h)."""
import os
import shutil
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libstrus_pattern_b200.so")
# the artefact names of the reference build (src/CMakeLists.txt:48-57, 72-74): SONAME 0.17 of libstrus_pattern and the
# analyzer module that strusModule's loader opens; both are this library under the names a strus installation looks for
ALIASES = ["libstrus_pattern.so.0.17", "modstrus_analyzer_pattern.so"]
SOURCES = ["lexer_capi.cu", "matcher_capi.cu", "strus_iface.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=default", "-shared", "-cudart", "shared"]


def nvcc_path():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def up_to_date():
    if not os.path.exists(LIB):
        return False
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "strus_pattern_b200.h")]
    return all(os.path.getmtime(d) <= t for d in deps)


def link_aliases():
    for name in ALIASES:
        path = os.path.join(PKG, name)
        if os.path.lexists(path):
            os.unlink(path)
        os.symlink(os.path.basename(LIB), path)


def build(force=False, verbose=False):
    """Compile every CUDA source into libstrus_pattern_b200.so (skipped when it is newer than its sources)."""
    if not force and up_to_date():
        if not all(os.path.lexists(os.path.join(PKG, a)) for a in ALIASES):
            link_aliases()
        return LIB
    nvcc = nvcc_path()
    if nvcc is None:
        if os.path.exists(LIB):
            return LIB  # GPU box without toolkit on PATH: use the shipped build
        raise RuntimeError("nvcc not found and no prebuilt %s" % LIB)
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [
        "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-o", LIB] + sources()
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout)
    if verbose:
        print(res.stdout)
    link_aliases()
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
