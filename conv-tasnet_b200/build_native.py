"""Builds libctn_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python conv-tasnet_b200/build_native.py [--force] [--verbose] [--fp16 | --all]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels to the
GPU box with the repo snapshot.
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, os.environ.get("CTN_LIB_NAME", "libctn_b200.so"))   # CTN_LIB_NAME / NVCC_FLAGS: experiment builds
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


# Variants of the library: name -> (file, extra nvcc flags).  "fp16" is the opt-in inference precision of
# DESIGN.md section 8 item 0 (fp16 operands, one MMA per product), selected per process with CTN_PRECISION=fp16.
VARIANTS = {
    None: ("libctn_b200.so", []),
    "fp16": ("libctn_b200_fp16.so", ["-DCTN_OPERAND_FP16=1"]),
}


def _stale(lib):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, variant=None):
    name, vflags = VARIANTS[variant]
    lib = LIB if variant is None else os.path.join(HERE, name)
    if not force and not _stale(lib):
        return lib
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in sources():
        obj = os.path.join(HERE, "build", os.path.basename(lib)[:-3] + "." + os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc, *ARCH, "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
               "--expt-relaxed-constexpr", *vflags, *os.environ.get("NVCC_FLAGS", "").split(), "-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write("== %s ==\n%s\n" % (os.path.basename(src), out))
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed building %s" % os.path.basename(lib))
    link = [nvcc, *ARCH, "-shared", "-o", lib, *objs, "-lcudart_static", "-ldl", "-lpthread", "-lrt"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout)
    return lib


if __name__ == "__main__":
    variants = [None, "fp16"] if "--all" in sys.argv else ["fp16"] if "--fp16" in sys.argv else [None]
    for v in variants:
        print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, variant=v))
