"""Build the CUDA library in-tree: parallelraytracing_b200/libprt_b200.so (sm_100a only)."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libprt_b200.so")
SOURCES = ["api.cu", "lbvh.cu", "wavefront.cu"]
HEADERS = ["prt_device.cuh", "prt_internal.h", os.path.join("..", "..", "include", "prt.h")]
NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-fmad=false",            # FP contract with the oracle: only explicit fmaf() fuses (Numerics)
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
]


def nvcc_path():
    for p in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if p and os.path.exists(p):
            return p
    raise RuntimeError("nvcc not found")


OUT_CHECKED = os.path.join(HERE, "libprt_b200_checked.so")


def source_hash(checked=False):
    """sha256 over the compiler flags and every source / header the library is built from."""
    import hashlib
    h = hashlib.sha256((" ".join(NVCC_FLAGS) + (" -DPRT_CHECKED=1" if checked else "")).encode())
    for d in SOURCES + HEADERS:
        with open(os.path.join(CSRC, d), "rb") as f:
            h.update(d.encode()); h.update(f.read())
    return h.hexdigest()


STAMP = OUT + ".srchash"


def needs_build(checked=False):
    """True when the library is missing or was built from other sources / flags than the ones in the tree (content hash,
    not mtimes: the tree is copied to the GPU box). checked=True: libprt_b200_checked.so (-DPRT_CHECKED=1, the bounds-checked
    build the tests run the small configurations through - compute-sanitizer is closed on the GPU pool)."""
    out = OUT_CHECKED if checked else OUT
    if not os.path.exists(out) or not os.path.exists(out + ".srchash"):
        return True
    with open(out + ".srchash") as f:
        return f.read().strip() != source_hash(checked)


def build(force=False, verbose=False, checked=False):
    out = OUT_CHECKED if checked else OUT
    if not force and not needs_build(checked):
        return out
    cmd = [nvcc_path()] + NVCC_FLAGS + (["-DPRT_CHECKED=1"] if checked else []) + (["-Xptxas", "-v"] if verbose else []) + ["-o", out] + SOURCES
    r = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if verbose:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    with open(out + ".srchash", "w") as f:
        f.write(source_hash(checked))
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build(force="--force" in sys.argv, checked=True))
