"""Build the in-tree C-ABI library for sm_100a with nvcc (no torch headers involved).

    python splat-loam_b200/build.py [--force]

Output: splat-loam_b200/lib/libsplat_b200.so (git-ignored; travels to the GPU box).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libsplat_b200.so")
SOURCES = [
    "dssr_api.cu",
    "dssr_preprocess.cu",
    "radix_sort.cu",
    "dssr_render_fwd.cu",
    "dssr_render_bwd.cu",
    "sknn.cu",
    "peer_exchange.cu",
    "render_post.cu",
    "adam_step.cu",
    "state_surgery.cu",
    "photometric.cu",
    "range_image.cu",
]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-lineinfo",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    # no --use_fast_math: the reference builds without it and parity depends on
    # IEEE division / sqrt and the accurate expf / atan2f / fmodf
]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "splat_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        objs.append(obj)
        # SPLAT_B200_EXTRA_NVCC_FLAGS: developer builds (-DRENDER_TRACE, -DSORT_TRACE)
        extra = os.environ.get("SPLAT_B200_EXTRA_NVCC_FLAGS", "").split()
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    failed = False
    for src, p in procs:
        out = p.communicate()[0].decode()
        if out.strip():
            print(f"--- {src}\n{out}", file=sys.stderr)
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
                           *objs, "-o", LIB, "-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
