"""Build liblt_b200.so (nvcc, sm_100a only) in-tree.   python -m legged_tracking_b200.build [--force]

nvcc cross-compiles without a GPU, so this also is the "does it build" check of __graft_entry__.build().
"""
import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
OUT = os.path.join(PKG_DIR, "liblt_b200.so")
SOURCES = ["lt_kernels.cu", "lt_torque_mma.cu", "lt_torque_tc5.cu", "lt_rollout.cu", "lt_terrain.cu", "lt_policy.cu", "lt_planner.cu", "lt_gait.cu", "lt_vel.cu"]
HEADERS = ["lt_core.cuh", "lt_torque_mma.h"]
# per-source header dependencies (a change to the GEMM header must not recompile the 3-minute step kernels)
DEPS = {"lt_policy.cu": [os.path.join(CSRC, "lt_gemm.cuh"), os.path.join(REPO, "include", "lt_policy.h"),
                         os.path.join(REPO, "include", "lt_b200.h")]}
for _src in ("lt_gait.cu", "lt_vel.cu"):          # the velocity-env sources also share lt_cmd.cuh
    DEPS[_src] = [os.path.join(CSRC, h) for h in HEADERS + ["lt_cmd.cuh"]] + [os.path.join(REPO, "include", "lt_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default", "-I", os.path.join(REPO, "include")]
OBJ_DIR = os.path.join(REPO, "build", "obj")


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), out=OUT, only=None):
    """Each .cu is compiled to its own object (in parallel; only stale ones) and the objects are linked into `out`.
    defines / out / only: tuning variants (profiles/variants.py); the product build uses none of them.  `only` names the
    sources the defines apply to -- the other objects are taken from the default build."""
    from concurrent.futures import ThreadPoolExecutor
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.join(REPO, "include", "lt_b200.h")]
    tag = "default" if not defines else "v_" + "_".join(d.replace("=", "-") for d in defines)
    odir = os.path.join(OBJ_DIR, tag)
    os.makedirs(odir, exist_ok=True)
    objs = [os.path.join(odir if (only is None or os.path.basename(s) in only) else os.path.join(OBJ_DIR, "default"),
                         os.path.splitext(os.path.basename(s))[0] + ".o") for s in srcs]
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    log = []

    def compile_one(so):
        src, obj = so
        mine = only is None or os.path.basename(src) in only
        if not mine and os.path.exists(obj):
            return
        if not force and not _stale(obj, [src] + DEPS.get(os.path.basename(src), hdrs)):
            return
        cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in (defines if mine else ())] + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        log.append(r.stdout + r.stderr)
    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        list(ex.map(compile_one, zip(srcs, objs)))
    if force or _stale(out, objs):
        cmd = [nvcc, "--shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose:
        print("".join(log))
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
