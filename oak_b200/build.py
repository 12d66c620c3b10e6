"""In-tree build of the CUDA library (sm_100a only). `python -m oak_b200.build` or build()."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "liboakb200.so")
# -jump-table-density=10: the engine's dispatches on a move's effect (35 handlers) become indexed branches (BRX) instead of
# the compare-and-branch trees cicc builds by default, which every divergent group of lanes walked on its own
# (2.7 % of the playout kernel's instructions): 3.29 -> 3.57 G battle-steps/s through the host API (profiles/r2p_*).
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-mavx2", "-Xcicc", "-jump-table-density=10", "-shared"]


def sources():
    return [os.path.join(CSRC, "oakb200.cu"), os.path.join(CSRC, "value_net.cu"), os.path.join(CSRC, "frame_log.cu"),
            os.path.join(CSRC, "pkmn_compat.cu"), os.path.join(CSRC, "search.cu"), os.path.join(CSRC, "train_rows.cu")]


def _deps():
    out = [os.path.join(HERE, "..", "include", "oakb200.h"), os.path.join(HERE, "..", "include", "oakb200_pkmn.h")]
    for f in os.listdir(CSRC):
        out.append(os.path.join(CSRC, f))
    return out


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in _deps())


def build(force=False, verbose=False, out=None, extra=None):
    """Compile oak_b200/liboakb200.so with nvcc for sm_100a (one object per source, compiled in parallel, then linked).
    Returns the library path. `out` / `extra`: a differently named library with extra nvcc flags (A/B variants)."""
    lib = out or LIB
    if out is None and not force and not stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    srcs = [s for s in sources() if os.path.exists(s)]
    extra = list(extra or []) + os.environ.get("OAKB200_NVCC_EXTRA", "").split()  # diagnostic builds (e.g. -DOAKB200_TC_TIMING)
    objdir = os.path.join(HERE, "_build", os.path.splitext(os.path.basename(lib))[0])
    os.makedirs(objdir, exist_ok=True)
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + extra + (["-Xptxas", "-v"] if verbose else [])

    def compile_one(src):
        obj = os.path.join(objdir, os.path.splitext(os.path.basename(src))[0] + ".o")
        cmd = [nvcc] + flags + ["-c", "-o", obj, src]
        print("[oak_b200.build]", " ".join(cmd), file=sys.stderr)
        subprocess.check_call(cmd)
        return obj
    with ThreadPoolExecutor(max_workers=len(srcs)) as pool:
        objs = list(pool.map(compile_one, srcs))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib] + objs
    print("[oak_b200.build]", " ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return lib


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
