"""In-tree build of libhexgnn.so (sm_100a only) with nvcc.  `python -m gnn_hex_b200.build`."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# HEXGNN_LIB: build / load another file instead (e.g. a -DHEXGNN_PHASES profiling build beside the product build)
LIB = os.environ.get("HEXGNN_LIB") or os.path.join(HERE, "libhexgnn.so")
SOURCES = ["api.cu", "csr_build.cu", "pack.cu", "sage_fwd_simt.cu", "sage_bwd_simt.cu", "tc_selftest.cu", "pack_tc.cu", "sage_fwd_tc.cu", "sage_bwd_tc.cu", "sage_dw_tc.cu", "td_loss.cu", "sage_big.cu", "optim.cu", "collate.cu", "noisy.cu", "lw_sage.cu", "lw_model.cu", "lw_bwd.cu", "api_lw.cu", "api_static.cu", "hexara.cu", "hexara_bwd.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--compiler-options", "-fPIC", "-Xptxas", "-v"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "hexgnn.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every CUDA source into gnn_hex_b200/libhexgnn.so (objects in build/, in parallel).
    HEXGNN_PHASES=1 in the environment compiles the per-phase cycle counters in (profiling builds only)."""
    if not force and not _stale():
        return LIB
    flags = NVCC_FLAGS + (["-DHEXGNN_PHASES"] if os.environ.get("HEXGNN_PHASES") == "1" else [])
    flags += ["-D" + m for m in os.environ.get("HEXGNN_DEFINES", "").split() if m]   # experiment switches
    objdir = os.path.join(HERE, "..", "build" if not os.environ.get("HEXGNN_LIB") else "build_alt")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for s in SOURCES:
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        cmd = [_nvcc()] + flags + ["-c", os.path.join(CSRC, s), "-o", obj]
        procs.append((s, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = []
    log = []
    for s, obj, p in procs:
        out, _ = p.communicate()
        log.append(out)
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError(f"nvcc failed on {s}")
        objs.append(obj)
    link = [_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    with open(os.path.join(objdir, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
