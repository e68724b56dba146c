"""In-tree builds: the CUDA product library (nvcc, sm_100a), the synthetic-frame generator (gcc) and,
for the checker only, the CPU oracle (make -C oracle).  Nothing here needs a GPU."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "liborbfe.so")
SYNTH = os.path.join(CSRC, "liborbfe_synth.so")
ORACLE = os.path.join(ROOT, "oracle", "_build", "liborbref.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-shared"]


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout


def build_cuda(force=False, defines=(), out=None):
    """One object per .cu file (compiled in parallel, only the stale ones), then one link.  `defines` / `out` build a
    variant library (tools/variant_sweep.sh); variants always recompile everything into their own object directory."""
    from concurrent.futures import ThreadPoolExecutor
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hdrs = glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.inc")) + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(ROOT, "include", "*.h"))
    lib = out or LIB
    variant = bool(defines) or out is not None
    if not force and not variant and _newer(LIB, srcs + hdrs):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(HERE, "_build", "obj" + ("_" + os.path.basename(lib) if variant else ""))
    os.makedirs(objdir, exist_ok=True)
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + ["-D" + d for d in defines]

    def one(src):
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        if force or variant or not _newer(obj, [src] + hdrs):
            _run([nvcc] + flags + ["-c", "-o", obj, src])
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        objs = list(pool.map(one, srcs))
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib] + objs)
    return lib


def build_synth(force=False):
    src = os.path.join(CSRC, "synth.c")
    if not force and _newer(SYNTH, [src]):
        return SYNTH
    _run(["gcc", "-O2", "-shared", "-fPIC", "-o", SYNTH, src])
    return SYNTH


def build_oracle(force=False):
    srcs = glob.glob(os.path.join(ROOT, "oracle", "cpu_ref", "*")) + [os.path.join(CSRC, "orb_pattern.inc")]
    if not force and _newer(ORACLE, srcs):
        return ORACLE
    _run(["make", "-C", os.path.join(ROOT, "oracle")] + (["-B"] if force else []))
    return ORACLE


def build_all(force=False):
    return build_cuda(force), build_synth(force), build_oracle(force)


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))


REFERENCE_SRC = "/root/reference/src/ORBextractor.cc"
REF_LIBS = [os.path.join(ROOT, "oracle", "_ref", n) for n in ("liborbslam3_orbextractor.so", "liborbslam3_orbextractor_mt.so", "libdbow2_orbvoc.so", "liborbslam3_orbmatcher.so", "liborbslam3_mappoint.so", "liborbslam3_frame.so")]


def build_reference(force=False):
    """oracle/_ref: the reference's own ORBextractor.cc compiled where it lies (make -C oracle ref).  Only possible in the
    container that has /root/reference; elsewhere the prebuilt libraries (they travel with the snapshot) are used as is."""
    if not os.path.exists(REFERENCE_SRC):
        return [p for p in REF_LIBS if os.path.exists(p)]
    _run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"] + (["-B"] if force else []))
    # the shim's ORBmatcher object layer behind the reference wrappers' C entry points (test artefact; links liborbfe.so)
    _run(["make", "-C", os.path.join(ROOT, "oracle"), "shim"] + (["-B"] if force else []))
    return REF_LIBS
