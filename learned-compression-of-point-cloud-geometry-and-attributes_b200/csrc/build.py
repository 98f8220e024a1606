"""Builds libpcc_b200.so (sm_100a) in-tree with nvcc.  Invoked by __graft_entry__.build()."""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
OUT = os.path.join(PKG, "libpcc_b200.so")
OBJ = os.path.join(HERE, "_obj")
NVCC = os.environ.get("PCC_NVCC", "/usr/local/cuda/bin/nvcc")
CCBIN = "/usr/bin/g++"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CUFLAGS = ARCH + ["-O3", "-std=c++17", "-lineinfo", "-ccbin", CCBIN, "-Xcompiler", "-fPIC,-O3",
                  "--expt-relaxed-constexpr", "-Xptxas", "-v"]
SOURCES = ["common.cu", "coords.cu", "conv_exact.cu", "conv_tf32.cu", "conv_dispatch.cu",
           "select_entropy.cu", "octree.cu", "train.cu", "wgrad_tf32.cu", "host_coder.cpp"]
HEADERS = ["pcc_common.cuh", os.path.join("..", "..", "include", "pcc_b200.h")]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src):
    obj = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
    deps = [os.path.join(HERE, src)] + [os.path.join(HERE, h) for h in HEADERS]
    if not _stale(obj, deps):
        return obj, ""
    cmd = [NVCC] + CUFLAGS + ["-x", "cu", "-c", os.path.join(HERE, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj, r.stderr


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    with cf.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        results = list(ex.map(_compile, SOURCES))
    objs = [o for o, _ in results]
    log = "".join(l for _, l in results)
    if log:
        with open(os.path.join(OBJ, "ptxas.log"), "a") as f:
            f.write(log)
        if verbose:
            print(log)
    if _stale(OUT, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-ccbin", CCBIN, "-o", OUT] + objs + ["-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
