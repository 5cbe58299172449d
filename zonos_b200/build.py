"""Builds zonos_b200/lib/libzonos_b200.so from zonos_b200/csrc/*.cu for sm_100a (nvcc cross-compiles without a GPU).

In-tree build: the .so travels to the GPU box with the repository snapshot.  `python -m zonos_b200.build`.
"""
import concurrent.futures
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libzonos_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "--use_fast_math=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--expt-relaxed-constexpr",
]
if os.environ.get("ZB200_DBG_NOA"):  # what-if experiment: the decode GEMMs skip their activation loads (wrong results)
    NVCC_FLAGS.append("-DSK_DBG_NOA")
if os.environ.get("ZB200_DBG_NOMMA"):  # what-if experiment: the decode GEMMs issue no MMAs (wrong results)
    NVCC_FLAGS.append("-DSK_DBG_NOMMA")
for _k in ("SK_SA128", "SK_SW128_KB"):  # ring-depth experiments
    if os.environ.get(_k):
        NVCC_FLAGS.append(f"-D{_k}={os.environ[_k]}")
if os.environ.get("ZB200_TRACE"):
    NVCC_FLAGS.append("-DSK_TRACE")
if os.environ.get("ZB200_TUNING"):  # experiment builds only: the library then honours its ZB200_* environment switches
    NVCC_FLAGS.append("-DZB200_TUNING")
if os.environ.get("ZB200_EXTRA_DEFS"):  # experiment builds: extra -D flags, written to lib/libzonos_b200<ZB200_LIB_SUFFIX>.so
    NVCC_FLAGS += os.environ["ZB200_EXTRA_DEFS"].split()
if os.environ.get("ZB200_LIB_SUFFIX"):
    LIB = os.path.join(LIBDIR, f"libzonos_b200{os.environ['ZB200_LIB_SUFFIX']}.so")
    OBJ = OBJ + os.environ["ZB200_LIB_SUFFIX"]
NVCC_FLAGS.remove("--use_fast_math=false")  # precise math on purpose: the sampler/Snake must match the reference


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    raise RuntimeError("nvcc not found")


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest():
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)) + ["../../include/zonos_b200.h"]:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(f.encode() + b"\0" + fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    stamp = LIB + ".sha256" if os.environ.get("ZB200_LIB_SUFFIX") else os.path.join(LIBDIR, "build.sha256")
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read().strip() == dig:
        return LIB
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(OBJ, src[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(compile_one, _sources()))
    r = subprocess.run([nvcc, "-shared", "-o", LIB, *objs, "-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as fh:
        fh.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
