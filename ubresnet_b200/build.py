"""In-tree build of libubresnet_b200.so (sm_100a only; nvcc cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIBPATH = os.path.join(LIBDIR, "libubresnet_b200.so")
SOURCES = ["runtime.cu", "fp32_path.cu", "batch_fill.cu", "conv_gemm.cu", "conv_halo.cu", "chain_halo.cu", "wgrad_gemm.cu", "wgrad_halo.cu", "pointwise.cu", "stem.cu", "stem_tc.cu", "optim.cu", "loss_tiling_pack.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--use_fast_math", "-Xptxas", "-v"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isfile(cand) or cand == "nvcc"):
            return cand
    return "nvcc"


HASHPATH = os.path.join(LIBDIR, "source.sha256")


def source_hash():
    """sha256 over every file of csrc/, the public header and the compiler flags: the library is rebuilt when (and only
    when) this changes -- file times do not survive a checkout or the copy to the GPU box"""
    import hashlib
    h = hashlib.sha256()
    deps = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + [os.path.join(HERE, "..", "include", "ubresnet_b200.h")]
    for d in deps:
        h.update(os.path.basename(d).encode())
        with open(d, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS + SOURCES).encode())
    return h.hexdigest()


def _stale():
    if not os.path.isfile(LIBPATH) or not os.path.isfile(HASHPATH):
        return True
    with open(HASHPATH) as fh:
        return fh.read().strip() != source_hash()


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link the shared library. Returns its path."""
    if not force and not _stale():
        return LIBPATH
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    logs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, obj, p in procs:
        out, _ = p.communicate()
        logs.append("== %s ==\n%s" % (src, out))
        if p.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, out))
        objs.append(obj)
    cmd = [_nvcc(), "-shared", "-o", LIBPATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static",
                                                       "-ldl", "-lrt", "-lpthread"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s" % r.stdout)
    with open(os.path.join(LIBDIR, "build.log"), "w") as fh:
        fh.write("\n".join(logs))
    with open(HASHPATH, "w") as fh:
        fh.write(source_hash() + "\n")
    if verbose:
        print("\n".join(logs))
    return LIBPATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
