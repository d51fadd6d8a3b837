"""Build the in-tree native libraries (sm_100a only).

    python -m mllm_b200.build            # kernels + host backend
Outputs (git-ignored, but they travel to the GPU box with the tree):
    mllm_b200/lib/libmllmcu.so         CUDA kernels + thin C ABI (include/mllmcu.h)
    mllm_b200/lib/libmllm_cuda_backend.so  C++ host side (CUDABackend / ops / ParamLoader / models)
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "lib")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_kernels(force=False, verbose=False):
    """One object per .cu (compiled in parallel, rebuilt only when it or a shared header changed), then one link."""
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(os.path.join(LIB, "obj"), exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))
    hdrs = glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + [os.path.join(ROOT, "include", "mllmcu.h")]
    out = os.path.join(LIB, "libmllmcu.so")
    base = [NVCC, "-O3", "-std=c++17", *ARCH, "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-I", os.path.join(ROOT, "include")]
    if verbose:
        base.insert(1, "-Xptxas=-v")

    def one(src):
        obj = os.path.join(LIB, "obj", os.path.basename(src)[:-3] + ".o")
        if force or _newer(obj, [src] + hdrs):
            subprocess.run(base + ["-c", src, "-o", obj], check=True)
            return obj, True
        return obj, False

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        res = list(ex.map(one, srcs))
    objs = [o for o, _ in res]
    if force or any(c for _, c in res) or _newer(out, objs):
        subprocess.run([NVCC, *ARCH, "-shared", "-o", out, *objs, "-lcuda"], check=True)
    return out


def build_host(force=False):
    os.makedirs(LIB, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(HERE, "csrc", "host", "*.cpp")))
    if not srcs:
        return None
    deps = srcs + glob.glob(os.path.join(HERE, "csrc", "host", "*.hpp")) + [os.path.join(ROOT, "include", "mllmcu.h")]
    out = os.path.join(LIB, "libmllm_cuda_backend.so")
    if not force and not _newer(out, deps):
        return out
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(HERE, "csrc", "host"), "-o", out, *srcs, "-L", LIB, "-lmllmcu", "-Wl,-rpath,$ORIGIN", "-lpthread"]
    subprocess.run(cmd, check=True)
    return out


def build_all(force=False, verbose=False):
    k = build_kernels(force, verbose)
    h = build_host(force)
    return k, h


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
