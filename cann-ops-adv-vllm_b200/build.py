"""In-tree build of the two native artefacts (no JIT cache: the .so files are git-ignored but travel with the
gpurun snapshot of the working tree):

  libspa_b200.so     hand-written sm_100a CUDA kernels + the C ABI (include/spa_b200.h); nvcc only,
                     no torch -- loadable from C, ctypes or another framework.
  custom_ops_lib.so  PyTorch binding (TORCH_LIBRARY myops + pybind module), g++ only.

Replaces the reference's extension/setup.py + CMake/CANN packaging for this path.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libspa_b200.so")
EXT = os.path.join(HERE, "custom_ops_lib.so")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build_kernels(force=False, verbose=False):
    import glob
    units = [os.path.join(CSRC, "spa_kernels.cu"), os.path.join(CSRC, "spa_kernels_opt.cu")]
    deps = sorted(glob.glob(os.path.join(CSRC, "*.cu*")))                  # every source and header
    deps.append(os.path.join(HERE, "..", "include", "spa_b200.h"))
    if not force and not _newer(LIB, deps):
        return LIB
    nvcc = [os.path.join(CUDA_HOME, "bin", "nvcc"), "-gencode", "arch=compute_100a,code=sm_100a", "-O3",
            "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC,-fvisibility=hidden"]
    if verbose:
        nvcc += ["-Xptxas", "-v"]
    # the two translation units (plain / options build of the chain kernel, csrc/spa_chain.cuh) compile in parallel
    objdir = os.path.join(HERE, "_obj")
    os.makedirs(objdir, exist_ok=True)
    objs, procs = [], []
    for u in units:
        o = os.path.join(objdir, os.path.basename(u) + ".o")
        cmd = nvcc + ["-c", "-o", o, u]
        print("+", " ".join(cmd), flush=True)
        procs.append((cmd, subprocess.Popen(cmd)))
        objs.append(o)
    for cmd, pr in procs:
        if pr.wait() != 0:
            raise subprocess.CalledProcessError(pr.returncode, cmd)
    _run(nvcc + ["-shared", "-o", LIB] + objs)
    return LIB


def build_binding(force=False):
    src = os.path.join(CSRC, "custom_ops_lib.cpp")
    if not force and not _newer(EXT, [src, os.path.join(HERE, "..", "include", "spa_b200.h"), LIB]):
        return EXT
    import torch
    from torch.utils import cpp_extension as ce
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = ce.include_paths() + [os.path.join(CUDA_HOME, "include"), sysconfig.get_paths()["include"]]
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
           "-DTORCH_EXTENSION_NAME=custom_ops_lib", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    for i in inc:
        cmd += ["-isystem", i]
    cmd += [src, "-o", EXT, f"-L{HERE}", "-lspa_b200", f"-L{tlib}", "-lc10", "-lc10_cuda", "-ltorch_cpu",
            "-ltorch_cuda", "-ltorch", "-ltorch_python", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
    _run(cmd)
    return EXT


def build_all(force=False, verbose=False):
    build_kernels(force, verbose)
    build_binding(force)
    return LIB, EXT


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
