"""In-tree build of the native pieces (no pip, no JIT cache):

  memb_b200/libmemb_b200.so                C ABI: host loader + sm_100a kernels (nvcc)
  memb_b200/_memb.cpython-*.so             pybind11 module on top of the C ABI (g++)
  oracle/liboracle.so                      CPU oracle (test infrastructure, g++)
  oracle/_ref/libmemb_ref.so               reference sources compiled in place, only
                                           where /root/reference exists (oracle/Makefile)

`python build_native.py [--force]` or `build_native.build_all()`.  Rebuilds a target
only when one of its inputs is newer than the output.
"""
import os
import shutil
import subprocess
import sys
import sysconfig

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "memb_b200")
CSRC = os.path.join(PKG, "csrc")
INCLUDE = os.path.join(ROOT, "include")
ORACLE = os.path.join(ROOT, "oracle")

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
LIB_NAME = "libmemb_b200.so"


def _newer(inputs, output):
    if not os.path.exists(output):
        return True
    out_time = os.path.getmtime(output)
    return any(os.path.getmtime(path) > out_time for path in inputs)


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build_library(force=False):
    out = os.path.join(PKG, LIB_NAME)
    headers = [os.path.join(CSRC, name) for name in ("kernels.cuh", "trained_rows.cuh", "fb_format.h", "memb_file.h", "builder.h")]
    headers.append(os.path.join(INCLUDE, "memb_b200.h"))
    sources = [os.path.join(CSRC, name) for name in ("device_reader.cu", "device_builder.cu", "memb_file.cpp", "builder.cpp")]
    if not force and not _newer(headers + sources, out):
        return out
    build_dir = os.path.join(PKG, "build")
    os.makedirs(build_dir, exist_ok=True)
    objects = []
    for src in sources:
        obj = os.path.join(build_dir, os.path.basename(src) + ".o")
        objects.append(obj)
        if not force and not _newer(headers + [src], obj):
            continue
        if src.endswith(".cu"):
            # MEMB_B200_NVCC_DEFS: extra -D switches for kernel experiments (development aid)
            _run([_nvcc(), *NVCC_ARCH, "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC,-Wall",
                  *os.environ.get("MEMB_B200_NVCC_DEFS", "").split(), "-c", src, "-o", obj])
        else:
            # -ffp-contract=off: the k-means running mean and the uniform quantiser
            # must round like the reference's plain x86-64 build (no FMA)
            _run(["g++", "-std=c++17", "-O3", "-fPIC", "-Wall", "-Wextra", "-ffp-contract=off", "-pthread",
                  "-c", src, "-o", obj])
    _run([_nvcc(), *NVCC_ARCH, "-shared", "-o", out, *objects, "-Xcompiler", "-pthread"])
    return out


def build_module(force=False):
    import pybind11

    suffix = sysconfig.get_config_var("EXT_SUFFIX")
    out = os.path.join(PKG, "_memb" + suffix)
    src = os.path.join(CSRC, "bindings.cpp")
    if not force and not _newer([src, os.path.join(INCLUDE, "memb_b200.h")], out):
        return out
    _run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-fvisibility=hidden",
          "-I" + pybind11.get_include(), "-I" + sysconfig.get_paths()["include"],
          src, "-o", out, "-L" + PKG, "-lmemb_b200", "-Wl,-rpath,$ORIGIN"])
    return out


def build_oracle(force=False):
    out = os.path.join(ORACLE, "liboracle.so")
    src = os.path.join(ORACLE, "oracle.cpp")
    if os.path.exists(src) and (force or _newer([src], out)):
        # -O3 like the reference's CMakeLists.txt:15; contraction off so uniform
        # dequant rounds like its plain x86-64 SSE build
        _run(["g++", "-std=c++14", "-O3", "-fPIC", "-shared", "-Wall", "-ffp-contract=off", "-pthread",
              src, "-o", out])
    if os.path.isdir("/root/reference/src") and os.path.exists(os.path.join(ORACLE, "Makefile")):
        _run(["make", "-s", "-C", ORACLE, "ref"])
    return out


def build_all(force=False):
    build_library(force)
    build_module(force)
    build_oracle(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
