"""Builds the in-tree native libraries (sm_100a only; nvcc cross-compiles without a GPU).

  chanamq_b200/libchanagpu.so   the product: C ABI of include/chanamq_gpu.h (CUDA kernels + host table)
  chanamq_b200/libchqgen.so     synthetic workload generators (tests / bench input only)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_product(force=False, verbose=False):
    out = os.path.join(HERE, "libchanagpu.so")
    srcs = [os.path.join(CSRC, f) for f in ("kernels.cu", "abi.cu", "decode.cu", "headers.cu", "fanout.cu", "encode.cu", "shard.cu", "pack.cu", "host_table.cpp", "host_headers.cpp", "host_inc.cpp", "numa.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "chanamq_gpu.h")]
    force = force or bool(os.environ.get("CHQ_NVCC_DEFS", "").strip())  # experiment builds never reuse a library
    if not force and not _newer(out, deps):
        return out
    cmd = [NVCC, *ARCH, "-O3", "-std=c++17", "-lineinfo", "-shared", "-Xcompiler", "-fPIC,-Wall",
           "-cudart", "static", "-o", out, *srcs, "-ldl"]
    cmd[1:1] = os.environ.get("CHQ_NVCC_DEFS", "").split()  # experiment switches, e.g. -DCHQ_PROBE_BATCH=2
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return out


def build_gen(force=False):
    out = os.path.join(HERE, "libchqgen.so")
    srcs = [os.path.join(HERE, "workload", f) for f in sorted(os.listdir(os.path.join(HERE, "workload"))) if f.endswith(".cpp")]
    if not force and not _newer(out, srcs):
        return out
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", out, *srcs])
    return out


def build_all(force=False, verbose=False):
    return build_product(force, verbose), build_gen(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
