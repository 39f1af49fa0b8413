"""Builds eds-bwt_b200/libedsbwt_cuda.so and the two CLI executables with nvcc for sm_100a (in-tree)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libedsbwt_cuda.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function", "-Xptxas", "-v"]
LIB_SRCS = ["engine.cu", "fused.cu", "items.cu", "multi.cu", "cache.cu", "move.cu", "api.cu", "gather_bench.cu", "static_scratch.cu", "builder.cu", "csv.cu", "host_io.cpp", "pack_host.cc"]
CLIS = {"EDSBWTsearch": "cli_edsbwtsearch.cpp", "MOVE_EDSBWTSearch": "cli_move_edsbwtsearch.cpp", "stringCheck": "cli_stringcheck.cpp", "build_MLF": "cli_build_mlf.cpp"}


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    hdrs = [os.path.join(CSRC, "internal.cuh"), os.path.join(HERE, "..", "include", "edsbwt.h"), os.path.join(CSRC, "cli_common.hpp")]
    objs = []
    stale = []
    for src in LIB_SRCS:
        sp = os.path.join(CSRC, src)
        obj = os.path.join(HERE, "build", src + ".o")
        objs.append(obj)
        if force or _newer(obj, [sp] + hdrs):
            stale.append((src, sp, obj))

    def compile_one(job):
        src, sp, obj = job
        if src.endswith(".cc"):       # host-only translation unit (AVX2 intrinsics): straight to g++
            cmd = ["g++", "-O3", "-std=c++17", "-fPIC", "-Wall", "-pthread", "-c", sp, "-o", obj]
        else:
            cmd = [NVCC] + ARCH + COMMON + ["-x", "cu", "-c", sp, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        with open(obj + ".log", "w") as f:
            f.write(log)
        return src, r.returncode, log

    # the translation units are independent: compile them side by side (a full build: 2 min one after the other, 30 s on 8 cores)
    with ThreadPoolExecutor(max_workers=max(1, min(len(stale), os.cpu_count() or 1))) as pool:
        for src, rc, log in pool.map(compile_one, stale):
            if rc != 0:
                sys.stderr.write(log)
                raise RuntimeError("nvcc failed on " + src)
            if verbose:
                print(log)
    if force or _newer(LIB, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-lcudart"]
        subprocess.run(cmd, check=True)
    for exe, src in CLIS.items():
        sp = os.path.join(CSRC, src)
        if not os.path.exists(sp):
            continue
        out = os.path.join(HERE, exe)
        if force or _newer(out, [sp, LIB] + hdrs):
            cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-o", out, sp, "-I", os.path.join(HERE, "..", "include"),
                   "-L", HERE, "-ledsbwt_cuda", "-Wl,-rpath,$ORIGIN"]
            subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
