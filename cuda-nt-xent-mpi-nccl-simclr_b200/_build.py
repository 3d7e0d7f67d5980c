"""In-tree builds: the C-ABI library (nvcc, sm_100a only) and the torch binding (g++).

Everything lands next to the sources so that `gpurun` ships the built `.so` files to the
GPU box (they are git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_DIR = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libntxent_b200.so")
LIB_PATH_TRACE = LIB_PATH[:-3] + "_trace.so"
BINDING_PATH = os.path.join(PKG_DIR, "ntxent_cuda.so")

KERNEL_SOURCES = ["capi.cu", "simt_kernels.cu", "epilogue_kernels.cu", "tc_kernels.cu", "tc_fwd_sym.cu", "comm.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]
# No --use_fast_math: it would change expf/exp2f/division accuracy in the fp32 tolerance path;
# the kernels call the fast intrinsics they want explicitly instead.


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the ntxent_b200 kernels cannot be built")


def _stale(target: str, sources) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_library(force: bool = False, verbose: bool = False, trace: bool = False,
                  variant: str | None = None, defines=()) -> str:
    """nvcc -> libntxent_b200.so (all kernels + the extern "C" boundary).  trace=True builds the
    variant with the pipeline time stamps compiled in (libntxent_b200_trace.so, for
    tools/trace_bwd.py; loaded instead of the product library when NTXENT_TRACE=1).
    variant="name", defines=["-DX=1"] builds libntxent_b200_name.so for A/B runs (loaded when
    NTXENT_LIBRARY points at it)."""
    if variant:
        trace = False
    srcs = [os.path.join(CSRC, s) for s in KERNEL_SOURCES]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(REPO_DIR, "include", "ntxent_b200.h"))
    lib_path = LIB_PATH_TRACE if trace else LIB_PATH
    if variant:
        lib_path = LIB_PATH[:-3] + f"_{variant}.so"
    if force or _stale(lib_path, deps):
        objs = []
        procs = []
        for s in srcs:
            o = os.path.join(CSRC, os.path.basename(s)[:-3] +
                             (f".{variant}.o" if variant else ".trace.o" if trace else ".o"))
            objs.append(o)
            cmd = [_nvcc(), *NVCC_FLAGS, "-I", os.path.join(REPO_DIR, "include"), "-c", s, "-o", o]
            if trace:
                cmd.insert(1, "-DNTXENT_TRACE=1")
            for d in defines:
                cmd.insert(1, d)
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                                text=True)))
        log = []
        for cmd, p in procs:
            out, _ = p.communicate()
            log.append(out)
            if p.returncode != 0:
                raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + out)
        # static cudart; libdl for the late-bound NCCL (comm.cu)
        link = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib_path, *objs, "-ldl"]
        subprocess.run(link, check=True)
        if not trace and not variant:
            with open(os.path.join(CSRC, "ptxas.log"), "w") as f:
                f.write("\n".join(log))
        if verbose:
            print("\n".join(log))
    return lib_path


def write_sass_summary(path: str | None = None) -> str:
    """cuobjdump -sass of the built kernel objects -> counts of the mnemonics that show what the
    kernels are made of (tcgen05.mma = UTC*MMA, tcgen05.ld/st = LDTM/STTM, TMA = UTMALDG,
    tcgen05.commit = UTCBAR, griddepcontrol = ACQBULK/..., peer flags = LD/ST .SYS), per object.
    Written to profiles/r2_sass_summary.txt by __graft_entry__.build()."""
    import re
    path = path or os.path.join(REPO_DIR, "profiles", "r2_sass_summary.txt")
    pats = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOMSWS",
            "MUFU.EX2", r"\bHMMA", r"\bHGMMA", r"\.SYS", "FENCE.VIEW.ASYNC", "SYNCS", "ELECT", "ACQBULK", "PREEXIT",
            "CCTL.IVALL"]
    cuobjdump = os.path.join(os.path.dirname(_nvcc()), "cuobjdump")
    lines = ["# cuobjdump -sass <object> | grep -c <mnemonic>   (sm_100a cubins, built by _build.py)",
             "# flags: " + " ".join(NVCC_FLAGS)]
    for src in KERNEL_SOURCES:
        obj = os.path.join(CSRC, src[:-3] + ".o")
        if not os.path.exists(obj):
            continue
        sass = subprocess.run([cuobjdump, "-sass", obj], stdout=subprocess.PIPE, text=True).stdout
        funcs = len(re.findall(r"^\s*Function :", sass, flags=re.M))
        counts = {p: len(re.findall(p, sass)) for p in pats}
        lines.append(f"{src[:-3]}.o: {funcs} kernels; " +
                     ", ".join(f"{k.replace(chr(92), '')} {v}" for k, v in counts.items() if v))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path


def build_binding(force: bool = False) -> str:
    """g++ -> ntxent_cuda.so, the torch extension with the reference's operator surface.

    Links against libntxent_b200.so through an $ORIGIN rpath; only this TU sees torch headers.
    """
    import torch
    from torch.utils import cpp_extension

    src = os.path.join(CSRC, "binding.cpp")
    deps = [src, os.path.join(REPO_DIR, "include", "ntxent_b200.h"),
            os.path.join(REPO_DIR, "include", "ntxent_kernel.cuh")]
    build_library()
    if force or _stale(BINDING_PATH, deps + [LIB_PATH]):
        inc = []
        for p in cpp_extension.include_paths("cuda"):
            inc += ["-isystem", p]
        inc += ["-isystem", sysconfig.get_paths()["include"]]
        torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
        abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
        cmd = [
            "/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=ntxent_cuda",
            "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
            "-I", os.path.join(REPO_DIR, "include"), *inc, src, "-o", BINDING_PATH,
            "-L", PKG_DIR, "-lntxent_b200", "-L", torch_lib, "-ltorch", "-ltorch_cpu", "-lc10",
            "-ltorch_cuda", "-lc10_cuda", "-ltorch_python", "-L", "/usr/local/cuda/lib64", "-lcudart",
            "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{torch_lib}",
        ]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("binding build failed:\n" + r.stdout[-4000:])
    return BINDING_PATH


def _torch_cxx_flags():
    import torch
    from torch.utils import cpp_extension
    inc = []
    for p in cpp_extension.include_paths("cuda"):
        inc += ["-isystem", p]
    inc += ["-isystem", sysconfig.get_paths()["include"]]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    return inc, torch_lib, abi


def build_cpp_tools(force: bool = False):
    """g++ -> tests/cpp/test_ntxent_ops (the reference's C++ operator tests, re-stated) and
    csrc/ntxent_benchmark (the C++ benchmark).  Both link ntxent_cuda.so through an rpath."""
    build_binding()
    inc, torch_lib, abi = _torch_cxx_flags()
    pyver = f"python{sys.version_info.major}.{sys.version_info.minor}"
    pylibdir = sysconfig.get_config_var("LIBDIR")
    targets = [
        (os.path.join(REPO_DIR, "tests", "cpp", "test_ntxent_ops"),
         [os.path.join(REPO_DIR, "tests", "cpp", "test_ntxent_ops.cpp")],
         [os.path.join(REPO_DIR, "tests", "cpp", "mini_test.hpp")]),
        (os.path.join(PKG_DIR, "ntxent_benchmark"),
         [os.path.join(CSRC, "benchmark.cpp"), os.path.join(CSRC, "dist_benchmark.cpp")], []),
    ]
    built = []
    for exe, srcs, extra_deps in targets:
        deps = srcs + extra_deps + [BINDING_PATH, os.path.join(REPO_DIR, "include", "ntxent_kernel.cuh")]
        if force or _stale(exe, deps):
            cmd = ["/usr/bin/g++", "-O2", "-std=c++17", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
                   "-I", os.path.join(REPO_DIR, "include"), "-I", os.path.join(REPO_DIR, "tests", "cpp"), *inc,
                   *srcs, "-o", exe, BINDING_PATH, "-L", PKG_DIR, "-lntxent_b200",
                   "-L", torch_lib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda", "-lc10_cuda",
                   "-ltorch_python", "-L", pylibdir, f"-l{pyver}", "-L", "/usr/local/cuda/lib64", "-lcudart",
                   f"-Wl,-rpath,{PKG_DIR}", f"-Wl,-rpath,{torch_lib}", f"-Wl,-rpath,{pylibdir}",
                   "-Wl,--no-as-needed", "-Wl,--allow-shlib-undefined"]
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"build of {exe} failed:\n" + r.stdout[-4000:])
        built.append(exe)
    built.append(build_dist_benchmark(force))
    return built


def build_dist_benchmark(force: bool = False) -> str:
    """g++ -> ntxent_dist_benchmark: the multi-GPU C++ benchmark on the C-ABI alone (no torch)."""
    build_library()
    exe = os.path.join(PKG_DIR, "ntxent_dist_benchmark")
    src = os.path.join(CSRC, "dist_benchmark.cpp")
    if force or _stale(exe, [src, LIB_PATH, os.path.join(REPO_DIR, "include", "ntxent_b200.h")]):
        cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-pthread", "-DNTXENT_DIST_BENCH_MAIN", "-I",
               os.path.join(REPO_DIR, "include"), "-isystem", "/usr/local/cuda/include", src, "-o", exe,
               "-L", PKG_DIR, "-lntxent_b200", "-L", "/usr/local/cuda/lib64", "-lcudart",
               f"-Wl,-rpath,{PKG_DIR}", "-Wl,-rpath,$ORIGIN"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("build of ntxent_dist_benchmark failed:\n" + r.stdout[-4000:])
    return exe


if __name__ == "__main__":
    build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    if "--binding" in sys.argv:
        build_binding(force="--force" in sys.argv)
    if "--tools" in sys.argv:
        print(build_cpp_tools(force="--force" in sys.argv))
