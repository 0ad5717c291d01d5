"""In-tree build of libmfa_b200.so (sm_100a only) with plain nvcc.

    python -m mini_flash_attention_b200.build [--force] [--verbose]

The shared library exports exactly the C-ABI declared in include/mfa_b200.h.  It is built next to this
file (git-ignored, but shipped to the GPU box by gpurun) so that the loaded .so is visibly in-tree.
"""
from __future__ import annotations

import os
import subprocess
import sys
import time
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
REPO = os.path.dirname(PKG_DIR)
LIB_PATH = os.path.join(PKG_DIR, "libmfa_b200.so")
OBJ_DIR = os.path.join(PKG_DIR, "_obj")

SOURCES = ["capi.cu", "tmap.cu", "decode_sm100.cu", "prefill_sm100.cu", "selftest_sm100.cu", "selftest2_sm100.cu", "kvcache_append_sm100.cu", "xchg_sm100.cu"]
HEADERS = ["kparams.h", "ptx.cuh", "tmap.h", "merge.cuh", os.path.join(REPO, "include", "mfa_b200.h")]

NVCC_FLAGS = [
    "-std=c++17", "-O3", "-lineinfo", "--use_fast_math",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    cand = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
    return cand if os.path.exists(cand) else "nvcc"


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_variant(name: str, defines, verbose: bool = False) -> str:
    """Tuning aid: build csrc/ with extra -D flags into mini_flash_attention_b200/variants/lib<name>.so."""
    vdir = os.path.join(PKG_DIR, "variants")
    odir = os.path.join(vdir, "_obj_" + name)
    os.makedirs(odir, exist_ok=True)
    out = os.path.join(vdir, f"lib{name}.so")
    nvcc = _nvcc()

    def one(src):
        obj = os.path.join(odir, os.path.basename(src) + ".o")
        cmd = [nvcc] + NVCC_FLAGS + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-I", CSRC, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(r.stdout + r.stderr)
        if verbose:
            print(r.stderr)
        return obj
    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(one, [os.path.join(CSRC, s) for s in SOURCES]))
    r = subprocess.run([nvcc, "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"], capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError(r.stdout + r.stderr)
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    hdrs = [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    if not force and not _stale(LIB_PATH, srcs + hdrs + [os.path.abspath(__file__)]):
        return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = _nvcc()
    t0 = time.time()

    def compile_one(src: str):
        obj = os.path.join(OBJ_DIR, os.path.basename(src) + ".o")
        if not force and not _stale(obj, [src] + hdrs + [os.path.abspath(__file__)]):
            return obj, 0, ""
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-I", CSRC, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return obj, r.returncode, r.stdout + r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(compile_one, srcs))
    for obj, rc, out in results:
        if verbose and out:
            print(out)
        if rc != 0:
            raise RuntimeError(f"nvcc failed for {obj}:\n{out}")
    link = [nvcc, "-shared", "-o", LIB_PATH] + [o for o, _, _ in results] + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(f"[mfa_b200] built {LIB_PATH} in {time.time() - t0:.1f}s")
    return LIB_PATH


SHIM_NAME = "_mfa_torch_shim"
SHIM_SRC = os.path.join(CSRC, "torch_shim.cpp")


def shim_path() -> str:
    import sysconfig
    return os.path.join(PKG_DIR, SHIM_NAME + sysconfig.get_config_var("EXT_SUFFIX"))


def build_shim(force: bool = False, verbose: bool = False) -> str:
    """Compiled host boundary (csrc/torch_shim.cpp): a torch extension module over libmfa_b200.so, built with plain
    g++ against this interpreter's torch headers.  Depends on the torch build, so it is rebuilt when torch changes."""
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    out = shim_path()
    lib = build(force=False, verbose=verbose)
    deps = [SHIM_SRC, os.path.join(REPO, "include", "mfa_b200.h"), os.path.abspath(__file__), torch.__file__]
    if not force and not _stale(out, deps):
        return out
    t0 = time.time()
    incs = ce.include_paths("cuda") + [sysconfig.get_paths()["include"]]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cuda_lib = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "lib64")
    cmd = (["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-Wno-attributes",
            f"-DTORCH_EXTENSION_NAME={SHIM_NAME}", "-DTORCH_API_INCLUDE_EXTENSION_H",
            f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
           + [f"-I{i}" for i in incs] + [SHIM_SRC, "-o", out]
           + [f"-L{PKG_DIR}", "-lmfa_b200", f"-L{torch_lib}", "-ltorch_python", "-ltorch", "-ltorch_cpu", "-ltorch_cuda",
              "-lc10", "-lc10_cuda", f"-L{cuda_lib}", "-lcudart",
              "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{torch_lib}", "-Wl,--no-as-needed"])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("torch shim build failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(f"[mfa_b200] built {out} in {time.time() - t0:.1f}s")
    return out


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv)
    print(LIB_PATH)
    if "--no-shim" not in sys.argv:
        print(build_shim(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv))
