#!/usr/bin/env python
"""Build the UNMODIFIED reference CUDA extension into oracle/_ref/ (test infrastructure only).

TEST INFRASTRUCTURE -- never imported by the product path.  Only tests/, bench.py
(--impl reference / cpu_baseline legs) and __graft_entry__ may touch oracle/.

The reference sources are compiled *where they lie* under /root/reference with an explicit
nvcc / g++ recipe (we do not run the reference's setup.py); nothing is copied into this repo.
Outputs go only to oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun):

    oracle/_ref/mini_flash_attention_ref_C.so    pybind module (reference csrc/api.cpp:4-9)

The module is named mini_flash_attention_ref_C so that it can coexist with our own package.
It needs torch at import time (the reference is a torch extension, csrc/mfa/api.cpp).

The reference has no CPU path (csrc/mfa/api.cpp:123 CUDAGuard), so this .so can only be
*executed* on the GPU box; here we only build and import it.
"""
import os
import subprocess
import sys
import sysconfig
import time
from concurrent.futures import ThreadPoolExecutor

REF = os.environ.get("MFA_REFERENCE_DIR", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
MODNAME = "mini_flash_attention_ref_C"


def main() -> int:
    if not os.path.isdir(os.path.join(REF, "csrc", "mfa")):
        print(f"[build_ref] {REF} not present; nothing to do (GPU box uses the prebuilt .so)")
        return 0
    so_path = os.path.join(OUT, MODNAME + ".so")
    if os.path.exists(so_path) and "--force" not in sys.argv:
        print(f"[build_ref] {so_path} exists; use --force to rebuild")
        return 0
    os.makedirs(OUT, exist_ok=True)
    import torch
    from torch.utils import cpp_extension as ce

    cuda_home = ce.CUDA_HOME or "/usr/local/cuda"
    inc = [
        os.path.join(REF, "csrc"), os.path.join(REF, "csrc", "mfa"),
        os.path.join(cuda_home, "include"), os.path.join(cuda_home, "include", "cccl"),
        os.path.join(REF, "3rd", "cutlass", "include"),
        sysconfig.get_paths()["include"],
    ] + ce.include_paths()
    iflags = [f"-I{p}" for p in inc]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    common = [f"-DTORCH_EXTENSION_NAME={MODNAME}", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={abi}"]
    # flags restated from the reference's setup.py:40-52 (plus an explicit sm_100a gencode,
    # which the reference leaves to TORCH_CUDA_ARCH_LIST)
    nvcc = [os.path.join(cuda_home, "bin", "nvcc"), "-std=c++20", "-O3", "--generate-line-info",
            "--use_fast_math", "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
            "-U__CUDA_NO_HALF2_OPERATORS__", "-U__CUDA_NO_BFLOAT16_CONVERSIONS__",
            "--expt-relaxed-constexpr", "--expt-extended-lambda",
            "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "--threads", "4"]
    cxx = ["g++", "-std=c++20", "-O3", "-fPIC"]
    jobs = [
        (nvcc + common + iflags + ["-c", os.path.join(REF, "csrc/mfa/flash.cu"), "-o", os.path.join(OUT, "flash.o")]),
        (cxx + common + iflags + ["-c", os.path.join(REF, "csrc/mfa/api.cpp"), "-o", os.path.join(OUT, "mfa_api.o")]),
        (cxx + common + iflags + ["-c", os.path.join(REF, "csrc/api.cpp"), "-o", os.path.join(OUT, "api.o")]),
    ]
    t0 = time.time()

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.returncode

    with ThreadPoolExecutor(3) as ex:
        rcs = list(ex.map(run, jobs))
    if any(rcs):
        print("[build_ref] compile failed")
        return 1
    libdirs = ce.library_paths(device_type="cuda")
    link = (["g++", "-shared", "-o", so_path] + [os.path.join(OUT, o) for o in ("flash.o", "mfa_api.o", "api.o")]
            + [f"-L{p}" for p in libdirs] + [f"-Wl,-rpath,{p}" for p in libdirs]
            + ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart"])
    if run(link):
        print("[build_ref] link failed")
        return 1
    for o in ("flash.o", "mfa_api.o", "api.o"):
        os.remove(os.path.join(OUT, o))
    print(f"[build_ref] built {so_path} in {time.time() - t0:.0f}s")
    return 0


if __name__ == "__main__":
    sys.exit(main())
