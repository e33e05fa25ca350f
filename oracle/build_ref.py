"""Compile the reference's own CPU attention kernels into oracle/_ref/ (authoring container only).

TEST INFRASTRUCTURE ONLY.  Sources are compiled WHERE THEY LIE under /root/reference
(sgl-kernel/csrc/cpu/decode.cpp, extend.cpp, gemm.cpp, gemm_fp8.cpp + headers); nothing is copied into this repo.  Only
the outputs (two .so files and build_info.json) land in oracle/_ref/, which is git-ignored but
travels to the GPU box.  Two ISA variants are built because the GPU box's host CPU may differ:
  native  : -march=native of the authoring container (AVX512-BF16 + AMX kernels, the reference's
            own CMake flag, sgl-kernel/csrc/cpu/CMakeLists.txt:46)
  generic : -march=x86-64-v3 (the reference's scalar/at::vec fallbacks)
oracle/ref_kernels.py picks the variant the running CPU supports.

The reference's own build system (CMake + FetchContent) is not used: plain g++ on the few files.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF_CPU = os.path.join(os.environ.get("B200_REFERENCE_ROOT", "/root/reference"), "sgl-kernel", "csrc",
                       "cpu")
# decode/extend attention + the W8A16 block-FP8 linear (fp8_scaled_mm_cpu) and its weight pre-pack
REF_SOURCES = ("decode.cpp", "extend.cpp", "gemm.cpp", "gemm_fp8.cpp")
ISA_FLAGS = ("avx2", "fma", "avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512_bf16",
             "avx512_vnni", "amx_tile", "amx_bf16", "amx_int8", "avx512_fp16")


def cpu_flags():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def _compile(variant: str, march: str) -> str:
    import torch
    from torch.utils import cpp_extension

    out = os.path.join(OUT, f"libsgl_cpu_ref_{variant}.so")
    inc = []
    for p in cpp_extension.include_paths():
        inc += ["-isystem", p]
    ref_root = os.path.dirname(os.path.dirname(REF_CPU))  # sgl-kernel
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-shared", "-fPIC", "-O3", f"-march={march}", "-fopenmp", "-std=c++17",
           "-Wno-unknown-pragmas", "-Wno-deprecated-declarations",
           f"-DB200_REF_NS=sgl_ref_{variant}", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           "-I", REF_CPU, "-I", os.path.join(ref_root, "include"), "-I", os.path.join(ref_root, "csrc"),
           *inc,
           os.path.join(HERE, "ref_shim.cpp"), *[os.path.join(REF_CPU, f) for f in REF_SOURCES],
           "-o", out, f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-lc10", "-l:libgomp.so.1",
           f"-Wl,-rpath,{torch_lib}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"reference CPU kernels ({variant}) failed to build:\n{res.stderr[-4000:]}")
    return out


def build(force: bool = False):
    if not os.path.isdir(REF_CPU):
        return None  # GPU box: only the prebuilt files are used
    os.makedirs(OUT, exist_ok=True)
    info_path = os.path.join(OUT, "build_info.json")
    srcs = [os.path.join(HERE, "ref_shim.cpp"), *[os.path.join(REF_CPU, f) for f in REF_SOURCES],
            os.path.abspath(__file__)]
    newest = max(os.path.getmtime(s) for s in srcs)
    outs = [os.path.join(OUT, f"libsgl_cpu_ref_{v}.so") for v in ("native", "generic")]
    if (not force and os.path.exists(info_path)
            and all(os.path.exists(o) and os.path.getmtime(o) > newest for o in outs)):
        return OUT
    with ThreadPoolExecutor(max_workers=2) as ex:
        list(ex.map(lambda a: _compile(*a), [("native", "native"), ("generic", "x86-64-v3")]))
    flags = cpu_flags()
    json.dump({"native_required_flags": sorted(f for f in ISA_FLAGS if f in flags),
               "sources": [f"sgl-kernel/csrc/cpu/{f}" for f in REF_SOURCES],
               "compiler": "g++ -O3 -fopenmp"}, open(info_path, "w"), indent=1)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
