"""TEST INFRASTRUCTURE ONLY — builds the *unmodified* reference CUDA extensions for sm_100a.

Sources are compiled where they lie under /root/reference (never copied); products go
to oracle/_ref/<ext>/ (git-ignored, NOT gpurun-ignored, so the .so files travel to the
GPU box).  They are used by tests/ (as the strongest parity oracle: identical tensors in,
bit-compare integer outputs) and by bench.py's reported "reference CUDA build" numbers.
The product path never imports anything from here.

The reference's own flags use -std=c++14 (e.g. gridencoder/backend.py:7-10), which
torch 2.11 headers reject; we substitute -std=c++17 and pin the arch to sm_100a.

Usage:  python oracle/build_ref.py [gridencoder raymarching shencoder ffmlp]
"""
import os
import sys

REF = os.environ.get("NERFNAV_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

EXTS = {
    # name -> (module name, subdir, sources, extra include dirs, extra nvcc flags)
    "gridencoder": ("_ref_gridencoder", "gridencoder", ["gridencoder.cu", "bindings.cpp"], [], []),
    "raymarching": ("_ref_raymarching", "raymarching", ["raymarching.cu", "bindings.cpp"], [], []),
    "shencoder": ("_ref_shencoder", "shencoder", ["shencoder.cu", "bindings.cpp"], [], []),
    "ffmlp": ("_ref_ffmlp", "ffmlp", ["ffmlp.cu", "bindings.cpp"],
              ["dependencies/cutlass/include", "dependencies/cutlass/tools/util/include"],
              ["--expt-extended-lambda", "-Xcompiler=-mf16c", "-Xcompiler=-Wno-float-conversion",
               "-Xcompiler=-fno-strict-aliasing"]),
}


def so_path(name):
    mod = EXTS[name][0]
    return os.path.join(OUT, name, mod + ".so")


def build(names=None, verbose=False):
    if not os.path.isdir(REF):
        return []
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils.cpp_extension import load
    built = []
    for name in (names or list(EXTS)):
        mod, sub, srcs, incs, flags = EXTS[name]
        if os.path.exists(so_path(name)):
            built.append(name)
            continue
        bdir = os.path.join(OUT, name)
        os.makedirs(bdir, exist_ok=True)
        load(name=mod,
             sources=[os.path.join(REF, sub, "src", s) for s in srcs],
             extra_include_paths=[os.path.join(REF, sub, i) for i in incs],
             extra_cflags=["-O3", "-std=c++17"],
             extra_cuda_cflags=["-O3", "-std=c++17", "--expt-relaxed-constexpr",
                                "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
                                "-U__CUDA_NO_HALF2_OPERATORS__"] + flags,
             build_directory=bdir, verbose=verbose, is_python_module=False)
        built.append(name)
    return built


def load_ref(name):
    """Import a prebuilt reference extension (tests/bench only). Returns None if absent."""
    p = so_path(name)
    if not os.path.exists(p):
        return None
    import importlib.util
    import torch  # noqa: F401  (libtorch symbols must be loaded first)
    spec = importlib.util.spec_from_file_location(EXTS[name][0], p)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


if __name__ == "__main__":
    print(build(sys.argv[1:] or None, verbose=True))
