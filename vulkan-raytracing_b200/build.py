"""Builds the native libraries in-tree (they travel to the GPU box with the snapshot):

  vulkan-raytracing_b200/librt_b200.so   CUDA kernels + C ABI (include/rt_b200.h), sm_100a only
  vulkan-raytracing_b200/librt_host.so   asset loading / camera (include/rt_host.h)
  vulkan-raytracing_b200/rt_headless     headless C++ host replacing the reference's main()
  oracle/liboracle.so                    CPU checker (test infrastructure)
  oracle/_ref/*                          pieces of the real reference, only where its checkout exists
  tests/emu/libemu.so                    serial CPU build of the device functions (tests only)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("RT_REFERENCE_DIR", "/root/reference")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",                    # numeric contract: fused ops only where fmaf() is written
    "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall,-Wno-unused-function",
    "--expt-relaxed-constexpr", "-Xptxas", "-v",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def run(cmd, **kw):
    print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if r.returncode != 0:
        print(r.stdout)
        raise RuntimeError(f"build step failed: {' '.join(cmd)}")
    return r.stdout


def build_cuda(force=False, verbose=False):
    csrc = os.path.join(HERE, "csrc")
    srcs = [os.path.join(csrc, f) for f in ("rt_api.cu", "rt_build.cu", "rt_sort.cu", "rt_normals.cu", "rt_render.cu")]
    deps = srcs + [os.path.join(csrc, f) for f in os.listdir(csrc)] + [os.path.join(ROOT, "include", "rt_b200.h")]
    out = os.path.join(HERE, "librt_b200.so")
    if not force and not _newer(out, deps):
        return out
    log = run(["nvcc", *NVCC_FLAGS, "-shared", "-o", out, *srcs, "-lcudart"])
    with open(os.path.join(HERE, "csrc", "ptxas.log"), "w") as f:
        f.write(log)
    if verbose:
        print(log)
    return out


def build_host(force=False):
    src = os.path.join(HERE, "host", "scene_io.cpp")
    deps = [src] + [os.path.join(HERE, "host", f) for f in os.listdir(os.path.join(HERE, "host"))]
    out = os.path.join(HERE, "librt_host.so")
    have_stb = os.path.exists(os.path.join(REF, "include", "stb_image.h"))
    if not force and not _newer(out, deps):
        return out
    if not have_stb and os.path.exists(out):
        return out  # keep a prebuilt library that has the JPEG decoder
    cmd = ["g++", "-O2", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-w", "-o", out, src]
    if have_stb:
        cmd += ["-DRTH_HAVE_STB", "-I", os.path.join(REF, "include")]
    run(cmd)
    return out


def build_headless(force=False):
    """rt_headless: the C++ host that replaces the reference's main() (links both C ABIs)."""
    src = os.path.join(HERE, "host", "rt_headless.cpp")
    out = os.path.join(HERE, "rt_headless")
    deps = [src, os.path.join(HERE, "host", "config.hpp"), os.path.join(ROOT, "include", "rt_b200.h"),
            os.path.join(ROOT, "include", "rt_host.h")]
    if not force and not _newer(out, deps):
        return out
    run(["g++", "-O2", "-std=c++17", "-Wall", "-o", out, src, "-L", HERE, "-l:librt_b200.so", "-l:librt_host.so",
         "-Wl,-rpath,$ORIGIN", "-Wl,-rpath-link," + HERE])
    return out


def build_oracle(force=False):
    d = os.path.join(ROOT, "oracle")
    out = os.path.join(d, "liboracle.so")
    if force or _newer(out, [os.path.join(d, "oracle.cpp"), os.path.join(ROOT, "include", "rt_b200.h")]):
        run(["make", "-C", d, "liboracle.so", "-B"])
    out64 = os.path.join(d, "liboracle64.so")
    if force or _newer(out64, [os.path.join(d, "oracle64.cpp"), os.path.join(ROOT, "include", "rt_b200.h")]):
        run(["make", "-C", d, "liboracle64.so", "-B"])
    if os.path.isdir(os.path.join(REF, "src")) and os.path.exists(os.path.join(d, "ref_build.sh")):
        run(["sh", os.path.join(d, "ref_build.sh"), REF], cwd=d)
    return out


def build_emu(force=False):
    d = os.path.join(ROOT, "tests", "emu")
    src = os.path.join(d, "emu.cpp")
    if not os.path.exists(src):
        return None
    out = os.path.join(d, "libemu.so")
    csrc = os.path.join(HERE, "csrc")
    deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")]
    if force or _newer(out, deps):
        run(["g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-mfma", "-fPIC", "-shared", "-Wall",
             "-Wno-unused-function", "-I", csrc, "-I", os.path.join(ROOT, "include"), "-o", out, src])
    return out


def build_all(force=False, verbose=False):
    build_host(force)
    build_oracle(force)
    build_emu(force)
    build_cuda(force, verbose)
    build_headless(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
