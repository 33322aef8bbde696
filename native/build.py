"""Build the native inference runtime: native/lib/libveles_native.so and native/lib/test_native.
Needs veles_b200/lib/libveles_b200.so (built by veles_b200.ops._lib.build)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)


def build(force=False):
    from veles_b200.ops import _lib
    kernel_lib = _lib.build()
    nvcc = _lib.find_nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found")
    out_dir = os.path.join(HERE, "lib")
    os.makedirs(out_dir, exist_ok=True)
    srcs = [os.path.join(HERE, "src", f) for f in sorted(os.listdir(os.path.join(HERE, "src"))) if f.endswith(".cc")]
    lib = os.path.join(out_dir, "libveles_native.so")
    newest = max(os.path.getmtime(p) for p in srcs + [os.path.join(HERE, "include", "veles", f)
                                                      for f in os.listdir(os.path.join(HERE, "include", "veles"))])
    common = ["-std=c++17", "-O2", "-I", os.path.join(HERE, "include"), "-Xcompiler", "-fPIC",
              "-gencode", "arch=compute_100a,code=sm_100a"]
    link = ["-L", os.path.dirname(kernel_lib), "-lveles_b200", "-lz", "-lcudart",
            "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../../veles_b200/lib"]
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < newest:
        subprocess.check_call([nvcc, "-x", "cu"] + common + ["-shared", "-o", lib] + srcs + link)
    test = os.path.join(out_dir, "test_native")
    test_src = os.path.join(HERE, "tests", "test_native.cc")
    if force or not os.path.exists(test) or os.path.getmtime(test) < max(newest, os.path.getmtime(test_src)):
        subprocess.check_call([nvcc, "-x", "cu"] + common + ["-o", test, test_src, "-L", out_dir, "-lveles_native"] +
                              link + ["-Xlinker", "-rpath", "-Xlinker", "$ORIGIN"])
    build_simd(force)
    return lib, test


def build_simd(force=False):
    """native/lib/libveles_simd.so: the host signal / array primitives (no CUDA dependency, plain g++)."""
    out_dir = os.path.join(HERE, "lib")
    os.makedirs(out_dir, exist_ok=True)
    src_dir = os.path.join(HERE, "simd")
    srcs = [os.path.join(src_dir, f) for f in sorted(os.listdir(src_dir)) if f.endswith(".cc")]
    hdrs = [os.path.join(HERE, "include", "veles", f) for f in ("simd.h", "wavelet_tables.h")]
    lib = os.path.join(out_dir, "libveles_simd.so")
    newest = max(os.path.getmtime(p) for p in srcs + hdrs)
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < newest:
        subprocess.check_call(["g++", "-std=c++17", "-O3", "-fno-math-errno", "-fPIC", "-shared",
                               "-I", os.path.join(HERE, "include"), "-o", lib] + srcs)
    test = os.path.join(out_dir, "test_simd")
    test_src = os.path.join(HERE, "tests", "test_simd.cc")
    if os.path.exists(test_src) and (force or not os.path.exists(test) or
                                     os.path.getmtime(test) < max(newest, os.path.getmtime(test_src))):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-I", os.path.join(HERE, "include"), "-o", test, test_src,
                               "-L", out_dir, "-lveles_simd", "-Wl,-rpath,$ORIGIN"])
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
