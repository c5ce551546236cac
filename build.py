#!/usr/bin/env python
"""Build every native artefact of rabit-b200 IN-TREE (lib/ and build/).

    python build.py            # host + sm_100a device code (needs nvcc), test programs
    python build.py --no-cuda  # host-only libraries (device_stub.cc instead of the kernels)
    python build.py --clean

Outputs (all git-ignored, all travel to the GPU box with the gpurun snapshot):
    lib/librabit.so        default engine: robust      (reference librabit)
    lib/librabit_mock.so   default engine: mock        (reference librabit_mock)
    lib/librabit_base.so   default engine: base        (reference librabit_base)
    lib/librabit_empty.so  default engine: empty       (reference librabit_empty)
    lib/librabit_nccl.so   default engine: nccl        (the library-backed comparator; reference librabit_mpi)
    build/bin/*            guide + test programs (basic, broadcast, lazy_allreduce,
                           model_recover, local_recover, lazy_recover, speed_test, ...)

The four libraries are the same objects; only engine.cc is compiled four times with a
different RABIT_DEFAULT_ENGINE (the reference builds them from different engine_*.cc
files, CMakeLists.txt:29-34).  CUDA code is compiled for sm_100a only:
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo
"""
import argparse
import concurrent.futures
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
HOST_SRCS = ["comm", "engine_base", "engine_robust", "engine_recover", "engine_ops",
             "engine_mock", "result_store", "c_api", "fake_device"]
ENGINES = {"librabit.so": "robust", "librabit_mock.so": "mock",
           "librabit_base.so": "base", "librabit_empty.so": "empty", "librabit_nccl.so": "nccl"}
CXXFLAGS = ["-std=c++17", "-O3", "-g", "-Wall", "-fPIC", "-pthread", "-I" + os.path.join(ROOT, "include")]
NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def run(cmd, quiet=False):
    if not quiet:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        sys.stderr.write(r.stdout.decode(errors="replace"))
        raise RuntimeError("command failed: " + " ".join(cmd))
    out = r.stdout.decode(errors="replace")
    if out.strip() and not quiet:
        print(out)
    return out


def find_nvcc():
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    return None


def headers():
    hs = []
    for pat in ("include/rabit/*.h", "include/rabit/internal/*.h", "include/*.h", "src/host/*.h",
                "src/device/*.h", "src/device/*.cuh", "src/device/*.inc", "test/*.h"):
        hs += glob.glob(os.path.join(ROOT, pat))
    return hs


def build(cuda=True, verbose=False, programs=True):
    nvcc = find_nvcc() if cuda else None
    if cuda and nvcc is None:
        raise RuntimeError("nvcc not found; pass --no-cuda for a host-only build")
    obj_dir = os.path.join(ROOT, "build", "obj_cuda" if cuda else "obj_host")
    # the two flavours share lib/: remember which one is there so that switching always relinks
    flavour = os.path.join(ROOT, "lib", ".flavour")
    want = "cuda" if cuda else "host"
    have = open(flavour).read().strip() if os.path.exists(flavour) else ""
    relink = have != want
    os.makedirs(obj_dir, exist_ok=True)
    os.makedirs(os.path.join(ROOT, "lib"), exist_ok=True)
    hdrs = headers()
    jobs = []   # (cmd, output)

    def cc(src, obj, extra=()):
        if newer([src] + hdrs, obj):
            jobs.append((["g++"] + CXXFLAGS + list(extra) + ["-c", src, "-o", obj], obj))

    objs = []
    for name in HOST_SRCS:
        o = os.path.join(obj_dir, name + ".o")
        cc(os.path.join(ROOT, "src/host", name + ".cc"), o)
        objs.append(o)
    engine_objs = {}
    for lib, eng in ENGINES.items():
        o = os.path.join(obj_dir, "select_%s.o" % eng)
        cc(os.path.join(ROOT, "src/host/engine.cc"), o, ['-DRABIT_DEFAULT_ENGINE="%s"' % eng])
        engine_objs[lib] = o
    if cuda:
        for src in sorted(glob.glob(os.path.join(ROOT, "src/device/*.cu"))):
            o = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".cu.o")
            if newer([src] + hdrs, o):
                jobs.append(([nvcc] + NVCC_ARCH + ["-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
                              "-Xcompiler", "-fPIC,-pthread,-Wall", "-Xptxas", "-v",
                              "-I" + os.path.join(ROOT, "include"), "-c", src, "-o", o], o))
            objs.append(o)
    else:
        o = os.path.join(obj_dir, "device_stub.o")
        cc(os.path.join(ROOT, "src/device/device_stub.cc"), o)
        objs.append(o)

    with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        futs = [ex.submit(run, cmd, not verbose) for cmd, _ in jobs]
        for f in futs:
            out = f.result()
            if cuda and "ptxas info" in out:
                with open(os.path.join(ROOT, "build", "ptxas.log"), "a") as fo:
                    fo.write(out)

    link_tail = ["-pthread", "-ldl", "-lrt"]
    for lib, eo in engine_objs.items():
        target = os.path.join(ROOT, "lib", lib)
        if relink or newer(objs + [eo], target):
            if cuda:
                run([nvcc] + NVCC_ARCH + ["-shared", "-o", target] + objs + [eo] +
                    ["-cudart", "static", "-Xcompiler", "-pthread"] + ["-ldl", "-lrt"], not verbose)
            else:
                run(["g++", "-shared", "-o", target] + objs + [eo] + link_tail, not verbose)

    with open(flavour, "w") as fo:
        fo.write(want)
    if cuda:
        build_torch_extension(verbose)
    if programs:
        bin_dir = os.path.join(ROOT, "build", "bin")
        os.makedirs(bin_dir, exist_ok=True)
        pjobs = []
        srcs = [(s, os.path.basename(s)[:-3]) for s in
                sorted(glob.glob(os.path.join(ROOT, "guide/*.cc")) + glob.glob(os.path.join(ROOT, "test/*.cc")))]
        # the harness of bench.py's reference arm, unchanged, against THIS library: the same program on both sides
        srcs.append((os.path.join(ROOT, "baseline", "ref_speed.cc"), "ours_speed"))
        for src, name in srcs:
            # programs that use mock= link the mock library, like the reference's test/Makefile
            lib = "rabit_mock"
            exe = os.path.join(bin_dir, name)
            extra = []
            if name.endswith("_cuda"):
                # programs that own device buffers: CUDA runtime API from plain C++
                if not cuda:
                    continue
                cuda_home = os.path.dirname(os.path.dirname(nvcc))
                extra = ["-I" + os.path.join(cuda_home, "include"), "-L" + os.path.join(cuda_home, "lib64"),
                         "-lcudart_static", "-ldl", "-lrt"]
            if newer([src] + hdrs + [os.path.join(ROOT, "lib", "lib%s.so" % lib)], exe):
                pjobs.append(["g++"] + CXXFLAGS + [src, "-o", exe, "-L" + os.path.join(ROOT, "lib"),
                                                   "-l" + lib, "-Wl,-rpath," + os.path.join("$ORIGIN", "..", "..", "lib")] + extra)
        with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
            for f in [ex.submit(run, c, not verbose) for c in pjobs]:
                f.result()
    return True


def build_torch_extension(verbose=False):
    """rabit_b200/_torch_ext*.so: the pybind11 / libtorch binding (src/python/torch_ext.cc). Host C++ only; it does not
    link the engine library (symbols are resolved from whichever lib/librabit*.so rabit.py loaded). Skipped with a note
    when torch is not importable: torch_binding.py then stays on its ctypes path."""
    try:
        import sysconfig
        import torch  # noqa: F401
        from torch.utils import cpp_extension as ce
        import pybind11
    except Exception as e:  # pylint: disable=broad-except
        print("torch extension skipped: %s" % e)
        return False
    src = os.path.join(ROOT, "src", "python", "torch_ext.cc")
    target = os.path.join(ROOT, "rabit_b200", "_torch_ext" + sysconfig.get_config_var("EXT_SUFFIX"))
    if not newer([src], target):
        return True
    import torch
    incs = ce.include_paths() + [pybind11.get_include(), sysconfig.get_paths()["include"],
                                 os.path.join(os.path.dirname(os.path.dirname(find_nvcc())), "include")]
    libdirs = ce.library_paths()
    cmd = (["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-Wno-unused-function",
            "-DTORCH_EXTENSION_NAME=_torch_ext", "-DTORCH_API_INCLUDE_EXTENSION_H",
            "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)] +
           ["-I" + i for i in incs] + [src, "-o", target] + ["-L" + d for d in libdirs] +
           ["-Wl,-rpath," + d for d in libdirs] + ["-ltorch_python", "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda", "-ldl"])
    run(cmd, not verbose)
    return True


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-cuda", action="store_true")
    ap.add_argument("--clean", action="store_true")
    ap.add_argument("-v", "--verbose", action="store_true")
    args = ap.parse_args()
    if args.clean:
        shutil.rmtree(os.path.join(ROOT, "build"), ignore_errors=True)
        for f in glob.glob(os.path.join(ROOT, "lib", "*.so")):
            os.remove(f)
        return
    build(cuda=not args.no_cuda, verbose=args.verbose)
    print("build ok")


if __name__ == "__main__":
    main()
