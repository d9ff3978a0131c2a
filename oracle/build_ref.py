"""Build the UNMODIFIED reference extensions into oracle/_ref/ (test infrastructure only).

This is test infrastructure, not product code.  It compiles the reference's own
sources where they lie under /root/reference (nothing is copied into this repo)
with ONE extra flag, `-gencode=arch=compute_100,code=sm_100`, because the
reference's setup.py (methods/dpvo/setup.py:10-18) ships no image that runs on a
B200.  Outputs go only into oracle/_ref/ (git-ignored, NOT gpurun-ignored, so the
built .so files travel to the GPU box where /root/reference does not exist).

Modules built (names are the ones the reference imports by bare name):
  cuda_corr          methods/dpvo/dpvo/altcorr/{correlation.cpp,correlation_kernel.cu}
  cuda_ba            methods/dpvo/dpvo/fastba/{ba.cpp,ba_cuda.cu,block_e.cu}
  lietorch_backends  methods/dpvo/dpvo/lietorch/src/{lietorch.cpp,lietorch_gpu.cu,lietorch_cpu.cpp}

Usage:  python oracle/build_ref.py [cuda_corr cuda_ba lietorch_backends]
Only tests/, __graft_entry__.smoke() and bench.py's reference legs may load the
results (see oracle/README.md).
"""
import os
import sys
import shutil

REF = "/root/reference/methods/dpvo"
EIGEN = "/root/reference/modules/eigen-3.4.0"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

SPECS = {
    "cuda_corr": dict(
        sources=["dpvo/altcorr/correlation.cpp", "dpvo/altcorr/correlation_kernel.cu"],
        includes=[]),
    "cuda_ba": dict(
        sources=["dpvo/fastba/ba.cpp", "dpvo/fastba/ba_cuda.cu", "dpvo/fastba/block_e.cu"],
        includes=[EIGEN]),
    "lietorch_backends": dict(
        sources=["dpvo/lietorch/src/lietorch.cpp", "dpvo/lietorch/src/lietorch_gpu.cu",
                 "dpvo/lietorch/src/lietorch_cpu.cpp"],
        includes=[os.path.join(REF, "dpvo/lietorch/include"), EIGEN]),
}


def build_one(name):
    from torch.utils.cpp_extension import load
    spec = SPECS[name]
    bdir = os.path.join(OUT, "_build_" + name)
    os.makedirs(bdir, exist_ok=True)
    load(name=name,
         sources=[os.path.join(REF, s) for s in spec["sources"]],
         extra_include_paths=spec["includes"],
         extra_cflags=["-O3"],
         extra_cuda_cflags=["-O3", "-Wno-deprecated-gpu-targets",
                            "-gencode=arch=compute_100,code=sm_100"],
         build_directory=bdir, verbose=False, is_python_module=False)
    so = os.path.join(bdir, name + ".so")
    shutil.copy2(so, os.path.join(OUT, name + ".so"))
    shutil.rmtree(bdir, ignore_errors=True)
    print("built", os.path.join(OUT, name + ".so"), flush=True)


def main(argv):
    if not os.path.isdir(REF):
        print("reference tree absent (GPU box): using prebuilt oracle/_ref/*.so")
        return 0
    os.makedirs(OUT, exist_ok=True)
    names = argv or list(SPECS)
    for n in names:
        if os.path.exists(os.path.join(OUT, n + ".so")) and not os.environ.get("REF_REBUILD"):
            print("up to date:", n)
            continue
        build_one(n)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
