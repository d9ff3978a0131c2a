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
  dpviewer_cuda      modules/DPViewer/dpviewer/viewer_cuda.cu behind oracle/viewer_ref_binding.cpp (this repo's
                     pybind wrapper: the reference only reaches these kernels through its OpenGL viewer class)

It also INSTALLS the reference's pure-Python package (what `pip install --target` of the
reference would place in site-packages: methods/dpvo/dpvo/**/*.py, utils/utils.py and
config/default.yaml) into oracle/_ref/site/, so that the unmodified `dpvo.dpvo.DPVO`,
`dpvo.net.VONet` and the loss of train.py can run on the GPU box on top of either the
reference extensions above or this repo's drop-in modules (tools/refdpvo.py).  The
install is git-ignored like the .so files; no reference source enters the history.

Usage:  python oracle/build_ref.py [cuda_corr cuda_ba lietorch_backends site]
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
    # the viewer's two plain CUDA kernels, headless (the window / GL interop in viewer.cpp is not built)
    "dpviewer_cuda": dict(
        sources=[os.path.join(HERE, "viewer_ref_binding.cpp"), "/root/reference/modules/DPViewer/dpviewer/viewer_cuda.cu"],
        includes=["/root/reference/modules/DPViewer/dpviewer"]),
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
         sources=[s if os.path.isabs(s) else os.path.join(REF, s) for s in spec["sources"]],
         extra_include_paths=spec["includes"],
         extra_cflags=["-O3"],
         extra_cuda_cflags=["-O3", "-Wno-deprecated-gpu-targets",
                            "-gencode=arch=compute_100,code=sm_100"],
         build_directory=bdir, verbose=False, is_python_module=False)
    so = os.path.join(bdir, name + ".so")
    shutil.copy2(so, os.path.join(OUT, name + ".so"))
    shutil.rmtree(bdir, ignore_errors=True)
    print("built", os.path.join(OUT, name + ".so"), flush=True)


def install_site():
    """Copy the reference's Python package (sources untouched) into oracle/_ref/site/."""
    site = os.path.join(OUT, "site")
    shutil.rmtree(site, ignore_errors=True)
    n = 0
    for base, sub in ((os.path.join(REF, "dpvo"), "dpvo"), (os.path.join(REF, "utils"), "utils")):
        for d, dirs, files in os.walk(base):
            dirs[:] = [x for x in dirs if x not in ("__pycache__", "include", "src", "retrieval", "test")]
            for f in files:
                if f.endswith(".py"):
                    rel = os.path.relpath(os.path.join(d, f), base)
                    dst = os.path.join(site, sub, rel)
                    os.makedirs(os.path.dirname(dst), exist_ok=True)
                    shutil.copy2(os.path.join(d, f), dst)
                    n += 1
    os.makedirs(os.path.join(site, "config"), exist_ok=True)
    shutil.copy2(os.path.join(REF, "config", "default.yaml"), os.path.join(site, "config", "default.yaml"))
    if not os.path.exists(os.path.join(site, "utils", "__init__.py")):
        open(os.path.join(site, "utils", "__init__.py"), "w").close()
    print(f"installed {n} reference .py files into {site}", flush=True)


def main(argv):
    if not os.path.isdir(REF):
        print("reference tree absent (GPU box): using prebuilt oracle/_ref/*.so")
        return 0
    os.makedirs(OUT, exist_ok=True)
    names = argv or (list(SPECS) + ["site"])
    if "site" in names:
        names = [n for n in names if n != "site"]
        if not os.path.isdir(os.path.join(OUT, "site", "dpvo")) or os.environ.get("REF_REBUILD"):
            install_site()
        else:
            print("up to date: site")
    for n in names:
        if os.path.exists(os.path.join(OUT, n + ".so")) and not os.environ.get("REF_REBUILD"):
            print("up to date:", n)
            continue
        build_one(n)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
