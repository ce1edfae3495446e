"""Recipe for oracle/_ref/: the reference's own tree_filter CUDA extension and Python glue -- TEST INFRASTRUCTURE ONLY.

Runs only where /root/reference is mounted (this container).  Outputs go to the git-ignored oracle/_ref/, which
travels to the GPU box with the repository snapshot:

  oracle/_ref/ltf/src/...                 working copy of lib_tree_filter/src (tree_filter.cpp, mst/, bfs/, refine/)
  oracle/_ref/tree_filter_cuda_ref.so     those sources built for sm_100a.  The ONLY edit is the removal of the line
                                          `#include <THC/THC.h>` (refine.cu:12, bfs.cu:12, mst.cu:9): that header no
                                          longer exists in torch >= 1.11 and nothing from it is used.  BASELINE.md §4.2
                                          names this rebuild as the labelled substitute for the shipped wheel
                                          (cp39 / CUDA 11.0 / torch 1.10 / sm_75), which cannot load here.
  oracle/_ref/py/...                      verbatim copies of gate_crf_loss.py, flower_common.py,
                                          flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py, networks/, dataloaders/ and
                                          the lib_tree_filter Python glue, in the reference's own package layout, so
                                          the GPU box can import the reference's call sites unchanged.

Nothing here is read by the product package; tests/, bench.py's reference legs and smoke() are the only users.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_CODE = "/root/reference/code"
REF_LTF = os.path.join(REF_CODE, "utils/TreeEnergyLoss/kernels/lib_tree_filter")
OUT = os.path.join(HERE, "_ref")
SO = os.path.join(OUT, "tree_filter_cuda_ref.so")

PY_FILES = [
    "utils/gate_crf_loss.py",
    "utils/losses.py",
    "utils/metrics.py",
    "utils/ramps.py",
    "val_2D.py",
    "flower_common.py",
    "flower_pCE_2D_GateCRFMsacleTreeEnergyLoss_Ours.py",
    "utils/TreeEnergyLoss/kernels/lib_tree_filter/modules/tree_filter.py",
    "utils/TreeEnergyLoss/kernels/lib_tree_filter/functions/bfs.py",
    "utils/TreeEnergyLoss/kernels/lib_tree_filter/functions/mst.py",
    "utils/TreeEnergyLoss/kernels/lib_tree_filter/functions/refine.py",
]
PY_DIRS = ["networks", "dataloaders"]


def available():
    return os.path.isdir(REF_LTF)


def _copy_sources():
    dst = os.path.join(OUT, "ltf", "src")
    if os.path.isdir(dst):
        _writable(dst)
        shutil.rmtree(dst)
    shutil.copytree(os.path.join(REF_LTF, "src"), dst)
    _writable(dst)
    stripped = 0
    for sub in ("refine/refine.cu", "bfs/bfs.cu", "mst/mst.cu"):
        path = os.path.join(dst, sub)
        os.chmod(path, 0o644)
        lines = open(path).read().split("\n")
        kept = [ln for ln in lines if ln.strip() != "#include <THC/THC.h>"]
        stripped += len(lines) - len(kept)
        open(path, "w").write("\n".join(kept))
    assert stripped == 3, f"expected to strip 3 include lines, stripped {stripped}"
    return dst


def _writable(root):
    for d, subdirs, files in os.walk(root):
        os.chmod(d, 0o755)
        for f in files:
            os.chmod(os.path.join(d, f), 0o644)


def _copy_python():
    dst = os.path.join(OUT, "py")
    if os.path.isdir(dst):
        _writable(dst)
        shutil.rmtree(dst)
    for rel in PY_FILES:
        os.makedirs(os.path.dirname(os.path.join(dst, rel)), exist_ok=True)
        shutil.copyfile(os.path.join(REF_CODE, rel), os.path.join(dst, rel))
    for d in PY_DIRS:
        shutil.copytree(os.path.join(REF_CODE, d), os.path.join(dst, d),
                        ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    _writable(dst)


def build(force=False):
    """Build oracle/_ref/tree_filter_cuda_ref.so and copy the reference Python.  No-op without /root/reference."""
    if not available():
        return False
    os.makedirs(OUT, exist_ok=True)
    _copy_python()
    srcs = [os.path.join(REF_LTF, "src", s) for s in
            ("tree_filter.cpp", "mst/mst.cu", "mst/boruvka.cpp", "bfs/bfs.cu", "refine/refine.cu")]
    if not force and os.path.exists(SO) and all(os.path.getmtime(SO) >= os.path.getmtime(s) for s in srcs) \
            and os.path.getmtime(SO) >= os.path.getmtime(__file__):
        return True
    src = _copy_sources()
    from torch.utils import cpp_extension as ce
    inc = [f"-I{p}" for p in ce.include_paths("cuda")] + [f"-I{src}"]
    import sysconfig
    inc.append(f"-I{sysconfig.get_paths()['include']}")
    common = ["-O3", "-std=c++17", "-DTORCH_EXTENSION_NAME=tree_filter_cuda_ref", "-DTORCH_API_INCLUDE_EXTENSION_H",
              "-D_GLIBCXX_USE_CXX11_ABI=1"]
    bdir = os.path.join(OUT, "ltf", "build")
    os.makedirs(bdir, exist_ok=True)
    objs = []
    procs = []
    for rel in ("tree_filter.cpp", "mst/boruvka.cpp", "mst/mst.cu", "bfs/bfs.cu", "refine/refine.cu"):
        obj = os.path.join(bdir, rel.replace("/", "_") + ".o")
        objs.append(obj)
        if rel.endswith(".cu"):
            cmd = ["/usr/local/cuda/bin/nvcc", *common, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
                   "-Xcompiler", "-fPIC", "-w", *inc, "-c", os.path.join(src, rel), "-o", obj]
        else:
            cmd = ["g++", *common, "-fPIC", "-w", *inc, "-c", os.path.join(src, rel), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("reference build failed: " + " ".join(cmd))
    libdirs = ce.library_paths("cuda")
    link = ["g++", "-shared", "-o", SO, *objs, *[f"-L{d}" for d in libdirs], *[f"-Wl,-rpath,{d}" for d in libdirs],
            "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart", "-pthread"]
    subprocess.run(link, check=True)
    return True


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv)
    print("built" if ok else "reference not mounted; nothing built", SO if ok else "")
