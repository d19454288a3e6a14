"""Compile the CUDA sources under csrc/ into the in-tree C-ABI library libmmd_b200.so (sm_100a only)."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmmd_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-pthread"]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = glob.glob(os.path.join(CSRC, "*")) + [os.path.join(HERE, "..", "include", "mmd_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, variant=None, defines=()):
    """`variant`/`defines`: development builds of kernel variants (libmmd_b200.<variant>.so, extra -D flags),
    loaded through the MMD_B200_LIB override of _lib.py; the product library is the default build."""
    lib = LIB if not variant else os.path.join(HERE, f"libmmd_b200.{variant}.so")
    if not variant and not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    procs = []
    for src in sorted(glob.glob(os.path.join(CSRC, "*.cu"))):
        obj = src[:-3] + (f".{variant}.o" if variant else ".o")
        cmd = ([nvcc] + NVCC_FLAGS + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else [])
               + ["-c", src, "-o", obj])
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(f"--- {os.path.basename(src)}\n{out}")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs + ["-lcudart"])
    if not variant:
        build_pyhelpers()
    return lib


PYHELPERS = os.path.join(HERE, "libmmd_pyhelpers.so")


def build_pyhelpers():
    """The CPython-API helper of the mirror (csrc/pydicts.c, plain C, no CUDA).  Optional: without Python.h the mirror
    builds its score dicts in Python."""
    import sysconfig
    inc = sysconfig.get_paths()["include"]
    if not os.path.exists(os.path.join(inc, "Python.h")):
        return None
    try:
        subprocess.check_call([os.environ.get("CC", "gcc"), "-O2", "-shared", "-fPIC", f"-I{inc}",
                               os.path.join(CSRC, "pydicts.c"), "-o", PYHELPERS])
    except (OSError, subprocess.CalledProcessError):
        return None
    return PYHELPERS


if __name__ == "__main__":
    var = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--variant=")]
    defs = [a[2:] for a in sys.argv if a.startswith("-D")]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, variant=var[0] if var else None, defines=defs))
