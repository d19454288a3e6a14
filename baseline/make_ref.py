"""Recipe for the CPU arm of bench.py: the UNMODIFIED reference files of the searchlight path, copied from /root/reference
into the git-ignored baseline/_ref/ (it travels to the GPU box with the snapshot, like the built .so files), plus on-disk
stub packages for the three imports those files make at module top but never use on this path and that are not installed
here (nilearn, nibabel, matplotlib).  `GroupIterator` -- the one nilearn name the path does use
(searchlight.py:130: contiguous chunks of centres per joblib worker) -- is restated from its documented behaviour.

    python baseline/make_ref.py          # run by __graft_entry__.build() when /root/reference is present

Nothing under baseline/_ref is product or test code: only `bench.py --impl reference` / its cpu_baseline leg import it.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("MMD_REFERENCE_ROOT", "/root/reference")
FILES = ["analyses/decoding/searchlight/searchlight.py", "analyses/decoding/ridge_regression_decoding.py", "eval.py", "data.py",
         "utils.py"]
STUBS = {
    "nilearn/__init__.py": "",
    "nilearn/datasets.py": "",
    "nilearn/decoding/__init__.py": "",
    "nilearn/decoding/searchlight.py": (
        "import numpy as np\n\n\n"
        "class GroupIterator:\n"
        "    \"\"\"nilearn.decoding.searchlight.GroupIterator: contiguous chunks of range(n_features), one per job.\"\"\"\n\n"
        "    def __init__(self, n_features, n_jobs=1):\n"
        "        self.n_features, self.n_jobs = n_features, n_jobs\n\n"
        "    def __iter__(self):\n"
        "        yield from np.array_split(np.arange(self.n_features), self.n_jobs)\n"),
    "nilearn/surface/__init__.py": "from . import surface  # noqa: F401\n",
    "nilearn/surface/surface.py": "",
    "nibabel/__init__.py": "GiftiImage = object\n",
    "nibabel/gifti.py": "GiftiDataArray = object\n",
    "nibabel/nifti1.py": "intent_codes = None\ndata_type_codes = None\n",
    "matplotlib/__init__.py": "",
    "matplotlib/pyplot.py": "",
}


def available():
    return all(os.path.exists(os.path.join(REF, f)) for f in FILES)


def make(force=False):
    if not os.path.isdir(REFERENCE_ROOT):
        return available()
    if available() and not force:
        return True
    for f in FILES:
        dst = os.path.join(REF, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REFERENCE_ROOT, f), dst)
    for f, body in STUBS.items():
        try:
            __import__(f.split("/")[0])
            continue                      # the real package is installed: no stub
        except ImportError:
            pass
        dst = os.path.join(REF, "_stubs", f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(dst, "w") as fh:
            fh.write(body)
    return True


def activate():
    """Put baseline/_ref (and its stubs) in front of sys.path and PYTHONPATH (joblib's loky workers are fresh interpreters
    that import the reference modules by name) and return the reference's searchlight module."""
    if not available():
        raise RuntimeError("baseline/_ref is missing: run `python baseline/make_ref.py` where /root/reference exists")
    paths = [REF, os.path.join(REF, "_stubs")]
    for name in ("eval", "data", "utils", "analyses"):
        for key in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
            del sys.modules[key]
    for p in reversed(paths):
        sys.path.insert(0, p)
    os.environ["PYTHONPATH"] = os.pathsep.join(paths + [os.environ.get("PYTHONPATH", "")]).rstrip(os.pathsep)
    from analyses.decoding.searchlight import searchlight as ref_sl
    return ref_sl


if __name__ == "__main__":
    print("baseline/_ref", "ready" if make(force="--force" in sys.argv) else "unavailable (no /root/reference here)")
