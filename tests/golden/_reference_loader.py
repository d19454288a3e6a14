"""Import the UNMODIFIED reference from /root/reference inside the build container.

Used only by `make_golden.py` (fixture generation) — never at test time, never on the GPU box
(where /root/reference does not exist).  nilearn, nibabel and matplotlib are not installed here and
are only needed by the reference at import time (`utils.py:3-8`, `data.py:9`,
`searchlight.py:8-10`), so they are replaced by empty stub modules; `GroupIterator` is restated
from its documented behaviour (contiguous `np.array_split` chunks).
"""
import sys
import types

import numpy as np

REFERENCE_ROOT = "/root/reference"


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class GroupIterator:
    def __init__(self, n_features, n_jobs=1):
        self.n_features, self.n_jobs = n_features, n_jobs

    def __iter__(self):
        yield from np.array_split(np.arange(self.n_features), self.n_jobs)


def load_reference():
    nl = _stub("nilearn")
    nl.datasets = _stub("nilearn.datasets")
    _stub("nilearn.decoding")
    _stub("nilearn.decoding.searchlight", GroupIterator=GroupIterator)
    surf = _stub("nilearn.surface")
    surf.surface = _stub("nilearn.surface.surface")
    _stub("nibabel", GiftiImage=object)
    _stub("nibabel.gifti", GiftiDataArray=object)
    _stub("nibabel.nifti1", intent_codes=None, data_type_codes=None)
    mpl = _stub("matplotlib")
    mpl.pyplot = _stub("matplotlib.pyplot")
    # the repo root also has eval.py/data.py/utils.py mirrors: make sure the reference wins here
    for name in ("eval", "data", "utils", "analyses"):
        for key in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
            del sys.modules[key]
    sys.path.insert(0, REFERENCE_ROOT)
    from analyses.decoding.searchlight import searchlight as ref_sl
    import eval as ref_eval
    import data as ref_data
    import analyses.decoding.ridge_regression_decoding as ref_ridge
    return ref_sl, ref_eval, ref_data, ref_ridge
