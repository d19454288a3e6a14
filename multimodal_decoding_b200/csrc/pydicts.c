/* Host helper of the Python mirror (no CUDA): the reference returns one dict of 8 scores per vertex
 * (searchlight.py:114, :149); for a 163,842-vertex hemisphere building them is the largest host cost of an end-to-end call.
 * This fills a numpy object array with those dicts through the CPython API (pre-sized dicts, cached key hashes): about twice
 * as fast as the fastest pure-Python construction.  Built by multimodal_decoding_b200/build.py with gcc against Python.h into
 * libmmd_pyhelpers.so and loaded with ctypes.PyDLL (the GIL is held); the mirror falls back to Python when it is absent. */
#include <Python.h>

/* out: n slots of a numpy object array (each holds a reference, normally None); scores: (n x n_keys) float64, row-major */
PyObject* mmd_build_score_dicts(PyObject* keys, const double* scores, Py_ssize_t n, PyObject** out) {
    if (!PyTuple_Check(keys)) {
        PyErr_SetString(PyExc_TypeError, "keys must be a tuple of str");
        return NULL;
    }
    const Py_ssize_t nk = PyTuple_GET_SIZE(keys);
    Py_hash_t hashes[64];
    if (nk > 64) {
        PyErr_SetString(PyExc_ValueError, "too many keys");
        return NULL;
    }
    for (Py_ssize_t j = 0; j < nk; ++j) {
        hashes[j] = PyObject_Hash(PyTuple_GET_ITEM(keys, j));
        if (hashes[j] == -1) return NULL;
    }
    for (Py_ssize_t i = 0; i < n; ++i) {
        PyObject* d = _PyDict_NewPresized(nk);
        if (!d) return NULL;
        for (Py_ssize_t j = 0; j < nk; ++j) {
            PyObject* v = PyFloat_FromDouble(scores[i * nk + j]);
            if (!v || _PyDict_SetItem_KnownHash(d, PyTuple_GET_ITEM(keys, j), v, hashes[j]) < 0) {
                Py_XDECREF(v);
                Py_DECREF(d);
                return NULL;
            }
            Py_DECREF(v);
        }
        PyObject* old = out[i];
        out[i] = d;
        Py_XDECREF(old);
    }
    Py_RETURN_NONE;
}

/* Two-phase construction of the same dicts, so that the allocation-heavy part can run on a helper thread WHILE the GPU
 * computes the scores (the ctypes call into libmmd_b200.so releases the GIL):
 *   mmd_prebuild_score_dicts  n dicts with the final keys, every value a fresh float object holding 0.0; the float objects'
 *                             addresses go to `slots` (n x n_keys pointers, row-major).  The dicts are stored into `out`.
 *   mmd_fill_score_floats     writes the scores into those float objects.
 * Writing `ob_fval` of a float is only sound because nothing but the dict built here references the object and nobody has
 * seen the dict yet (CPython does not cache or share float objects and a float does not cache its hash); the mirror fills
 * the values before it hands the array out and never calls the second phase twice on the same slots. */
PyObject* mmd_prebuild_score_dicts(PyObject* keys, Py_ssize_t n, PyObject** out, PyObject** slots) {
    if (!PyTuple_Check(keys)) {
        PyErr_SetString(PyExc_TypeError, "keys must be a tuple of str");
        return NULL;
    }
    const Py_ssize_t nk = PyTuple_GET_SIZE(keys);
    Py_hash_t hashes[64];
    if (nk > 64) {
        PyErr_SetString(PyExc_ValueError, "too many keys");
        return NULL;
    }
    for (Py_ssize_t j = 0; j < nk; ++j) {
        hashes[j] = PyObject_Hash(PyTuple_GET_ITEM(keys, j));
        if (hashes[j] == -1) return NULL;
    }
    for (Py_ssize_t i = 0; i < n; ++i) {
        PyObject* d = _PyDict_NewPresized(nk);
        if (!d) return NULL;
        for (Py_ssize_t j = 0; j < nk; ++j) {
            PyObject* v = PyFloat_FromDouble(0.0);
            if (!v || _PyDict_SetItem_KnownHash(d, PyTuple_GET_ITEM(keys, j), v, hashes[j]) < 0) {
                Py_XDECREF(v);
                Py_DECREF(d);
                return NULL;
            }
            slots[i * nk + j] = v;                 /* borrowed: the dict holds the only reference */
            Py_DECREF(v);
        }
        PyObject* old = out[i];
        out[i] = d;
        Py_XDECREF(old);
    }
    Py_RETURN_NONE;
}

PyObject* mmd_fill_score_floats(PyObject** slots, const double* scores, Py_ssize_t n_values) {
    for (Py_ssize_t i = 0; i < n_values; ++i) {
        if (!slots[i] || !PyFloat_CheckExact(slots[i]) || Py_REFCNT(slots[i]) != 1) {
            PyErr_SetString(PyExc_RuntimeError, "score slot is not an unshared float object");
            return NULL;
        }
        ((PyFloatObject*)slots[i])->ob_fval = scores[i];
    }
    Py_RETURN_NONE;
}
