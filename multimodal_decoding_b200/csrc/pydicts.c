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
