"""ctypes binding of the C ABI declared in include/mmd_b200.h (libmmd_b200.so, built in-tree)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MMD_B200_LIB: development override used to A/B kernel variants built with build.py --variant
LIB_PATH = os.environ.get("MMD_B200_LIB") or os.path.join(_HERE, "libmmd_b200.so")

N_SCORES = 8
N_STAGES = 10
N_COUNTERS = 8
N_NULL_SCORES = 6
EVAL_PAD = 144

EXPORTS = [
    "mmd_create", "mmd_destroy", "mmd_last_error", "mmd_version", "mmd_score_name", "mmd_knn", "mmd_radius_neighbors", "mmd_upload", "mmd_upload_active", "mmd_mark_vertices",
    "mmd_searchlight", "mmd_searchlight_i64", "mmd_searchlight_ragged", "mmd_predict_centre", "mmd_get_stats", "mmd_get_counters", "mmd_set_group_size", "mmd_reset_stats", "mmd_stage_name", "mmd_configure", "mmd_stream",
    "mmd_searchlight_null", "mmd_searchlight_null_pc", "mmd_searchlight_null_ragged", "mmd_searchlight_cv", "mmd_ridge_load", "mmd_ridge_cv", "mmd_ridge_fit", "mmd_ridge_predict", "mmd_cosine_distances", "mmd_dist_mat_to_pairwise_acc", "mmd_pairwise_accuracy", "mmd_pairwise_accuracy_static",
    "mmd_ttest_1samp", "mmd_tfce", "mmd_cluster_labels", "mmd_check_guards", "mmd_guard_selftest",
    "mmd_set_predict_mode", "mmd_set_syrk_mode", "mmd_enable_margins", "mmd_get_margins",
]

_lib = None


def load():
    """Load the shared library; raises if it has not been built (there is no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m multimodal_decoding_b200.build` "
            "(the searchlight has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, ip, i32, i64, dbl = C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.c_int64, C.c_double
    lib.mmd_create.argtypes = [i32, C.POINTER(vp)]
    lib.mmd_create.restype = i32
    lib.mmd_destroy.argtypes = [vp]
    lib.mmd_destroy.restype = None
    lib.mmd_last_error.argtypes = [vp]
    lib.mmd_last_error.restype = C.c_char_p
    lib.mmd_version.restype = C.c_char_p
    lib.mmd_check_guards.argtypes = [vp]
    lib.mmd_check_guards.restype = i32
    lib.mmd_guard_selftest.argtypes = [vp]
    lib.mmd_guard_selftest.restype = i32
    lib.mmd_stream.argtypes = [vp]
    lib.mmd_stream.restype = vp
    lib.mmd_score_name.argtypes = [i32]
    lib.mmd_score_name.restype = C.c_char_p
    lib.mmd_stage_name.argtypes = [i32]
    lib.mmd_stage_name.restype = C.c_char_p
    lib.mmd_knn.argtypes = [vp, vp, i32, i32, i32, i32, vp, vp]
    lib.mmd_knn.restype = i32
    lib.mmd_radius_neighbors.argtypes = [vp, vp, i32, dbl, i32, i32, vp, vp, vp]
    lib.mmd_radius_neighbors.restype = i32
    lib.mmd_upload.argtypes = [vp, vp, i32, i32, vp, i32, vp, i32, vp, i32, vp, vp, i32]
    lib.mmd_upload.restype = i32
    lib.mmd_upload_active.argtypes = [vp, vp, i32, i32, vp, i32, vp, i32, vp, i32, vp, vp, i32, vp, i32]
    lib.mmd_upload_active.restype = i32
    lib.mmd_mark_vertices.argtypes = [vp, i32, i64, i32, vp]
    lib.mmd_mark_vertices.restype = i32
    lib.mmd_searchlight.argtypes = [vp, vp, i32, i32, dbl, vp, vp]
    lib.mmd_searchlight.restype = i32
    lib.mmd_searchlight_i64.argtypes = [vp, vp, i32, i32, dbl, vp, vp]
    lib.mmd_searchlight_i64.restype = i32
    lib.mmd_searchlight_ragged.argtypes = [vp, vp, vp, vp, i32, dbl, vp, vp]
    lib.mmd_searchlight_ragged.restype = i32
    lib.mmd_searchlight_null.argtypes = [vp, vp, i32, i32, dbl, vp, vp, i32, vp, vp, vp]
    lib.mmd_searchlight_null.restype = i32
    lib.mmd_searchlight_null_pc.argtypes = [vp, vp, i32, i32, dbl, vp, vp, vp, i32, vp, vp, vp]
    lib.mmd_searchlight_null_pc.restype = i32
    lib.mmd_searchlight_null_ragged.argtypes = [vp, vp, vp, vp, i32, dbl, vp, vp, vp, i32, vp, vp, vp]
    lib.mmd_searchlight_null_ragged.restype = i32
    lib.mmd_searchlight_cv.argtypes = [vp, vp, i32, i32, vp, i32, vp, i32, vp, vp]
    lib.mmd_searchlight_cv.restype = i32
    lib.mmd_ridge_load.argtypes = [vp, vp, i32, i32, i32, vp, i32, i32]
    lib.mmd_ridge_load.restype = i32
    lib.mmd_ridge_cv.argtypes = [vp, vp, i32, vp, i32, vp]
    lib.mmd_ridge_cv.restype = i32
    lib.mmd_ridge_fit.argtypes = [vp, dbl, vp]
    lib.mmd_ridge_fit.restype = i32
    lib.mmd_ridge_predict.argtypes = [vp, vp, i32, i32, vp]
    lib.mmd_ridge_predict.restype = i32
    lib.mmd_cosine_distances.argtypes = [vp, vp, i32, vp, i32, i32, vp]
    lib.mmd_cosine_distances.restype = i32
    lib.mmd_dist_mat_to_pairwise_acc.argtypes = [vp, vp, i32, i32, C.POINTER(dbl)]
    lib.mmd_dist_mat_to_pairwise_acc.restype = i32
    lib.mmd_pairwise_accuracy.argtypes = [vp, vp, i32, vp, i32, i32, i32, i32, C.POINTER(dbl)]
    lib.mmd_pairwise_accuracy.restype = i32
    lib.mmd_pairwise_accuracy_static.argtypes = [i32, vp, vp, i32, vp, i32, i32, i32, i32, vp, C.c_char_p, i32]
    lib.mmd_pairwise_accuracy_static.restype = i32
    lib.mmd_ttest_1samp.argtypes = [vp, vp, i32, i32, dbl, dbl, i32, vp]
    lib.mmd_ttest_1samp.restype = i32
    lib.mmd_tfce.argtypes = [vp, vp, i32, i32, vp, i32, dbl, dbl, dbl, vp]
    lib.mmd_tfce.restype = i32
    lib.mmd_cluster_labels.argtypes = [vp, vp, i32, vp, i32, dbl, vp]
    lib.mmd_cluster_labels.restype = i32
    lib.mmd_predict_centre.argtypes = [vp, vp, i32, dbl, vp]
    lib.mmd_predict_centre.restype = i32
    lib.mmd_get_stats.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(i64), C.POINTER(i64), C.POINTER(dbl)]
    lib.mmd_get_stats.restype = i32
    lib.mmd_get_counters.argtypes = [vp, C.POINTER(i64), C.POINTER(dbl)]
    lib.mmd_get_counters.restype = i32
    lib.mmd_set_predict_mode.argtypes = [vp, i32]
    lib.mmd_set_predict_mode.restype = i32
    lib.mmd_set_syrk_mode.argtypes = [vp, i32]
    lib.mmd_set_syrk_mode.restype = i32
    lib.mmd_enable_margins.argtypes = [vp, i32]
    lib.mmd_enable_margins.restype = i32
    lib.mmd_get_margins.argtypes = [vp, vp, i32]
    lib.mmd_get_margins.restype = i32
    lib.mmd_set_group_size.argtypes = [vp, i32]
    lib.mmd_set_group_size.restype = i32
    lib.mmd_reset_stats.argtypes = [vp]
    lib.mmd_reset_stats.restype = None
    lib.mmd_configure.argtypes = [vp, i32, i32, i32]
    lib.mmd_configure.restype = i32
    _lib = lib
    return lib
