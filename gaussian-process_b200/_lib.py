"""ctypes binding of libgpk.so (C ABI declared in include/gpk.h).  Fails loudly when the library
has not been built: there is no fallback path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GPK_LIB") or os.path.join(_HERE, "libgpk.so")     # GPK_LIB: instrumented builds (tools/)

_vp, _i, _ll, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t

# name -> (restype, argtypes); mirrors include/gpk.h one to one
SIGNATURES = {
    "gpk_version": (_i, []),
    "gpk_last_error_string": (C.c_char_p, []),
    "gpk_launch_count": (C.c_ulonglong, []),
    "gpk_padded_n": (_i, [_i]),
    "gpk_cov_batched": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _ll, _ll, _vp]),
    "gpk_potrf_workspace_bytes": (_sz, [_i, _i]),
    "gpk_potrf_batched": (_i, [_vp, _i, _i, _vp, _vp, _sz, _vp]),
    "gpk_staged_workspace_bytes": (_sz, [_i, _i]),
    "gpk_fit_alpha_batched": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _sz, _vp]),
    "gpk_selfpredict_counts": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gpk_loss_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "gpk_loss_batched": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, C.c_double, _vp, _vp, _vp, _vp, _vp,
                              _vp, _vp, _sz, _vp]),
    "gpk_fit_workspace_bytes": (_sz, [_i, _i, _i]),
    "gpk_fit": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gpk_predict_workspace_bytes": (_sz, [_i, _i]),
    "gpk_predict": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _ll, _vp, _vp, _i, _vp, _sz, _vp]),
    "gpk_predict_static": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _ll, _vp, _vp, _vp]),
    "gpk_predict_state_bytes": (_sz, [_i]),
    "gpk_predict_prepare": (_i, [_vp, _vp, _i, _vp, _sz, _vp]),
    "gpk_predict_emu_workspace_bytes": (_sz, [_i, _i]),
    "gpk_predict_emu": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _ll, _vp, _vp, _i, _vp, _sz, _vp]),
    "gpk_kmeans_step": (_i, [_vp, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "gpk_pso_layout": (_sz, [_i, _i, _i, _i, _i, _i, _ll, _vp]),
    "gpk_pso_step": (_i, [_vp, _sz, _i, _i, _i, _i, _i, _i, _ll, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gpk_prob_surface": (_i, [_vp, _ll, _i, _i, _vp, _vp, _vp, _vp]),
    "gpk_surface_scan": (_i, [_vp, _ll, _vp, _vp, _vp]),
    "gpk_surface_gather": (_i, [_vp, _vp, _ll, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gpk_set_groups": (_i, [_i, _i]),
    "gpk_oz_slice_bytes": (_sz, [_i, _i]),
    "gpk_oz_slice": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp]),
    "gpk_oz_gemm_tn": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _vp]),
    "gpk_csv_bound": (_sz, [_ll, _i]),
    "gpk_csv_format_f64": (_ll, [_vp, _ll, _i, _ll, _vp, _sz, _i]),
    "gpk_csv_surface_bound": (_sz, [_ll, _i, _i]),
    "gpk_csv_format_surface": (_ll, [_vp, _vp, _vp, _ll, _i, _i, _vp, _sz, _i]),
    "gpk_finish_wasserstein": (_i, [_vp, _ll, _i, _vp, _i, _vp, _vp, _vp, _vp]),
    "gpk_debug_trap": (_i, [_vp]),
    "gpk_timing_begin": (_i, [_vp]),
    "gpk_timing_end": (_i, []),
    "gpk_timing_query": (_i, [_i, C.c_char_p, _i, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]),
    "gpk_sum_gaussians": (_i, [_vp, _i, _vp, _i, C.c_double, _vp, _vp]),
}

_lib = None


class GpkError(RuntimeError):
    pass


def load():
    """Loads libgpk.so (once) and declares every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or gaussian-process_b200/csrc/build.sh).  This package has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = library / header mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, what):
    if status != 0:
        msg = load().gpk_last_error_string()
        raise GpkError(f"{what} failed (status {status}): {msg.decode() if msg else ''}")


def launch_count():
    return int(load().gpk_launch_count())


_timing = False


def timing_active():
    return _timing


def timing_begin(stream):
    global _timing
    check(load().gpk_timing_begin(stream), "gpk_timing_begin")
    _timing = True


def timing_end():
    """-> {kernel class: (total ms, launches)} since timing_begin."""
    global _timing
    lib = load()
    _timing = False
    check(lib.gpk_timing_end(), "gpk_timing_end")
    out, i = {}, 0
    name = C.create_string_buffer(128)
    ms, cnt = C.c_double(), C.c_longlong()
    while lib.gpk_timing_query(i, name, 128, C.byref(ms), C.byref(cnt)) == 0:
        if cnt.value:
            out[name.value.decode()] = (ms.value, cnt.value)
        i += 1
    return out
