"""ctypes loader for libhydia_b200.so (C ABI in include/hydia_b200.h). No fallback: a missing or
unloadable library is an ImportError-grade failure, and creating a context without a CUDA device
raises (the product has no CPU path)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HYDIA_B200_LIB") or os.path.join(_HERE, "libhydia_b200.so")   # the override is for diagnostic builds only

U64P = C.POINTER(C.c_uint64)
F64P = C.POINTER(C.c_double)
I32P = C.POINTER(C.c_int)
VP = C.c_void_p

# name -> (restype, argtypes); must list every symbol include/hydia_b200.h declares
SIGNATURES = {
    "hx_last_error": (C.c_char_p, []),
    "hx_device_count": (C.c_int, []),
    "hx_ctx_create": (VP, [C.c_int, C.c_int, C.c_int, C.c_int, U64P, U64P, U64P, U64P, C.c_int, C.c_int]),
    "hx_ctx_destroy": (None, [VP]),
    "hx_ctx_sync": (C.c_int, [VP]),
    "hx_ctx_launches": (C.c_uint64, [VP]),
    "hx_kernel_stamp": (C.c_char_p, []),
    "hx_diag_mac_partial": (VP, [VP, VP, C.c_int, C.c_int, I32P, C.c_int]),
    "hx_ct_fold_mod": (C.c_int, [VP]),
    "hx_ct_alloc": (VP, [VP, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "hx_ct_device_view": (C.c_int, [VP, C.POINTER(C.c_void_p), U64P, I32P]),
    "hx_stream_order": (C.c_int, [VP, VP, C.c_int]),
    "hx_ctx_set_option": (C.c_int, [VP, C.c_char_p, C.c_int]),
    "hx_prof_kinds": (C.c_int, []),
    "hx_prof_name": (C.c_char_p, [C.c_int]),
    "hx_prof_enable": (C.c_int, [VP, C.c_int]),
    "hx_prof_reset": (C.c_int, [VP]),
    "hx_prof_read": (C.c_int, [VP, C.c_int, U64P, F64P, F64P]),
    "hx_timer_start": (C.c_int, [VP]),
    "hx_timer_stop": (C.c_int, [VP, F64P]),
    "hx_galois_for_rotation": (C.c_uint32, [VP, C.c_int]),
    "hx_key_load_mult": (C.c_int, [VP, U64P, U64P]),
    "hx_key_load_auto": (C.c_int, [VP, C.c_uint32, U64P, U64P]),
    "hx_ct_upload": (VP, [VP, C.c_int, C.c_int, U64P, C.c_int, C.c_int]),
    "hx_ct_download": (C.c_int, [VP, U64P]),
    "hx_ct_upload_dev": (VP, [VP, C.c_int, C.c_int, VP, C.c_int, C.c_int]),
    "hx_ct_download_dev": (C.c_int, [VP, VP]),
    "hx_ct_download_dev_async": (C.c_int, [VP, VP, VP]),
    "hx_ct_upload_dev_async": (VP, [VP, C.c_int, C.c_int, VP, C.c_int, C.c_int, VP]),
    "hx_ct_info": (C.c_int, [VP, I32P, I32P, I32P, I32P]),
    "hx_ct_clone": (VP, [VP]),
    "hx_ct_download_many": (C.c_int, [C.POINTER(VP), C.c_int, VP]),
    "hx_trace_start": (C.c_int, [VP]),
    "hx_trace_mark": (C.c_int, [VP, C.c_char_p, VP]),
    "hx_trace_dump": (C.c_int, [VP, C.c_char_p, C.c_int]),
    "hx_rescale_noalloc": (C.c_int, [VP, VP, C.c_int, C.c_int, VP]),
    "hx_ct_free": (None, [VP]),
    "hx_add": (VP, [VP, VP]),
    "hx_sub": (VP, [VP, VP]),
    "hx_add_const": (VP, [VP, C.c_double]),
    "hx_mult_const": (VP, [VP, C.c_double]),
    "hx_mult_norelin": (VP, [VP, VP]),
    "hx_relin": (VP, [VP]),
    "hx_mult": (VP, [VP, VP]),
    "hx_square": (VP, [VP]),
    "hx_rescale": (VP, [VP]),
    "hx_level_reduce": (VP, [VP, C.c_int]),
    "hx_rotate": (VP, [VP, C.c_int]),
    "hx_eval_sum": (VP, [VP, C.c_int]),
    "hx_add_many": (VP, [C.POINTER(VP), C.c_int]),
    "hx_rot_precompute": (VP, [VP]),
    "hx_rot_fast": (VP, [VP, C.c_int, VP]),
    "hx_digits_free": (None, [VP]),
    "hx_cheby_series": (VP, [VP, F64P, C.c_int, C.c_double, C.c_double]),
    "hx_poly": (VP, [VP, F64P, C.c_int]),
    "hx_compare": (VP, [VP, C.c_double, C.c_int]),
    "hx_db_register": (C.c_int, [VP, C.c_int, C.c_int, U64P]),
    "hx_db_blocks": (C.c_int, [VP]),
    "hx_score_diag": (C.c_int, [VP, VP, C.c_int, C.POINTER(VP), C.c_int]),
    "hx_membership": (VP, [VP, VP, C.c_int, C.c_double, C.c_int]),
    "hx_index": (C.c_int, [VP, VP, C.c_int, C.c_double, C.c_int, C.POINTER(VP), C.c_int]),
    "hx_db_register_ct": (C.c_int, [VP, C.c_int, C.c_int, VP]),
    "hx_db_get": (VP, [VP, C.c_int, C.c_int]),
    "hx_rot_fast_all": (VP, [VP, C.c_int]),
    "hx_rotset_get": (VP, [VP, C.c_int]),
    "hx_rotset_free": (None, [VP]),
    "hx_score_blocks": (C.c_int, [VP, I32P, C.c_int, C.POINTER(VP)]),
    "hx_ct_stack": (VP, [C.POINTER(VP), C.c_int]),
    "hx_ct_unstack": (C.c_int, [VP, C.POINTER(VP), C.c_int]),
    "hx_mult_plain": (VP, [VP, U64P, C.c_int]),
    "hx_inner_product": (VP, [VP, VP, C.c_int]),
    "hx_client_encode": (C.c_int, [VP, F64P, C.c_int, U64P]),
    "hx_score_hers": (C.c_int, [VP, C.POINTER(VP), C.c_int, C.POINTER(VP), C.c_int]),
    "hx_score_hers_blocks": (C.c_int, [VP, C.POINTER(VP), C.c_int, I32P, C.c_int, C.POINTER(VP)]),
    "hx_compare_many": (C.c_int, [VP, C.POINTER(VP), C.c_int, C.c_double, C.c_int, C.POINTER(VP)]),
    "hx_ntt": (C.c_int, [VP, U64P, C.c_int, I32P, C.c_int, C.c_int]),
    "hx_params_generate": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, I32P, I32P, I32P, I32P, U64P, U64P, U64P, U64P]),
    "hx_client_create": (VP, [VP, C.c_uint64]),
    "hx_client_destroy": (None, [VP]),
    "hx_client_gen_mult_key": (C.c_int, [VP]),
    "hx_client_gen_rot_keys": (C.c_int, [VP, I32P, C.c_int]),
    "hx_client_gen_sum_keys": (C.c_int, [VP]),
    "hx_client_encrypt": (VP, [VP, F64P, C.c_int]),
    "hx_client_enroll": (C.c_int, [VP, C.c_int, C.c_int, F64P, C.c_int]),
    "hx_client_decrypt": (C.c_int, [VP, VP, F64P]),
    "hx_client_secret": (C.c_int, [VP, C.POINTER(C.c_int8)]),
    "hx_ctx_get_params": (C.c_int, [VP, I32P, I32P, I32P, I32P, I32P, U64P, U64P, U64P, U64P]),
    "hx_key_count": (C.c_int, [VP, I32P]),
    "hx_key_list": (C.c_int, [VP, C.POINTER(C.c_uint32), C.c_int]),
    "hx_key_export": (C.c_int, [VP, C.c_uint32, U64P, U64P]),
    "hx_client_export_public": (C.c_int, [VP, U64P, U64P]),
    "hx_client_import": (VP, [VP, C.c_uint64, C.POINTER(C.c_int8), U64P, U64P]),
}

_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C image_matching_b200/csrc). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)      # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class HxError(RuntimeError):
    """Raised where OpenFHE would throw OpenFHEException (missing key, level exhausted, ...)."""


def check_ptr(p):
    if not p:
        raise HxError(load().hx_last_error().decode())
    return p


def check_int(rc):
    if rc < 0:
        raise HxError(load().hx_last_error().decode())
    return rc
