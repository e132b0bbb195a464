"""ctypes binding of the C ABI (include/asset_b200.h).

The CUDA library is mandatory: there is no CPU fallback on this path.  Loading fails loudly if the
in-tree ``libasset_b200.so`` has not been built (``python -c "import __graft_entry__ as g; g.build()"``).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libasset_b200.so")


class AssetB200Error(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise AssetB200Error(
            f"{LIB_PATH} is missing: build it with `make -C asset_asrl_b200/csrc` "
            "(or __graft_entry__.build()); the KKT path has no CPU fallback")
    return C.CDLL(LIB_PATH)


lib = _load()

_i, _d, _ll, _vp, _cp = C.c_int, C.c_double, C.c_longlong, C.c_void_p, C.c_char_p
_ip = C.POINTER(C.c_int)
_dp = C.POINTER(C.c_double)
_llp = C.POINTER(C.c_longlong)

# name -> (restype, argtypes); mirrors include/asset_b200.h one to one
SIGNATURES = {
    "ab_last_error": (_cp, []),
    "ab_version": (_i, []),
    "ab_device_count": (_i, []),
    "ab_graph_create": (_vp, []),
    "ab_graph_destroy": (None, [_vp]),
    "ab_graph_node": (_i, [_vp, _i, _i, _i, _ip, _i, _dp, _i, _ip, _i, _i, _i]),
    "ab_graph_node_info": (_i, [_vp, _i, _ip, _ip, _ip, _ip]),
    "ab_graph_describe": (_ll, [_vp, _i, _cp, _ll]),
    "ab_problem_create": (_vp, [_i, _i, _i]),
    "ab_problem_destroy": (None, [_vp]),
    "ab_problem_set_dims": (_i, [_vp, _i, _i, _i]),
    "ab_problem_add_tape": (_i, [_vp, _vp, _i, _i]),
    "ab_problem_add_function": (_i, [_vp, _i, _i, _i, _i, _i, _ip, _ip, _dp]),
    "ab_problem_analyze": (_i, [_vp]),
    "ab_problem_dims": (_i, [_vp, _llp]),
    "ab_problem_pattern": (_i, [_vp, _ip, _ip]),
    "ab_problem_kkt_locations": (_i, [_vp, _ip]),
    "ab_problem_function_info": (_i, [_vp, _i, _ip]),
    "ab_problem_function_indices": (_i, [_vp, _i, _ip, _ip]),
    "ab_problem_function_appconst": (_i, [_vp, _i, _dp]),
    "ab_problem_tape_info": (_i, [_vp, _i, _ip]),
    "ab_problem_tape_describe": (_ll, [_vp, _i, _cp, _ll]),
    "ab_problem_program_stats": (_i, [_vp, _i, _i, _llp]),
    "ab_problem_program_code": (_i, [_vp, _i, _i, _vp, _dp]),
    "ab_problem_program_tables": (_ll, [_vp, _i, _i, _dp, _ll]),
    "ab_problem_kernel_ptx": (_ll, [_vp, _i, _i, _i, _cp, _ll]),
    "ab_ptx_assemble": (_i, [_cp, _llp]),
    "ab_problem_tape_code": (_i, [_vp, _i, _vp]),
    "ab_phase_create": (_vp, [_vp, _i, _i, _i, _i, _i, _i]),
    "ab_phase_destroy": (None, [_vp]),
    "ab_phase_set_control_mode": (_i, [_vp, _i]),
    "ab_phase_set_integral_mode": (_i, [_vp, _i]),
    "ab_phase_set_static_params": (_i, [_vp, _i]),
    "ab_phase_set_mesh": (_i, [_vp, _i, _dp, _ip]),
    "ab_phase_add_function": (_i, [_vp, _i, _i, _i, _ip, _i, _ip, _i, _ip, _i, _ip, _i]),
    "ab_phase_clear_functions": (_i, [_vp]),
    "ab_phase_dims": (_i, [_vp, _ip]),
    "ab_phase_state_spacing": (_i, [_vp, _dp]),
    "ab_phase_transcribe": (_i, [_vp, _vp, _i, _i, _i, _ip]),
    "ab_phase_region_vindex": (_i, [_vp, _i, _ip, _i, _ip, _i, _ip, _i, _ip, _i, _ip]),
    "ab_phase_make_input": (_i, [_vp, _dp, _dp, _dp]),
    "ab_problem_fused_info": (_ll, [_vp, _i, _llp, _cp, _ll]),
    "ab_problem_fused_ptx": (_ll, [_vp, _i, _i, _cp, _ll]),
    "ab_problem_fused_residual_ptx": (_ll, [_vp, _i, _i, _cp, _ll]),
    "ab_problem_fused_mode_ptx": (_ll, [_vp, _i, _i, _i, _cp, _ll]),
    "ab_problem_fused_layout": (_i, [_vp, _i, _llp, _ip, _ip]),
    "ab_problem_dest_maps": (_i, [_vp, _i, _ip, _ip, _llp]),
    "ab_plan_create": (_i, [_vp, _i, C.POINTER(_vp)]),
    "ab_plan_destroy": (None, [_vp]),
    "ab_plan_create_sharded": (_i, [_vp, _i, _i, _i, C.POINTER(_vp)]),
    "ab_plan_cut_count": (_ll, [_vp]),
    "ab_plan_fused_segments": (_i, [_vp]),
    "ab_plan_output_ranges": (_i, [_vp, _i, _llp, _i]),
    "ab_plan_get_outputs_packed": (_i, [_vp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "ab_plan_cut_buffer": (_vp, [_vp]),
    "ab_plan_cut_pack": (_i, [_vp]),
    "ab_plan_cut_unpack": (_i, [_vp]),
    "ab_partition_create": (_vp, [_vp, _i, _i]),
    "ab_partition_destroy": (None, [_vp]),
    "ab_partition_app_range": (_i, [_vp, _i, _ip]),
    "ab_partition_gather_sizes": (_i, [_vp, _i, _i, _llp]),
    "ab_partition_gather": (_i, [_vp, _i, _i, _llp, _ip, _ip]),
    "ab_partition_cut_size": (_ll, [_vp, _i]),
    "ab_partition_cut": (_i, [_vp, _i, _ip, _ip, _ip]),
    "ab_problem_gather_sizes": (_i, [_vp, _i, _i, _llp]),
    "ab_problem_gather": (_i, [_vp, _i, _i, _llp, _ip, _ip]),
    "ab_plan_set_inputs": (_i, [_vp, _dp, _dp, _dp]),
    "ab_plan_set_inputs_sliced": (_i, [_vp, _dp, _dp, _dp]),
    "ab_plan_input_ranges": (_ll, [_vp, _i, _llp, _ll]),
    "ab_problem_input_ranges": (_ll, [_vp, _i, _llp, _ll]),
    "ab_plan_set_solver_coeffs": (_i, [_vp, _dp]),
    "ab_plan_eval": (_i, [_vp, _i, _d]),
    "ab_plan_sync": (_i, [_vp]),
    "ab_plan_get_outputs": (_i, [_vp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "ab_plan_eval_host": (_i, [_vp, _i, _d, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "ab_plan_profile": (_i, [_vp, _i]),
    "ab_plan_profile_read": (_i, [_vp, C.POINTER(C.c_float), _i]),
    "ab_plan_profile_label": (_ll, [_vp, _i, _cp, _ll]),
    "ab_plan_device_ptr": (_vp, [_vp, _i]),
    "ab_plan_check_guards": (_i, [_vp]),
    "ab_problem_window": (_i, [_vp, _ll, _i, _ll, _ll, _ll, _i]),
    "ab_problem_window_info": (_i, [_vp, _i, _ip]),
    "ab_problem_window_far_vars": (_i, [_vp, _ip, _i]),
    "ab_plan_stream": (_vp, [_vp]),
    "ab_plan_launches": (_ll, [_vp]),
    "ab_plan_specialised_launches": (_ll, [_vp]),
    "ab_plan_kernel_info": (_i, [_vp, _i, _i, _ip]),
}
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


def last_error():
    return lib.ab_last_error().decode()


def check(rc):
    """Raise on a negative / non-zero status; returns rc for id-returning calls."""
    if rc < 0:
        msg = last_error()
        if rc == -3:
            raise AssetB200Error(msg)
        raise ValueError(msg)
    return rc


def check0(rc):
    if rc != 0:
        raise AssetB200Error(f"status {rc}: {last_error()}")


def iarr(a):
    a = np.ascontiguousarray(a, dtype=np.int32).ravel()
    return a, a.ctypes.data_as(_ip)


def darr(a):
    a = np.ascontiguousarray(a, dtype=np.float64).ravel()
    return a, a.ctypes.data_as(_dp)


def dptr(a):
    """Pointer to an existing contiguous float64 array (no copy) or NULL."""
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)
