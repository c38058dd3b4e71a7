"""ctypes loader of libthetal_b200.so (the product).  No CPU fallback: a missing library or a
missing GPU is a hard error."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# THETAL_B200_LIB: alternative build of the same library (kernel tuning variants, tools/bench_dirk.py)
LIB_PATH = os.environ.get("THETAL_B200_LIB") or os.path.join(_HERE, "csrc", "libthetal_b200.so")
_lib = None

# every symbol include/thetal_b200.h declares
SYMBOLS = [
    "initialize_hommexx_session", "finalize_hommexx_session", "reset_cxx_comm",
    "init_connectivity", "add_connection", "finalize_connectivity",
    "init_simulation_params_c", "init_elements_c", "init_hvcoord_c", "init_time_level_c",
    "init_reference_element_c", "init_elements_2d_c", "init_tensorvisc_c", "init_tensorvisc2_c",
    "init_geopotential_c", "init_diagnostics_c", "init_elements_states_c", "init_reference_states_c",
    "init_functors_c", "init_boundary_exchanges_c", "initialize_dp3d_from_ps_c",
    "push_forcing_to_c", "prim_run_subcycle_c", "cxx_push_results_to_f90", "push_test_state_to_c",
    "sync_diagnostics_to_host_c",
    "thetal_b200_set_nlev", "thetal_b200_init_comm", "thetal_b200_get_nccl_id",
    "thetal_b200_set_abort_on_error", "thetal_b200_last_error", "thetal_b200_last_error_msg", "thetal_b200_clear_error",
    "thetal_b200_set_extra_params", "prim_advance_exp_c", "caar_c", "dirk_c", "advance_hypervis_c",
    "thetal_b200_get_derived", "thetal_b200_zero_derived", "thetal_b200_sync",
    "thetal_b200_time_steps", "thetal_b200_launch_count", "thetal_b200_profile_step",
    "thetal_b200_push_level", "thetal_b200_pull_level", "thetal_b200_pin_host", "thetal_b200_enqueue_step",
    "thetal_b200_halo_counts", "thetal_b200_halo_lists", "thetal_b200_dss_table", "thetal_b200_halo_stats", "thetal_b200_global_diagnostics", "vertical_remap_c", "thetal_b200_pipeline_step", "thetal_b200_pipeline_flush", "thetal_b200_set_vertical_remap",
    "thetal_b200_time_dirk", "thetal_b200_measure_fp64_peak", "apply_cam_forcing_dynamics_c",
    "init_elements_states_strided_c", "dss_c", "thetal_b200_field_extrema",
    "thetal_b200_global_diagnostics_noalloc",
]


def _preload_bundled_nccl():
    """The library links against libnccl.so.2; PyTorch ships its own copy under the same soname and needs the symbols of
    that version.  Whichever copy a process loads first serves both, so when a PyTorch-bundled NCCL is installed it is
    loaded (globally) before the library: `import torch` keeps working whatever the import order.  A Fortran host
    without Python is not concerned: there the system NCCL is the only one."""
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            path = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(path):
                ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
                return path
    except Exception:
        pass
    return None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        _preload_bundled_nccl()
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.thetal_b200_last_error.restype = ctypes.c_int
        _lib.thetal_b200_last_error_msg.restype = ctypes.c_char_p
        _lib.thetal_b200_time_steps.restype = ctypes.c_double
        _lib.thetal_b200_time_dirk.restype = ctypes.c_double
        _lib.thetal_b200_measure_fp64_peak.restype = ctypes.c_double
        _lib.thetal_b200_launch_count.restype = ctypes.c_long
    return _lib
