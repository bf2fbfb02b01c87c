"""ctypes binding of the C-ABI library (include/venture_b200.h).

The library is the product: if it cannot be loaded, or no CUDA device is
present, calls fail loudly -- there is no CPU or PyTorch fallback.
"""
import ctypes
import os

from .exception import VentureException

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libventure_b200.so")

VK_OK, VK_EINVAL, VK_ECUDA, VK_EUNSUPPORTED, VK_ESTATE = 0, 1, 2, 3, 4
BLOCK_ONE, BLOCK_ALL, BLOCK_EACH, BLOCK_ORDERED = -1, -2, -3, -4
FP64, FP32 = 0, 1
RESAMPLE_MULTINOMIAL, RESAMPLE_SYSTEMATIC = 0, 1

c = ctypes
_vp = c.c_void_p
_i32, _i64, _u32, _u64, _f64 = c.c_int32, c.c_int64, c.c_uint32, c.c_uint64, c.c_double

# name -> argtypes (restype is int status unless listed in _RESTYPES)
PROTOTYPES = {
    "vk_last_error": [],
    "vk_version": [],
    "vk_device_count": [_vp],
    "vk_create_normal_normal": [_i64, _u64, c.c_int, _f64, _f64, _f64, _i64, _vp],
    "vk_create_program": [_i64, _u64, c.c_int, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp],
    "vk_program_resimulate": [_vp, _i32],
    "vk_create_blr": [_i64, _u64, c.c_int, _i32, _i64, _f64, _f64, _vp],
    "vk_create_hmm": [_i64, _u64, c.c_int, _i32, _i32, _i32, _vp, _vp, _i32, _i32, _vp],
    "vk_create_lda": [_i64, _u64, c.c_int, _i32, _i32, _i32, _i32, _f64, _f64, _vp],
    "vk_destroy": [_vp],
    "vk_kind": [_vp],
    "vk_num_particles": [_vp],
    "vk_set_seed": [_vp, _u64],
    "vk_get_seed": [_vp, _vp],
    "vk_get_step": [_vp, _vp],
    "vk_set_step": [_vp, _u32],
    "vk_set_particle_base": [_vp, _i64],
    "vk_reset_to_prior": [_vp],
    "vk_set_stream": [_vp, _vp],
    "vk_synchronize": [_vp],
    "vk_nn_observe": [_vp, _vp, _i64],
    "vk_blr_observe": [_vp, _vp, _vp, _i64],
    "vk_hmm_observe": [_vp, _i32],
    "vk_lda_observe": [_vp, _vp, _vp, _i64],
    "vk_forget_observations": [_vp, _vp],
    "vk_make_consistent": [_vp, _vp],
    "vk_likelihood_weight": [_vp],
    "vk_get_log_weights": [_vp, _vp],
    "vk_set_log_weights": [_vp, _vp],
    "vk_increment_log_weights": [_vp, _vp],
    "vk_logsumexp": [_vp, _vp],
    "vk_normalized_probs": [_vp, _vp],
    "vk_logsumexp_array": [_vp, _i64, c.c_int, _vp],
    "vk_resample": [_vp, _i64, c.c_int, _vp, _vp],
    "vk_mh": [_vp, _i32, _i32, _i32, _vp, _vp],
    "vk_gibbs": [_vp, _i32, _i32, _i32, _vp, _vp],
    "vk_emap": [_vp, _i32, _i32, _i32, _vp, _vp],
    "vk_slice": [_vp, _i32, _i32, _f64, _f64, _i32, _vp, _vp],
    "vk_slice_doubling": [_vp, _i32, _i32, _f64, _i32, _i32, _vp, _vp],
    "vk_pgibbs": [_vp, _i32, _i32, _i32, _i32, _vp, _vp],
    "vk_p2p_export": [_vp, _vp, _i32, _vp],
    "vk_p2p_open": [_vp, _i32, _i32, _vp, _i32],
    "vk_p2p_resample": [_vp, _i32],
    "vk_p2p_status": [_vp, _vp, _vp],
    "vk_logsumexp_static": [_vp, _i64, _vp, _vp],
    "vk_subsampled_mh": [_vp, _i32, _i32, _i64, _i32, _f64, _i32, _f64, _i32, _i32, _vp, _vp],
    "vk_log_likelihood_at": [_vp, _i32, _i32, _vp],
    "vk_log_joint_at": [_vp, _i32, _i32, _vp],
    "vk_num_sites": [_vp, _vp],
    "vk_read_site": [_vp, _i64, _vp],
    "vk_write_site": [_vp, _i64, _vp],
    "vk_refresh": [_vp],
    "vk_set_precision": [_vp, _i32],
    "vk_get_precision": [_vp, _vp],
    "vk_num_blocks": [_vp, _i32, _vp],
    "vk_device_buffer": [_vp, _i32, _vp, _vp],
    "vk_shard_weight_max": [_vp, _vp],
    "vk_shard_weight_stats": [_vp, _vp],
    "vk_shard_build_cdf": [_vp, _f64, _i32, _vp, _vp],
    "vk_shard_ancestors": [_vp, _u64, _u64, _i64, _i64, _i64, c.c_int, _u32, _vp, _vp],
    "vk_state_row_bytes": [_vp, _vp],
    "vk_pack_state": [_vp, _vp, _i64, _vp],
    "vk_unpack_state": [_vp, _vp, _i64, _vp],
    "vk_fill_log_weights": [_vp, _f64],
    "vk_blr_set_observation_shard": [_vp, _i64, _i64],
    "vk_ssmh_begin": [_vp],
    "vk_ssmh_propose": [_vp, _i32, _i32, _i32, _f64],
    "vk_ssmh_round": [_vp, _i64, _i64, _vp],
    "vk_ssmh_decide": [_vp, _vp, _i64, _i32, _f64, _vp],
    "vk_ssmh_commit": [_vp],
    "vk_ssmh_stats": [_vp, _i32, _vp],
    "vk_sp_simulate": [_i32, _vp, _i32, _i64, _i32, _u64, _u32, c.c_int, _vp],
    "vk_sp_log_density": [_i32, _vp, _vp, _i32, _i64, _i32, c.c_int, _vp],
    "vk_sp_log_density_of_data": [_i32, _vp, _i32, _i64, _i32, c.c_int, _vp],
    "vk_sp_aaa_simulate": [_i32, _vp, _i32, _i64, _i32, _u64, _u32, c.c_int, _vp, _vp],
}
_RESTYPES = {"vk_last_error": c.c_char_p, "vk_num_particles": _i64}

_LIB = None


def load():
    """Load libventure_b200.so (built by venturecxx_b200.build / __graft_entry__.build)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(SO_PATH):
        raise VentureException(
            "evaluation",
            "libventure_b200.so is not built (%s); run `python -m venturecxx_b200.build`. "
            "There is no CPU fallback." % SO_PATH)
    lib = c.CDLL(SO_PATH)
    for name, argtypes in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, c.c_int)
    _LIB = lib
    return lib


def check(rc):
    if rc == VK_OK:
        return
    msg = load().vk_last_error().decode("utf-8", "replace")
    kind = "invalid_argument" if rc == VK_EINVAL else "evaluation"
    raise VentureException(kind, msg, status=rc)


def call(name, *args):
    check(getattr(load(), name)(*args))


def ptr(a):
    """void* of a numpy array (None -> NULL)."""
    if a is None:
        return None
    return a.ctypes.data_as(c.c_void_p)
