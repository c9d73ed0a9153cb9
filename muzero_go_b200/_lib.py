"""ctypes binding of libmzgo.so (include/mzgo.h).  No CPU fallback: if the library is missing it is
built with nvcc; if that fails, or a compute call is made without a B200-class device, we raise."""
import ctypes as C
import os

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))
BOARD_WORDS = 64
MAX_N = 19


class MzgoError(RuntimeError):
    pass


vp = C.c_void_p


class TreeDesc(C.Structure):
    """mzgo_tree_t"""
    _fields_ = [("G", C.c_int32), ("A", C.c_int32), ("M", C.c_int32), ("move_index", C.c_int32),
                ("root_prior", vp), ("root_mask", vp), ("prior", vp), ("edge_visit", vp), ("edge_vsum", vp),
                ("child", vp), ("parent", vp), ("root_visit", vp), ("root_vsum", vp), ("n_nodes", vp),
                ("leaf_parent", vp), ("leaf_action", vp), ("leaf_node", vp), ("leaf_depth", vp),
                ("leaf_in_slot", vp), ("leaf_out_slot", vp), ("choice_ctr", vp), ("depth_sum", vp),
                ("sims_done", vp), ("dropped", vp), ("active", vp), ("move_index_dev", vp)]


class SearchParams(C.Structure):
    """mzgo_search_params_t"""
    _fields_ = [("c_puct", C.c_double), ("discount", C.c_double), ("dirichlet_epsilon", C.c_double),
                ("pass_epsilon", C.c_double), ("dirichlet_alpha", C.c_float), ("c_puct_is_int", C.c_int32),
                ("seed", C.c_uint64), ("variant", C.c_int32), ("game_base", C.c_int32)]


NET_PARAM_FIELDS = ["rep_w1", "rep_b1", "rep_w2", "rep_b2", "rep_w3", "rep_b3", "emb", "dyn_w", "dyn_b",
                    "rew_w", "rew_b", "rh_w", "rh_b", "ro_w", "ro_b", "val_w", "val_b", "vfc_w", "vfc_b",
                    "pol_w", "pol_b", "pass_logit"]


class NetDesc(C.Structure):
    """mzgo_net_t"""
    _fields_ = ([("n", C.c_int32), ("C", C.c_int32), ("A", C.c_int32), ("precision", C.c_int32),
                 ("max_batch", C.c_int32), ("reserved", C.c_int32)]
                + [(f, vp) for f in NET_PARAM_FIELDS]
                + [("workspace", vp), ("workspace_bytes", C.c_int64)])


class SelfPlayDesc(C.Structure):
    """mzgo_selfplay_t"""
    _fields_ = [("tree", TreeDesc), ("params", SearchParams), ("net", NetDesc),
                ("n", C.c_int32), ("num_simulations", C.c_int32), ("max_game_moves", C.c_int32),
                ("temperature_moves", C.c_int32), ("temperature", C.c_float), ("komi", C.c_float),
                ("restart_finished", C.c_int32), ("traj_capacity", C.c_int32),
                ("boards", vp), ("arena", vp), ("root_policy", vp), ("root_value_net", vp), ("leaf_policy", vp),
                ("leaf_reward", vp), ("leaf_value", vp), ("in_slot", vp), ("out_slot", vp), ("action", vp),
                ("move_count", vp), ("pass_streak", vp), ("game_temperature", vp), ("score_winner", vp),
                ("traj_boards", vp), ("traj_action", vp), ("traj_policy", vp), ("traj_root_value", vp),
                ("traj_reward", vp), ("traj_done", vp), ("traj_winner", vp), ("counters", vp)]


i32, u32, f32, i64 = C.c_int, C.c_uint32, C.c_float, C.c_int64
_PROTOS = {
    "mzgo_last_error": (C.c_char_p, []),
    "mzgo_abi_version": (i32, []),
    "mzgo_device_ok": (i32, []),
    "mzgo_launch_count": (i64, [i32]),
    "mzgo_graph_replays": (C.c_longlong, []),
    "mzgo_debug_fastdiv_check": (C.c_longlong, [C.c_ulonglong, i32, C.c_longlong]),
    "mzgo_debug_rules_tpb_min": (i32, [i32]),
    "mzgo_debug_tc_trace": (i32, [vp]),
    "mzgo_debug_tree_trace": (i32, [vp]),
    "mzgo_sizeof": (i64, [i32]),
    "mzgo_rules_reset": (i32, [vp, i32, i32, vp]),
    "mzgo_rules_step": (i32, [vp, vp, vp, i32, i32, vp]),
    "mzgo_rules_score": (i32, [vp, vp, vp, f32, i32, i32, vp]),
    "mzgo_rules_expand_planes": (i32, [vp, vp, i32, i32, i32, vp]),
    "mzgo_rules_pack_planes": (i32, [vp, vp, i32, i32, vp]),
    "mzgo_rules_playout": (i32, [vp, vp, i32, vp, vp, vp, i32, i32, vp]),
    "mzgo_tree_reset_roots": (i32, [C.POINTER(TreeDesc), C.POINTER(SearchParams), vp, vp, i32, vp, u32, vp]),
    "mzgo_tree_select": (i32, [C.POINTER(TreeDesc), C.POINTER(SearchParams), vp, i32, vp]),
    "mzgo_tree_expand_backup": (i32, [C.POINTER(TreeDesc), C.POINTER(SearchParams), vp, vp, vp, vp]),
    "mzgo_tree_step": (i32, [C.POINTER(TreeDesc), C.POINTER(SearchParams), vp, vp, vp, vp, i32, i32, i32, vp]),
    "mzgo_tree_visit_counts": (i32, [C.POINTER(TreeDesc), vp, vp]),
    "mzgo_tree_choose_action": (i32, [C.POINTER(TreeDesc), C.POINTER(SearchParams), f32, vp, u32, vp, vp, vp, vp]),
    "mzgo_net_workspace_bytes": (i64, [i32, i32, i32, i32]),
    "mzgo_net_slot_bytes": (i64, [i32, i32, i32]),
    "mzgo_net_prepare": (i32, [C.POINTER(NetDesc), vp]),
    "mzgo_net_initial": (i32, [C.POINTER(NetDesc), vp, vp, vp, vp, vp, vp, i32, i32, vp]),
    "mzgo_net_recurrent": (i32, [C.POINTER(NetDesc), vp, vp, vp, vp, vp, vp, vp, i32, i32, vp]),
    "mzgo_net_recurrent_tree": (i32, [C.POINTER(NetDesc), vp, C.POINTER(TreeDesc), C.POINTER(SearchParams), vp, i32, vp, vp,
                                      vp, i32, vp]),
    "mzgo_net_predict": (i32, [C.POINTER(NetDesc), vp, vp, vp, vp, i32, i32, vp]),
    "mzgo_net_export_latent": (i32, [C.POINTER(NetDesc), vp, vp, vp, i32, vp]),
    "mzgo_net_import_latent": (i32, [C.POINTER(NetDesc), vp, vp, vp, i32, vp]),
    "mzgo_selfplay_run": (i32, [C.POINTER(SelfPlayDesc), i32, i32, u32, vp]),
    "mzgo_search_run": (i32, [C.POINTER(SelfPlayDesc), vp, i32, vp]),
    "mzgo_train_conv3x3_prep_w": (i32, [vp, vp, vp, i32, i32, vp]),
    "mzgo_train_conv3x3_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]),
    "mzgo_train_conv3x3_bwd_data": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, vp]),
    "mzgo_train_wgrad_scratch_floats": (i64, [i32, i32, i32]),
    "mzgo_train_conv3x3_bwd_weight": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]),
    "mzgo_train_wgrad_static_max_floats": (i64, []),
    "mzgo_train_conv3x3_bwd_weight_static": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]),
    "mzgo_selfplay_pack_rows": (i32, [C.POINTER(SelfPlayDesc), i32, i32, vp, i32, i32, vp]),
}

_lib = None


def so_path():
    return os.path.join(_HERE, "libmzgo.so")


def load():
    """Load (building if needed) libmzgo.so.  Raises MzgoError when it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    path = so_path()
    # MZGO_KEEP_SO=1: use the library as built (the checked build of `python -m muzero_go_b200.build --checked`)
    stale = os.path.isdir(os.path.join(_HERE, "csrc")) and not _build.up_to_date() and not os.environ.get("MZGO_KEEP_SO")
    if not os.path.exists(path) or stale:
        try:
            _build.build()
        except Exception as e:  # no nvcc on the box: use the shipped .so if there is one
            if not os.path.exists(path):
                raise MzgoError(f"libmzgo.so is missing and could not be built: {e}") from e
    lib = C.CDLL(path)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    for which, struct in enumerate((TreeDesc, SearchParams, NetDesc, SelfPlayDesc)):
        if lib.mzgo_sizeof(which) != C.sizeof(struct):
            raise MzgoError(f"{struct.__name__}: ctypes mirror is {C.sizeof(struct)} B, library says "
                            f"{lib.mzgo_sizeof(which)} B (include/mzgo.h and _lib.py disagree)")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise MzgoError(f"mzgo error {rc}: {load().mzgo_last_error().decode()}")


def require_device():
    if not load().mzgo_device_ok():
        raise MzgoError("no sm_100 CUDA device: muzero_go_b200 has no CPU fallback")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def norm_device(device):
    """torch.device with an explicit index ('cuda' -> 'cuda:<current>'), so device comparisons are exact."""
    import torch
    d = torch.device(device)
    if d.type == "cuda" and d.index is None:
        d = torch.device("cuda", torch.cuda.current_device())
    return d
