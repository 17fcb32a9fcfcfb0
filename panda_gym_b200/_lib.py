"""ctypes binding of libpanda_b200.so (C ABI in include/panda_b200.h).

The shared library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no CPU or
PyTorch fallback: importing this module without the library raises.
"""
import ctypes
import os

c_int, c_double, c_void_p, c_u64, c_i64, c_float, c_u8 = (ctypes.c_int, ctypes.c_double, ctypes.c_void_p,
                                                          ctypes.c_uint64, ctypes.c_int64, ctypes.c_float, ctypes.c_uint8)
PG_MAX_OBSTACLES = 8
LIB_NAME = "libpanda_b200.so"
LIB_PATH = os.environ.get("PANDA_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)  # override: tuning variants


class Stage(ctypes.Structure):
    """pg_stage: scenario / curriculum-stage parameter block."""
    _fields_ = [
        ("goal_sampler", c_int), ("goal_lo", c_double * 3), ("goal_hi", c_double * 3),
        ("goal_rmin", c_double), ("goal_rmax", c_double),
        ("goal_upper_half", c_int), ("goal_front_half", c_int), ("goal_three_quarter_front", c_int),
        ("n_obstacles", c_int), ("obstacle_type", c_int * PG_MAX_OBSTACLES),
        ("obstacle_size", (c_double * 3) * PG_MAX_OBSTACLES),
        ("obstacle_sampler", c_int), ("randomize_obstacle_position", c_int), ("allow_overlapping_obstacles", c_int),
        ("random_num_obs", c_int), ("sample_size_lo", c_int), ("sample_size_hi", c_int),
        ("max_episode_steps", c_int), ("ee_error_threshold", c_double), ("speed_threshold", c_double),
        ("pose_randomizer", c_int), ("force_randomize_pose", c_int), ("pose_lo", c_double * 3), ("pose_hi", c_double * 3),
        ("random_size_cuboids", c_int), ("custom_neutral", c_int), ("neutral_q", c_double * 7), ("needs_scene", c_int),
    ]


class Config(ctypes.Structure):
    """pg_config: environment configuration."""
    _fields_ = [
        ("n_envs", c_int), ("task", c_int), ("control", c_int), ("reward_type", c_int), ("goal_condition", c_int),
        ("action_limiter", c_int), ("obs_js", c_int), ("obstacle_obs", c_int), ("n_substeps", c_int),
        ("truncate_on_collision", c_int), ("terminate_on_success", c_int),
        ("collision_reward", c_double), ("safety_distance", c_double), ("base_position", c_double * 3),
        ("auto_reset", c_int), ("randomize_robot_pose", c_int), ("stage", Stage),
        ("has_fixed_target", c_int), ("fixed_target", c_double * 3), ("pose_retry_limit", c_int),
    ]


class View(ctypes.Structure):
    _fields_ = [("ptr", c_void_p), ("dtype", c_int), ("ndim", c_int), ("shape", c_i64 * 3)]


class HostOut(ctypes.Structure):
    _fields_ = [("observation", c_void_p), ("achieved_goal", c_void_p), ("desired_goal", c_void_p),
                ("reward", c_void_p), ("done", c_void_p), ("flags", c_void_p)]


class HerBatch(ctypes.Structure):
    _fields_ = [(k, c_void_p) for k in ("observation", "achieved_goal", "desired_goal", "next_observation", "next_achieved_goal",
                                        "next_desired_goal", "action", "reward", "done")] + [("batch", c_i64), ("n_virtual", c_i64), ("row_stride", c_i64)]


VIEW_IDS = {name: i for i, name in enumerate([
    "observation", "achieved_goal", "desired_goal", "reward", "done", "flags", "terminal_observation",
    "terminal_achieved", "ep_return", "ep_length", "q", "qd", "goal", "obstacle_pos", "stats", "step_count", "d_link", "object", "obstacle_size", "joint_acc", "joint_jerk"])}

# every symbol include/panda_b200.h declares: (name, restype, argtypes)
P = c_void_p
SYMBOLS = [
    ("pg_default_config", c_int, [P, c_int, ctypes.c_char_p]),
    ("pg_stage_from_scenario", c_int, [P, ctypes.c_char_p]),
    ("pg_create", c_int, [P, c_int, P]),
    ("pg_destroy", c_int, [P]),
    ("pg_set_stage", c_int, [P, P, P]),
    ("pg_seed", c_int, [P, c_u64, c_u64]),
    ("pg_reset", c_int, [P, P, P, P]),
    ("pg_step", c_int, [P, P, P]),
    ("pg_step_host", c_int, [P, P, P, P]),
    ("pg_reset_host", c_int, [P, P, P]),
    ("pg_copy_out", c_int, [P, P, P]),
    ("pg_step_id", c_i64, [P]),
    ("pg_finished_count", c_int, [P, c_i64]),
    ("pg_fetch_finished", c_int, [P, c_i64, P, c_int, P, P]),
    ("pg_host_layout", c_int, [P, P, P]),
    ("pg_compute_reward", c_int, [P, P, P, P, c_i64, c_int, P]),
    ("pg_random_actions", c_int, [P, c_u64, c_u64, P, P]),
    ("pg_save_state", c_int, [P, P, P]),
    ("pg_restore_state", c_int, [P, c_int, P]),
    ("pg_remove_state", c_int, [P, c_int]),
    ("pg_get_view", c_int, [P, c_int, P]),
    ("pg_obs_dim", c_int, [P]),
    ("pg_act_dim", c_int, [P]),
    ("pg_num_envs", c_int, [P]),
    ("pg_launch_count", c_i64, [P]),
    ("pg_set_joint_state", c_int, [P, P, P, P]),
    ("pg_set_goal", c_int, [P, P, P]),
    ("pg_set_obstacles", c_int, [P, P, P]),
    ("pg_set_object_state", c_int, [P, P, P]),
    ("pg_observe", c_int, [P, P]),
    ("pg_manipulability", c_int, [P, P, c_i64, P, P]),
    ("pg_set_scene", c_int, [P, c_int, P, P]),
    ("pg_render", c_int, [P, c_int, c_int, c_int, P, c_float, c_float, c_float, c_float, P, P]),
    ("pg_set_eval_metrics", c_int, [P, c_int, P]),
    ("pg_get_eval_metrics", c_int, [P, P, P]),
    ("pg_debug_substeps", c_int, [c_int, c_int, P, P, P, P, c_int, c_int, P]),
    ("pg_debug_minv", c_int, [c_int, c_int, P, P, P]),
    ("pg_debug_ik", c_int, [c_int, c_int, P, P, P, P, P]),
    ("pg_debug_fma_peak", c_int, [c_int, c_int, c_int, P]),
    ("pg_debug_step_kernel_attrs", c_int, [c_int, P]),
    ("pg_her_create", c_int, [c_int, c_i64, c_int, c_int, c_int, c_int, c_i64, P]),
    ("pg_her_destroy", c_int, [P]),
    ("pg_her_add", c_int, [P] * 12),
    ("pg_her_set_mt_state", c_int, [P, P, c_int]),
    ("pg_her_get_mt_state", c_int, [P, P, P]),
    ("pg_her_sample", c_int, [P, P, c_i64, c_int, c_u64, c_int, P, P]),
    ("pg_her_sample_static", c_int, [P, P, c_i64, c_int, c_u64, c_int, P, P, P]),
    ("pg_her_last_indices", c_int, [P, c_i64, P, P, P, P, P]),
    ("pg_her_info", c_int, [P, P, P, P, P]),
    ("pg_last_error", ctypes.c_char_p, []),
    ("pg_version", ctypes.c_char_p, []),
]

_lib = None


class PandaB200Error(RuntimeError):
    pass


def load():
    """Load the CUDA library; raise loudly when it has not been built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PandaB200Error(
            "%s is missing: build the sm_100a CUDA extension first (python -c 'import __graft_entry__ as g; g.build()'). "
            "panda_gym_b200 has no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        f = getattr(lib, name)  # AttributeError if the .so is stale
        f.restype = res
        f.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise PandaB200Error("panda_b200 error %d: %s" % (rc, load().pg_last_error().decode()))
