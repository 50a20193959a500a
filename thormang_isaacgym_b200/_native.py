"""ctypes binding of ``libthormang_b200.so`` (C ABI: include/thormang_b200.h).

There is NO fallback: if the shared library is missing or a call fails, this module
raises.  Nothing here computes anything; it only marshals raw device pointers.
"""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("THORMANG_B200_LIB") or os.path.join(_PKG, "libthormang_b200.so")   # env: developer A/B builds

TK_MAX_DOF = 64
TK_STATS = 8
NOISE_NONE, NOISE_COMPACT, NOISE_DENSE, NOISE_PHILOX = 0, 1, 2, 3
VARIANT_HUMANOID, VARIANT_ANT = 0, 1

_f, _d, _i32, _i64, _u64, _p = C.c_float, C.c_double, C.c_int32, C.c_int64, C.c_uint64, C.c_void_p


class HumanoidParams(C.Structure):
    _fields_ = [
        ("device", _i32), ("num_envs", _i32), ("num_dof", _i32), ("variant", _i32),
        ("dt", _f), ("dof_vel_scale", _f), ("angular_velocity_scale", _f), ("contact_force_scale", _f),
        ("power_scale", _f), ("heading_weight", _f), ("up_weight", _f), ("actions_cost_scale", _f),
        ("energy_cost_scale", _f), ("joints_at_limit_cost_scale", _f), ("death_cost", _f),
        ("termination_height", _f), ("max_episode_length", _f), ("clip_actions", _f), ("max_motor_effort", _f),
        ("reserved1", _f),
        ("targets", _f * 3), ("inv_start_rot", _f * 4), ("basis_vec0", _f * 3), ("basis_vec1", _f * 3),
        ("reserved2", _f * 3),
        ("dof_limits_lower", _f * TK_MAX_DOF), ("dof_limits_upper", _f * TK_MAX_DOF),
        ("motor_efforts", _f * TK_MAX_DOF),
    ]


class HumanoidBuffers(C.Structure):
    _fields_ = [(name, _p) for name in (
        "root_states", "dof_state", "dof_force", "sensors", "actions", "obs_buf", "rew_buf", "reset_buf",
        "progress_buf", "timeout_buf", "randomize_buf", "potentials", "prev_potentials", "up_vec", "heading_vec",
        "initial_root_states", "initial_dof_pos", "episode_return")]


class ResetNoise(C.Structure):
    _fields_ = [("mode", _i32), ("reserved", _i32), ("unit_a", _p), ("unit_b", _p), ("rows", _i64),
                ("seed", _u64), ("step", _u64), ("env_id_offset", _i64)]


class GogoroParams(C.Structure):
    _fields_ = [
        ("device", _i32), ("num_envs", _i32), ("num_dof", _i32), ("steering_joint", _i32),
        ("rear_wheel_joint", _i32), ("max_episode_length", _i32), ("speed_freq_update", _i32),
        ("yaw_freq_update", _i32), ("max_steering", _f), ("max_steering_change", _f),
        ("imu_filter_noise", _d * 2), ("imu_noise", _d * 2), ("speed_sensor_noise", _d * 2),
        ("steering_action_noise", _d * 2), ("speed_range", _d * 2), ("seat_offset_x_range", _d * 2),
        ("seat_offset_y_range", _d * 2), ("seat_offset_z_range", _d * 2), ("seat_offset_xr_range", _d * 2),
        ("steering_offset", _d * 2), ("speed_sensor_offset", _d * 2),
    ]


class GogoroBuffers(C.Structure):
    _fields_ = [(name, _p) for name in (
        "root_tensor", "state_dof", "obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf", "timeout_buf",
        "yaw_command", "curent_command", "curent_speed", "steer_offsets", "imu_offsets", "curent_speed_offset",
        "action_history", "config_vector", "thormang_pose", "root_reset_tensor", "dof_pos_command",
        "dof_speed_command", "episode_return")]


class GogoroNoise(C.Structure):
    _fields_ = [("reset", ResetNoise), ("sensor_unit", _p), ("speed_unit", _p), ("yaw_unit", _p)]


GOGORO_PAPER_VARIANT_PAPER, GOGORO_PAPER_VARIANT_TURNING = 0, 1


class GogoroPaperParams(C.Structure):
    _fields_ = [
        ("device", _i32), ("num_envs", _i32), ("num_dof", _i32), ("steering_joint", _i32),
        ("rear_wheel_joint", _i32), ("max_episode_length", _i32), ("speed_freq_update", _i32),
        ("yaw_freq_update", _i32), ("buff_size", _i32), ("delay_width", _i32), ("use_steer_delay", _i32),
        ("variant", _i32), ("max_steering", _f), ("start_speed", _f),
        ("imu_filter_noise", _d * 2), ("imu_noise", _d * 2), ("speed_sensor_noise", _d * 2),
        ("speed_sensor_offset", _d * 2), ("imu_x_offset", _d * 2), ("steering_offset", _d * 2),
        ("command_delay", _d * 2), ("speed_range", _d * 2),
    ]


class GogoroPaperBuffers(C.Structure):
    _fields_ = [(name, _p) for name in (
        "root_tensor", "state_dof", "obs_buf", "buffer_obs_ring", "command_ring", "rew_buf", "reset_buf",
        "progress_buf", "timeout_buf", "yaw_command", "curent_command", "curent_speed", "steer_offsets",
        "curent_speed_offset", "curent_imu_x_offset", "speed_no_noise", "steer_delay", "command_history",
        "thormang_pose", "root_reset_tensor", "dof_pos_command", "dof_speed_command", "episode_return")]


class GogoroPaperNoise(C.Structure):
    _fields_ = [("mode", _i32), ("reserved", _i32), ("sensor_unit", _p), ("reset_unit", _p), ("speed_unit", _p),
                ("yaw_unit", _p), ("seed", _u64), ("step", _u64), ("env_id_offset", _i64)]


class AnymalParams(C.Structure):
    _fields_ = [
        ("device", _i32), ("num_envs", _i32), ("num_dof", _i32), ("num_bodies", _i32), ("max_episode_length", _i32),
        ("base_index", _i32), ("num_knee", _i32), ("knee_indices", _i32 * 8), ("reserved", _i32),
        ("dof_pos_scale", _f), ("dof_vel_scale", _f), ("action_scale", _f), ("clip_actions", _f),
        ("dt", _d), ("rew_lin_vel_xy", _d), ("rew_ang_vel_z", _d), ("rew_torque", _d),
        ("command_x_range", _d * 2), ("command_y_range", _d * 2), ("command_yaw_range", _d * 2),
        ("default_dof_pos", _f * 12), ("base_init_state", _f * 13), ("gravity_vec", _f * 3),
    ]


class AnymalBuffers(C.Structure):
    _fields_ = [(name, _p) for name in (
        "root_states", "dof_state", "torques", "contact_forces", "commands", "actions", "obs_buf", "rew_buf",
        "reset_buf", "progress_buf", "timeout_buf", "episode_return")]


class AnymalNoise(C.Structure):
    _fields_ = [("mode", _i32), ("reserved", _i32), ("unit_pos", _p), ("unit_vel", _p), ("unit_cmd", _p),
                ("seed", _u64), ("step", _u64), ("env_id_offset", _i64)]


OP3_AGENTS, OP3_DOF, OP3_OBS, OP3_MAX_FEET, OP3_MAX_GRIPPERS = 2, 22, 88, 8, 4


class Op3Params(C.Structure):
    _fields_ = [
        ("device", _i32), ("num_envs", _i32), ("num_agents", _i32), ("num_dof", _i32), ("num_bodies_total", _i32),
        ("max_episode_length", _i32), ("num_feet", _i32), ("num_grippers", _i32),
        ("feet_indices", (_i32 * OP3_MAX_FEET) * OP3_AGENTS), ("gripper_indices", (_i32 * OP3_MAX_GRIPPERS) * OP3_AGENTS),
        ("lin_vel_scale", _f), ("ang_vel_scale", _f), ("dof_pos_scale", _f), ("dof_vel_scale", _f), ("action_scale", _f),
        ("clip_actions", _f),
        ("dt", _d), ("rew_torque", _d), ("rew_up", _d), ("rew_air_time", _d), ("rew_syns_hip", _d), ("rew_no_fly", _d),
        ("rew_stand", _d), ("rew_action_rate", _d),
        ("command_x_range", _d * 2), ("command_y_range", _d * 2), ("command_yaw_range", _d * 2),
        ("default_dof_pos", _f * OP3_DOF), ("inv_start_rot", _f * 4), ("gravity_vec", _f * 3), ("up_vec", _f * 3),
        ("targets", _f * 3), ("goal_pos", _f * 3),
    ]


class Op3Buffers(C.Structure):
    _fields_ = [(name, _p) for name in (
        "root_states", "initial_root_states", "dof_state", "torques", "prev_torques", "contact_forces", "commands",
        "actions", "last_actions", "feet_air_time", "last_contacts", "potentials", "prev_potentials", "table_potentials",
        "prev_table_potentials", "obs_buf", "rew_buf", "reset_buf", "progress_buf", "timeout_buf", "episode_return",
        "grip_counts")]


class Op3Noise(C.Structure):
    _fields_ = [("mode", _i32), ("reserved", _i32), ("unit_cmd", _p), ("seed", _u64), ("step", _u64),
                ("env_id_offset", _i64)]


class NoiseLambda(C.Structure):
    _fields_ = [("distribution", _i32), ("operation", _i32), ("range", _d * 2), ("range_correlated", _d * 2)]


# every symbol include/thormang_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "tk_last_error": (C.c_char_p, []),
    "tk_abi_version": (C.c_int, []),
    "tk_device_info": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "tk_humanoid_create": (C.c_int, [C.POINTER(HumanoidParams), C.POINTER(_p)]),
    "tk_humanoid_destroy": (C.c_int, [_p]),
    "tk_humanoid_bind": (C.c_int, [_p, C.POINTER(HumanoidBuffers)]),
    "tk_humanoid_pre_physics": (C.c_int, [_p, _p, _p, _i32, _p]),
    "tk_humanoid_compute_observations": (C.c_int, [_p, _p]),
    "tk_humanoid_compute_reward": (C.c_int, [_p, _p]),
    "tk_humanoid_reset_idx": (C.c_int, [_p, _p, _i32, _p, _p, _p]),
    "tk_humanoid_post_physics": (C.c_int, [_p, C.POINTER(ResetNoise), _p]),
    "tk_humanoid_reward_terms": (C.c_int, [_p, _p, _p]),
    "tk_humanoid_wm_reward": (C.c_int, [_p, _p, _i64, _p, _p]),
    "tk_humanoid_philox_noise": (C.c_int, [_p, _u64, _u64, _i64, _p, _p, _p]),
    "tk_humanoid_stats": (C.c_int, [_p, _p, _i32, _p]),
    "tk_humanoid_advance_epoch": (C.c_int, [_p, _p]),
    "tk_gogoro_advance_epoch": (C.c_int, [_p, _p]),
    "tk_gogoro_create": (C.c_int, [C.POINTER(GogoroParams), C.POINTER(_p)]),
    "tk_gogoro_destroy": (C.c_int, [_p]),
    "tk_gogoro_bind": (C.c_int, [_p, C.POINTER(GogoroBuffers)]),
    "tk_gogoro_pre_physics": (C.c_int, [_p, _p, _p, _u64, _u64, _i64, _i32, _p]),
    "tk_gogoro_compute_observations": (C.c_int, [_p, _p, _p]),
    "tk_gogoro_compute_reward": (C.c_int, [_p, _p, _p]),
    "tk_gogoro_compute_obs_rwd": (C.c_int, [_p, _p, _u64, _u64, _i64, _p]),
    "tk_gogoro_reset_idx": (C.c_int, [_p, _p, _i32, _p, _p]),
    "tk_gogoro_post_physics": (C.c_int, [_p, C.POINTER(GogoroNoise), _p]),
    "tk_gogoro_stats": (C.c_int, [_p, _p, _i32, _p]),
    "tk_gogoro_philox_noise": (C.c_int, [_p, _u64, _u64, _i64, _p, _p, _p, _p, _p, _p]),
    "tk_gogoro_paper_create": (C.c_int, [C.POINTER(GogoroPaperParams), C.POINTER(_p)]),
    "tk_gogoro_paper_destroy": (C.c_int, [_p]),
    "tk_gogoro_paper_bind": (C.c_int, [_p, C.POINTER(GogoroPaperBuffers)]),
    "tk_gogoro_paper_pre_physics": (C.c_int, [_p, _p, _i32, _p]),
    "tk_gogoro_paper_compute_observations": (C.c_int, [_p, _p, _p]),
    "tk_gogoro_paper_compute_obs_rwd": (C.c_int, [_p, C.POINTER(GogoroPaperNoise), _p]),
    "tk_gogoro_paper_post_physics": (C.c_int, [_p, C.POINTER(GogoroPaperNoise), _p]),
    "tk_gogoro_paper_diff_sums": (C.c_int, [_p, _p, _p]),
    "tk_gogoro_paper_diff_sums_static": (C.c_int, [_p, _i64, _i32, _i32, _p, _p]),
    "tk_gogoro_paper_finalize_reward": (C.c_int, [_p, _p, _p]),
    "tk_gogoro_paper_reset_idx": (C.c_int, [_p, _p, _i32, _p, _p]),
    "tk_gogoro_paper_clean_frames": (C.c_int, [_p, _p, _p]),
    "tk_gogoro_paper_advance_epoch": (C.c_int, [_p, _p]),
    "tk_gogoro_paper_stats": (C.c_int, [_p, _p, _i32, _p]),
    "tk_gogoro_paper_philox_noise": (C.c_int, [_p, _u64, _u64, _i64, _p, _p, _p, _p, _p]),
    "tk_anymal_create": (C.c_int, [C.POINTER(AnymalParams), C.POINTER(_p)]),
    "tk_anymal_destroy": (C.c_int, [_p]),
    "tk_anymal_bind": (C.c_int, [_p, C.POINTER(AnymalBuffers)]),
    "tk_anymal_pre_physics": (C.c_int, [_p, _p, _p, _i32, _p]),
    "tk_anymal_compute_observations": (C.c_int, [_p, _p]),
    "tk_anymal_compute_reward": (C.c_int, [_p, _p]),
    "tk_anymal_reset_idx": (C.c_int, [_p, _p, _i32, _p, _p, _p, _p]),
    "tk_anymal_post_physics": (C.c_int, [_p, C.POINTER(AnymalNoise), _p]),
    "tk_anymal_advance_epoch": (C.c_int, [_p, _p]),
    "tk_anymal_stats": (C.c_int, [_p, _p, _i32, _p]),
    "tk_anymal_philox_noise": (C.c_int, [_p, _u64, _u64, _i64, _p, _p, _p, _p]),
    "tk_ma_op3_create": (C.c_int, [C.POINTER(Op3Params), C.POINTER(_p)]),
    "tk_ma_op3_destroy": (C.c_int, [_p]),
    "tk_ma_op3_bind": (C.c_int, [_p, C.POINTER(Op3Buffers)]),
    "tk_ma_op3_pre_physics": (C.c_int, [_p, _p, _p, _i32, _p]),
    "tk_ma_op3_grip_counts": (C.c_int, [_p, _p]),
    "tk_ma_op3_compute_observations": (C.c_int, [_p, _p]),
    "tk_ma_op3_compute_reward": (C.c_int, [_p, _p]),
    "tk_ma_op3_reset_idx": (C.c_int, [_p, _p, _i32, _p, _p]),
    "tk_ma_op3_post_physics": (C.c_int, [_p, C.POINTER(Op3Noise), _p]),
    "tk_ma_op3_finalize_reward": (C.c_int, [_p, _p]),
    "tk_ma_op3_advance_epoch": (C.c_int, [_p, _p]),
    "tk_ma_op3_stats": (C.c_int, [_p, _p, _i32, _p]),
    "tk_ma_op3_philox_noise": (C.c_int, [_p, _u64, _u64, _i64, _p, _p]),
    "tk_step_noise_clamp": (C.c_int, [_p, _i64, C.POINTER(NoiseLambda), _p, _p, _u64, _u64, _i64, _f, _p, _p, _p]),
    "tk_step_noise_units": (C.c_int, [_i64, _i32, _u64, _u64, _i64, _p, _p]),
    "tk_humanoid_set_obs_copy": (C.c_int, [_p, _p]),
    "tk_reset_ids_scratch_bytes": (_i64, [_i64]),
    "tk_reset_ids": (C.c_int, [_p, _i64, _p, _p, _p, _p]),
}

_lib = None


def lib() -> C.CDLL:
    """Load the extension (once).  Raises ImportError if it was not built: there is no CPU path."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -m thormang_isaacgym_b200.build` "
                "(thormang_isaacgym_b200 has no CPU fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)   # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        if handle.tk_abi_version() != 1:
            raise ImportError("libthormang_b200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().tk_last_error()
        raise RuntimeError(f"{what or 'thormang_b200'} failed (code {rc}): {msg.decode() if msg else '?'}")


def ptr(t) -> int:
    """Raw device pointer of a torch tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()


_device_index = {}


def current_stream_ptr(device) -> int:
    """cudaStream_t of torch's CURRENT stream on ``device`` ('cuda:1' | torch.device | int), as an integer.
    Goes through the raw getter (no torch.cuda.Stream object is built): this sits on the per-step launch path."""
    import torch
    idx = _device_index.get(device)
    if idx is None:
        d = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        idx = d.index if d.index is not None else torch.cuda.current_device()
        if d.index is not None:             # 'cuda' without an index follows the current device: never cached
            _device_index[device] = idx
    return torch._C._cuda_getCurrentRawStream(idx)
