"""Shared helpers of the parity tests: build the CUDA task and the CPU oracle from one synthetic state."""
import math

import numpy as np
import torch

from oracle import gogoro_oracle as go
from oracle import humanoid_oracle as ho
from thormang_isaacgym_b200 import config as tcfg
from thormang_isaacgym_b200 import synthetic as syn

# Parity tolerances (north_star: fp32 observations and rewards within 1e-5 relative).
# rtol = 1e-5 everywhere.  atol covers values near zero, where "relative" has no meaning:
#   * 2e-6 on ordinary columns (about 16 ulp at magnitude 1: sums with cancellation such as R^-1 v);
#   * angle columns are compared modulo 2*pi with atol 5e-6: the reference itself quantises angles to
#     ulp(2*pi) = 4.8e-7 through `% (2*pi)` (get_euler_xyz), and a 1-ulp libm difference in atan2 next to the
#     cut flips the representative by 2*pi (SURVEY.md section 7 "fp32 transcendental parity").
RTOL = 1e-5
ATOL = 2e-6
ANGLE_ATOL = 5e-6


def close(got, want, rtol=RTOL, atol=ATOL, what=""):
    got, want = got.detach().cpu().float(), want.detach().cpu().float()
    assert got.shape == want.shape, f"{what}: shape {tuple(got.shape)} vs {tuple(want.shape)}"
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    bad = err > tol
    if bool(bad.any()):
        idx = torch.nonzero(bad)[0].tolist()
        raise AssertionError(f"{what}: {int(bad.sum())} / {bad.numel()} outside tolerance; first at {idx}: "
                             f"got {got[tuple(idx)].item():.9g} want {want[tuple(idx)].item():.9g}; max err {err.max().item():.3g}")


PI_F32 = float(np.float32(math.pi))          # 3.14159274...: the largest magnitude fp32 atan2 / a +-2*pi wrap can return
CUT_TOL = 1e-4


TWO_PI_F32 = float(np.float32(2 * math.pi))


def angle_close(got, want, atol=ANGLE_ATOL, what="", bound=PI_F32, cut_tol=CUT_TOL, lo=None, flip_ok=None):
    """Angles agree modulo 2*pi within atol, AND (the modulo is only legitimate next to the cut)
    * every output lies in [lo, bound] (default [-pi_f32, pi_f32]; bound=None for channels whose range is open), and
    * got and want pick different representatives (|got - want| > pi) only where the reference value itself sits within
      cut_tol of an end of the range (or where ``flip_ok`` says the cut of an upstream angle is near) -- a kernel
      returning yaw + 2*pi away from the cut fails."""
    got, want = got.detach().cpu().double(), want.detach().cpu().double()
    d = torch.remainder(got - want + math.pi, 2 * math.pi) - math.pi
    assert float(d.abs().max()) <= atol, f"{what}: max angular error {float(d.abs().max()):.3g}"
    flipped = (got - want).abs() > math.pi
    allowed = torch.zeros_like(flipped) if flip_ok is None else flip_ok.detach().cpu().bool().expand_as(flipped)
    if bound is not None:
        lo = -bound if lo is None else lo
        eps = 2.0 ** -22 * max(abs(bound), abs(lo), 1.0)
        over = (got > bound + eps) | (got < lo - eps)
        assert not bool(over.any()), f"{what}: {int(over.sum())} outputs outside [{lo:.7g}, {bound:.7g}], e.g. {float(got[over][0]):.9g}"
        allowed = allowed | ((want - bound).abs() <= cut_tol) | ((want - lo).abs() <= cut_tol)
    elif flip_ok is None:
        return                      # open range and no knowledge of the upstream cut: modulo comparison only
    off_cut = flipped & ~allowed
    assert not bool(off_cut.any()), (f"{what}: {int(off_cut.sum())} values differ from the reference by 2*pi away from "
                                     f"the cut, e.g. got {float(got[off_cut][0]):.9g} want {float(want[off_cut][0]):.9g}")


def exact(got, want, what=""):
    got, want = got.detach().cpu(), want.detach().cpu()
    assert got.shape == want.shape and got.dtype == want.dtype, f"{what}: {got.dtype}{tuple(got.shape)} vs {want.dtype}{tuple(want.shape)}"
    if not torch.equal(got, want):
        bad = torch.nonzero(got != want)
        raise AssertionError(f"{what}: {bad.shape[0]} mismatches, first at {bad[0].tolist()}: "
                             f"{got[tuple(bad[0].tolist())].item()} vs {want[tuple(bad[0].tolist())].item()}")


HUMANOID_ANGLE_COLS = [7, 8, 9]


def humanoid_obs_close(got, want, what="obs"):
    rest = [c for c in range(got.shape[1]) if c not in HUMANOID_ANGLE_COLS]
    close(got[:, rest], want[:, rest], what=what)
    angle_close(got[:, HUMANOID_ANGLE_COLS], want[:, HUMANOID_ANGLE_COLS], what=what + "[angles]")


def gogoro_obs_close(got, want, what="obs", noisy=False):
    lin = [1, 2, 5] if noisy else [1, 2, 3, 5]
    close(got[:, lin], want[:, lin], what=what)
    # noisy rows carry N(0, sigma) on both angle channels: their range is open and the cut of the clean angle is not
    # visible in them (the clean rows of the same step are compared through buffer_obs with the range check on)
    angle_close(got[:, [0, 4]], want[:, [0, 4]], what=what + "[angles]", bound=None if noisy else PI_F32)
    if noisy:   # channel 3 = round(delta_yaw): exact unless delta_yaw sits within the angle tolerance of k + 0.5
        g, w = got[:, 3].cpu(), want[:, 3].cpu()
        assert float((g - w).abs().max()) <= 1.0 and float((g != w).float().mean()) < 1e-3, what + "[round]"


# ------------------------------------------------------------------------------------------------
def make_humanoid_pair(n, seed=1234, device="cuda:0", state=None, asset=None, start_rotation=None, sim_cls=None, **env_over):
    """(cuda task, cpu oracle, state) initialised from the same synthetic snapshot."""
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.humanoid import Humanoid
    asset = asset or syn.humanoid_asset()
    state = state or syn.humanoid_state(n, seed=seed, asset=asset)
    cfg = tcfg.humanoid_cfg(n, seed=seed)
    cfg["env"]["trackRandomizeBuf"] = True
    cfg["env"].update(env_over)
    sim = (sim_cls or FakeSim)(n, 21, device, num_sensor_floats=12, asset=asset)
    sim.load_state(state)
    cfg["sim_backend"] = sim
    env = Humanoid(cfg, device, device, -1, True)
    env.actions.copy_(state["actions"].to(device))
    env.progress_buf.copy_(state["progress_buf"].to(device))
    env.reset_buf.copy_(state["reset_buf"].to(device))
    env.potentials.copy_(state["potentials"].to(device))
    consts = ho.HumanoidConsts.from_cfg(cfg)
    orc = ho.HumanoidOracle(consts, state["root_states"].clone(), state["dof_state"].clone(), state["dof_force"].clone(),
                            state["sensors"].clone(), asset["dof_limits_lower"], asset["dof_limits_upper"],
                            asset["motor_efforts"])
    orc.actions = state["actions"].clone()
    orc.progress_buf = state["progress_buf"].clone()
    orc.reset_buf = state["reset_buf"].clone()
    orc.potentials = state["potentials"].clone()
    return env, orc, state


def make_gogoro_pair(n, seed=1234, device="cuda:0", state=None, sim_cls=None, **env_over):
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.gogoro_new import Gogoro
    state = state or syn.gogoro_state(n, seed=seed)
    cfg = tcfg.gogoro_cfg(n, seed=seed)
    cfg["env"].update(env_over)
    cfg["env"]["resetOnInit"] = False
    names = ["steering_joint", "rear_wheel_joint"] + [f"j{i}" for i in range(37)]
    sim = (sim_cls or FakeSim)(n, 39, device, num_sensor_floats=0, dof_names=names)
    sim.load_state(state)
    cfg["sim_backend"] = sim
    env = Gogoro(cfg, device, device, -1, True)
    consts = go.GogoroConsts.from_cfg(cfg, steering_joint=0, rear_wheel_joint=1)
    orc = go.GogoroOracle(consts, state["root_tensor"].clone(), state["state_dof"].clone(), state["thormang_pose"].clone())
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "imu_offsets",
                 "curent_speed_offset", "action_history", "progress_buf", "reset_buf", "thormang_pose"):
        getattr(env, name).copy_(state[name].to(device))
        setattr(orc, name, state[name].clone())
    # the oracle's views must alias its own dof tensor after the attribute swaps above
    orc.thormang_pose = state["thormang_pose"].clone()
    return env, orc, state


ANT_ANGLE_COLS = [7, 8, 9]


def ant_obs_close(got, want, what="obs"):
    """ant.py:395-400 concatenates yaw, roll (get_euler_xyz: [0, 2*pi], cut at 0 / 2*pi) and the raw angle_to_target
    (atan2 - yaw, open range) WITHOUT normalize_angle: the representative may flip only where yaw / roll sit at their cut."""
    rest = [c for c in range(got.shape[1]) if c not in ANT_ANGLE_COLS]
    close(got[:, rest], want[:, rest], what=what)
    angle_close(got[:, [7, 8]], want[:, [7, 8]], what=what + "[yaw, roll]", lo=0.0, bound=TWO_PI_F32)
    wy = want[:, 7].detach().cpu().double()
    yaw_at_cut = (wy <= CUT_TOL) | (wy >= TWO_PI_F32 - CUT_TOL)
    angle_close(got[:, 9], want[:, 9], what=what + "[angle_to_target]", bound=None, flip_ok=yaw_at_cut)


def make_ant_pair(n, seed=1234, device="cuda:0", state=None, asset=None, **env_over):
    """(cuda Ant task, cpu AntOracle, state) from one synthetic snapshot."""
    from oracle import ant_oracle as ao
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.ant import Ant
    asset = asset or syn.ant_asset()
    state = state or syn.ant_state(n, seed=seed, asset=asset)
    cfg = tcfg.ant_cfg(n, seed=seed)
    cfg["env"]["trackRandomizeBuf"] = True
    cfg["env"].update(env_over)
    sim = FakeSim(n, 8, device, num_sensor_floats=24, asset=asset)
    sim.load_state(state)
    cfg["sim_backend"] = sim
    env = Ant(cfg, device, device, -1, True)
    env.actions.copy_(state["actions"].to(device))
    env.progress_buf.copy_(state["progress_buf"].to(device))
    env.reset_buf.copy_(state["reset_buf"].to(device))
    env.potentials.copy_(state["potentials"].to(device))
    env.prev_potentials.copy_(state["prev_potentials"].to(device))
    consts = ao.AntConsts.from_cfg(cfg)
    orc = ao.AntOracle(consts, state["root_states"].clone(), state["dof_state"].clone(), state["sensors"].clone(),
                       asset["dof_limits_lower"], asset["dof_limits_upper"], asset["joint_gears"])
    orc.actions = state["actions"].clone()
    orc.progress_buf = state["progress_buf"].clone()
    orc.reset_buf = state["reset_buf"].clone()
    orc.potentials = state["potentials"].clone()
    orc.prev_potentials = state["prev_potentials"].clone()
    orc.randomize_buf = torch.zeros(n, dtype=torch.long)
    return env, orc, state


def gogoro_paper_frame_close(got, want, what="frame", noisy=False):
    """[N,8] frames: roll, yaw, d_roll, d_yaw, speed, delta_yaw, command, delay.  Noisy frames are normalised
    (angles / pi, :534-538), so the angle tolerance shrinks with them and the wrap period becomes 2."""
    if not noisy:
        close(got[:, [2, 3, 4, 6, 7]], want[:, [2, 3, 4, 6, 7]], what=what)
        angle_close(got[:, [0, 1, 5]], want[:, [0, 1, 5]], what=what + "[angles]")
    else:
        close(got[:, [1, 2, 3, 4, 6, 7]], want[:, [1, 2, 3, 4, 6, 7]], what=what)
        angle_close(got[:, [0, 5]] * math.pi, want[:, [0, 5]] * math.pi, atol=ANGLE_ATOL + 2e-6, what=what + "[angles/pi]",
                    bound=None)      # noise + offsets added after the wrap (:525-538): open range


def make_gogoro_paper_pair(n, seed=1234, device="cuda:0", state=None, turning=False, sim_cls=None, **env_over):
    """(cuda task, cpu oracle, state) of the 8x20 Gogoro variant from one synthetic snapshot;
    ``turning``: the gogoro_realistic_turning.py sibling (task ``GogoroTurning``, oracle ``GogoroTurningOracle``)."""
    from oracle import gogoro_paper_oracle as gp
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    if turning:
        from oracle import gogoro_turning_oracle as gt
        from thormang_isaacgym_b200.tasks.gogoro_realistic_turning import Gogoro as GogoroPaper
        env_over.setdefault("useSteerDelay", True)
    else:
        from thormang_isaacgym_b200.tasks.gogoro_realistic_turning_sim_paper import Gogoro as GogoroPaper
    state = state or syn.gogoro_paper_state(n, seed=seed)
    cfg = tcfg.gogoro_paper_cfg(n, seed=seed)
    cfg["env"].update(env_over)
    cfg["env"]["resetOnInit"] = False
    names = ["steering_joint", "rear_wheel_joint"] + [f"j{i}" for i in range(37)]
    sim = (sim_cls or FakeSim)(n, 39, device, num_sensor_floats=0, dof_names=names)
    sim.load_state(state)
    cfg["sim_backend"] = sim
    env = GogoroPaper(cfg, device, device, -1, True)
    consts = gp.GogoroPaperConsts.from_cfg(cfg, steering_joint=0, rear_wheel_joint=1)
    orc_cls = gt.GogoroTurningOracle if turning else gp.GogoroPaperOracle
    orc = orc_cls(consts, state["root_tensor"].clone(), state["state_dof"].clone(), state["thormang_pose"].clone())
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "curent_speed_offset",
                 "curent_imu_x_offset", "steer_delay", "command_history", "progress_buf", "reset_buf", "thormang_pose"):
        getattr(env, name).copy_(state[name].to(device))
        setattr(orc, name, state[name].clone())
    orc.buffer_obs = state["buffer_obs"].clone()
    orc.buffer_obs_noisy = state["buffer_obs_noisy"].clone()
    orc.obs_buf = orc.buffer_obs_noisy.flatten(start_dim=1)
    # device tick is 0: ring slot t holds the frame of time index t
    env.buffer_obs_ring.copy_(state["buffer_obs"].to(device))
    env.command_ring.copy_(state["buffer_obs"][:, :, 6].to(device))
    env.obs_buf.copy_(state["buffer_obs_noisy"].flatten(start_dim=1).to(device))
    return env, orc, state


def make_anymal_pair(n, seed=1234, device="cuda:0", state=None, sim_cls=None, **env_over):
    """(cuda Anymal task, cpu AnymalOracle, state) from one synthetic snapshot."""
    from oracle import anymal_oracle as ao
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.anymal import Anymal
    state = state or syn.anymal_state(n, seed=seed)
    cfg = tcfg.anymal_cfg(n, seed=seed)
    cfg["env"].update(env_over)
    cfg["env"]["resetOnInit"] = False
    sim = (sim_cls or FakeSim)(n, 12, device, num_sensor_floats=0, dof_names=list(tcfg.ANYMAL_DOF_NAMES), num_bodies=13)
    sim.load_state(state)
    cfg["sim_backend"] = sim
    env = Anymal(cfg, device, device, -1, True)
    for name in ("commands", "actions", "progress_buf", "reset_buf"):
        getattr(env, name).copy_(state[name].to(device))
    consts = ao.AnymalConsts.from_cfg(cfg)
    default_row = torch.tensor([cfg["env"]["defaultJointAngles"][k] for k in tcfg.ANYMAL_DOF_NAMES])
    b = cfg["env"]["baseInitState"]
    orc = ao.AnymalOracle(consts, state["root_states"].clone(), state["dof_state"].clone(), state["torques"].clone(),
                          state["contact_forces"].clone(), default_row,
                          torch.tensor(b["pos"] + b["rot"] + b["vLinear"] + b["vAngular"]))
    orc.commands = state["commands"].clone()
    orc.actions = state["actions"].clone()
    orc.progress_buf = state["progress_buf"].clone()
    orc.reset_buf = state["reset_buf"].clone()
    return env, orc, state


OP3_STATE_NAMES = ("prev_torques", "commands", "actions", "last_actions", "feet_air_time", "last_contacts", "potentials",
                   "prev_potentials", "table_potentials", "progress_buf", "reset_buf")


def make_ma_op3_pair(n, seed=1234, device="cuda:0", state=None, feet=None, **env_over):
    """(cuda MA_OP3 task, cpu MAOP3Oracle, state) from one synthetic snapshot.  ``feet``: [2][F] body
    indices replacing the four "*ank*" links (exercises the run-time foot count of the kernel)."""
    from oracle import ma_op3_oracle as mo
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.MA_OP3 import MA_OP3
    state = state or syn.ma_op3_state(n, seed=seed)
    cfg = tcfg.ma_op3_cfg(n, seed=seed)
    cfg["env"].update(env_over)
    cfg["env"]["resetOnInit"] = False
    if feet is not None:
        cfg["env"]["feetIndices"] = [list(map(int, f)) for f in feet]
        nf = len(feet[0])
        state = dict(state)
        state["feet_air_time"] = state["feet_air_time"][:, :, :1].repeat(1, 1, nf).contiguous()
        state["last_contacts"] = state["last_contacts"][:, :, :1].repeat(1, 1, nf).contiguous()
    nb = state["contact_forces"].numel() // (n * 3)
    sim = FakeSim(n, 44, device, num_sensor_floats=0, dof_names=list(tcfg.OP3_DOF_NAMES), num_bodies=nb, actors_per_env=3)
    init = state.get("initial_root_states")
    if init is None:
        # a distinct "spawn" snapshot so that reset writes are visible: the initial pose of MA_OP3.py:221-231 per env
        init = torch.zeros(n, 3, 13)
        init[:, :, 0:3] = torch.tensor([[-0.31, 0.0, 0.27], [0.30, 0.0, 0.27], [0.0, 0.0, 0.30]])
        init[:, :, 0:2] += state["root_states"].view(n, 3, 13)[:, 2:3, 0:2] * 0.5
        init[:, :, 6] = 1.0
    sim.load_state({"root_states": init.view(n * 3, 13)})
    cfg["sim_backend"] = sim
    env = MA_OP3(cfg, device, device, -1, True)           # initial_root_states := clone of the spawn snapshot (:99-100)
    sim.load_state(state)
    for name in OP3_STATE_NAMES:
        getattr(env, name).copy_(state[name].to(device))
    default_feet, grip = syn.ma_op3_body_indices()
    feet = default_feet if feet is None else torch.tensor(cfg["env"]["feetIndices"], dtype=torch.long)
    st = dict(state)
    st["initial_root_states"] = init.clone()
    orc = mo.oracle_from_state(st, consts=mo.OP3Consts.from_cfg(cfg), feet=feet, grip=grip)
    return env, orc, state
