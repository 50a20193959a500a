"""Seeded synthetic env states for tests and benchmarks.

There is no PhysX here, so "state after a physics step" is drawn from the
distributions listed in SURVEY.md 8(d): they put about a third of the humanoids
below the termination height, ~1 % of the envs into reset, a few joints beyond
98 % of their range, ~5 % of the scooters past the tilt limit, and force the
progress counters onto the episode-end / command-resample boundaries.

Everything is generated on the CPU with an explicit ``torch.Generator`` so that
rank r of a sharded run (seed 1234 + r) and the CPU oracle see the same numbers,
then moved to ``device``.
"""
import math
from typing import Dict

import torch

HUMANOID_DOF = 21
HUMANOID_SENSOR = 12
GOGORO_DOF = 39
GOGORO_HISTORY = 5


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return g


def _u(g, shape, lo, hi):
    return lo + (hi - lo) * torch.rand(shape, generator=g, dtype=torch.float32)


def _n(g, shape, std=1.0):
    return std * torch.randn(shape, generator=g, dtype=torch.float32)


def humanoid_asset(seed: int = 7, num_dof: int = HUMANOID_DOF) -> Dict[str, torch.Tensor]:
    """Per-task constants the reference reads from the MJCF asset (humanoid.py:160-171,205-215)."""
    g = _gen(seed)
    lower = -_u(g, (num_dof,), 0.5, 1.5)
    upper = _u(g, (num_dof,), 0.5, 1.5)
    efforts = _u(g, (num_dof,), 22.5, 135.0)
    return {"dof_limits_lower": lower, "dof_limits_upper": upper, "motor_efforts": efforts}


def humanoid_state(n: int, seed: int = 1234, asset: Dict[str, torch.Tensor] = None, dt: float = 0.0166,
                   device="cpu", reset_frac: float = 0.01, fallen_frac: float = None) -> Dict[str, torch.Tensor]:
    """One post-physics snapshot for ``n`` humanoids (config 1/2/4 of BASELINE.json).

    ``fallen_frac=None`` keeps the SURVEY 8(d) height distribution U(0.3, 1.8) (about a third below the
    termination height: good for parity coverage).  A number puts that fraction of the torsos below 0.8 and the
    rest in U(0.85, 1.8): the steady-state reset rate of a training run (0.1-3 % per step), used
    by the benchmarks."""
    asset = asset if asset is not None else humanoid_asset()
    lo, hi = asset["dof_limits_lower"], asset["dof_limits_upper"]
    d = lo.shape[0]
    g = _gen(seed)
    root = torch.empty((n, 13), dtype=torch.float32)
    root[:, 0:2] = _u(g, (n, 2), -5.0, 5.0)
    root[:, 2] = _u(g, (n,), 0.3, 1.8)
    if fallen_frac is not None:
        low = torch.rand((n,), generator=g) < fallen_frac
        root[:, 2] = torch.where(low, _u(g, (n,), 0.3, 0.79), _u(g, (n,), 0.85, 1.8))
    q = _n(g, (n, 4))
    root[:, 3:7] = q / q.norm(dim=-1, keepdim=True).clamp(min=1e-6)
    root[:, 7:10] = _n(g, (n, 3), 1.0)
    root[:, 10:13] = _n(g, (n, 3), 2.0)
    span = hi - lo
    dof_pos = (lo - 0.05 * span) + torch.rand((n, d), generator=g) * (1.1 * span)
    dof_vel = _n(g, (n, d), 5.0)
    dof_state = torch.stack((dof_pos, dof_vel), dim=-1).reshape(n * d, 2).contiguous()
    dof_force = _n(g, (n, d), 50.0)
    sensors = _n(g, (n, HUMANOID_SENSOR), 100.0)
    actions = _u(g, (n, d), -1.0, 1.0)
    progress = torch.randint(0, 1000, (n,), generator=g, dtype=torch.int64)
    forced = torch.rand((n,), generator=g) < 0.01
    progress = torch.where(forced, 998 + (torch.arange(n) % 2), progress)
    reset = (torch.rand((n,), generator=g) < reset_frac).to(torch.int64)
    potentials = -_u(g, (n,), 900.0, 1100.0) / dt
    out = dict(root_states=root, dof_state=dof_state, dof_force=dof_force, sensors=sensors, actions=actions,
               progress_buf=progress, reset_buf=reset, potentials=potentials)
    return {k: v.to(device) for k, v in out.items()}


def ant_asset(seed: int = 9) -> Dict[str, torch.Tensor]:
    """Per-task constants of the Ant asset (ant.py:160-162,199-209): 8 DOFs, joint gears as motor efforts."""
    a = humanoid_asset(seed=seed, num_dof=8)
    a["joint_gears"] = a.pop("motor_efforts")
    return a


def ant_state(n: int, seed: int = 1234, asset: Dict[str, torch.Tensor] = None, dt: float = 0.0166, device="cpu",
              fallen_frac: float = None) -> Dict[str, torch.Tensor]:
    """One post-physics snapshot for ``n`` ants: the Humanoid distributions at D = 8, 24 sensor floats, torso
    heights around the 0.31 termination height."""
    asset = asset if asset is not None else ant_asset()
    h = humanoid_state(n, seed=seed, asset={"dof_limits_lower": asset["dof_limits_lower"],
                                            "dof_limits_upper": asset["dof_limits_upper"]}, dt=dt)
    g = _gen(seed + 5)
    h["root_states"][:, 2] = _u(g, (n,), 0.1, 0.9)
    if fallen_frac is not None:
        low = torch.rand((n,), generator=g) < fallen_frac
        h["root_states"][:, 2] = torch.where(low, _u(g, (n,), 0.1, 0.3), _u(g, (n,), 0.35, 0.9))
    h["sensors"] = _n(g, (n, 24), 10.0)
    h["prev_potentials"] = h["potentials"] + _n(g, (n,), 5.0)
    h.pop("dof_force")
    return {k: v.to(device) for k, v in h.items()}


def humanoid_reset_noise(n: int, seed: int, num_dof: int = HUMANOID_DOF, device="cpu"):
    """U(0,1) draws for up to ``n`` resets: the [K, D] x 2 that humanoid.py:258-259 consume."""
    g = _gen(seed + 77)
    return (torch.rand((n, num_dof), generator=g).to(device), torch.rand((n, num_dof), generator=g).to(device))


def _euler_to_quat(roll, pitch, yaw):
    sr, cr = torch.sin(roll / 2), torch.cos(roll / 2)
    sp, cp = torch.sin(pitch / 2), torch.cos(pitch / 2)
    sy, cy = torch.sin(yaw / 2), torch.cos(yaw / 2)
    return torch.stack((sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy,
                        cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy), dim=-1)


def gogoro_state(n: int, seed: int = 1234, device="cpu", num_dof: int = GOGORO_DOF) -> Dict[str, torch.Tensor]:
    """One post-physics snapshot for ``n`` scooters (config 3 of BASELINE.json)."""
    g = _gen(seed)
    roll = _n(g, (n,), 0.15)
    pitch = _n(g, (n,), 0.05)
    yaw = _u(g, (n,), -math.pi, math.pi)
    root = torch.empty((n, 13), dtype=torch.float32)
    root[:, 0:2] = _u(g, (n, 2), -50.0, 50.0)
    root[:, 2] = _u(g, (n,), 0.0, 0.1)
    root[:, 3:7] = _euler_to_quat(roll, pitch, yaw)
    speed = _u(g, (n,), 4.0, 13.0)
    root[:, 7] = torch.cos(yaw) * speed
    root[:, 8] = torch.sin(yaw) * speed
    root[:, 9] = 0.0
    root[:, 7:10] += _n(g, (n, 3), 0.2)
    root[:, 10:13] = _n(g, (n, 3), 0.5)
    dof_state = _n(g, (n * num_dof, 2), 0.3)
    progress = torch.randint(0, 1000, (n,), generator=g, dtype=torch.int64)
    forced = torch.rand((n,), generator=g) < 0.02
    ladder = torch.tensor([298, 299, 300, 997, 998, 999], dtype=torch.int64)
    progress = torch.where(forced, ladder[torch.arange(n) % 6], progress)
    out = dict(
        root_tensor=root, state_dof=dof_state, progress_buf=progress,
        reset_buf=(torch.rand((n,), generator=g) < 0.01).to(torch.int64),
        yaw_command=_u(g, (n,), -math.pi, math.pi),
        curent_command=_u(g, (n,), -0.5, 0.5),
        curent_speed=_u(g, (n,), 4.0, 13.0),
        steer_offsets=_n(g, (n,), 0.01),
        imu_offsets=_n(g, (n,), 0.05),
        curent_speed_offset=_u(g, (n,), -0.5, 0.5),
        action_history=_u(g, (n, GOGORO_HISTORY), -1.0, 1.0),
        actions=_u(g, (n, 1), -1.5, 1.5),
        thormang_pose=_u(g, (1, num_dof), -1.5, 1.5).repeat(n, 1),
    )
    return {k: v.to(device) for k, v in out.items()}


def gogoro_noise(n: int, seed: int, device="cpu") -> Dict[str, torch.Tensor]:
    """Unit draws for one Gogoro step: N(0,1) steer [N], N(0,1) sensors [N,5], reset draws [N,10]
    (cols 0,2,3,4 uniform, others normal), U(0,1) speed [N] and yaw [N]."""
    g = _gen(seed + 991)
    reset_unit = torch.randn((n, 10), generator=g)
    uni = torch.rand((n, 10), generator=g)
    for col in (0, 2, 3, 4):
        reset_unit[:, col] = uni[:, col]
    out = dict(steer_unit=torch.randn((n,), generator=g), sensor_unit=torch.randn((n, 5), generator=g),
               reset_unit=reset_unit, speed_unit=torch.rand((n,), generator=g),
               yaw_unit=torch.rand((n,), generator=g))
    return {k: v.to(device) for k, v in out.items()}


def gogoro_paper_state(n: int, seed: int = 1234, device="cpu", num_dof: int = GOGORO_DOF, max_steps: int = 3600,
                       fallen_frac=None) -> Dict[str, torch.Tensor]:
    """One post-physics snapshot for ``n`` scooters of the 8-channel x 20-frame variant: the
    ``gogoro_state`` distributions plus the delay line, the per-env offsets and 20 frames of clean / noisy history
    (time-ordered; channel 6 of the clean frames is the command history the reward differentiates)."""
    base = gogoro_state(n, seed=seed, num_dof=num_dof)
    g = _gen(seed + 4001)
    if fallen_frac is not None:    # control the share of envs beyond the 0.38 rad fall threshold
        roll = torch.where(torch.rand((n,), generator=g) < fallen_frac, _u(g, (n,), 0.4, 0.6), _n(g, (n,), 0.1).clamp(-0.3, 0.3))
        yaw = _u(g, (n,), -math.pi, math.pi)
        base["root_tensor"][:, 3:7] = _euler_to_quat(roll, _n(g, (n,), 0.03), yaw)
    progress = torch.randint(0, max_steps, (n,), generator=g, dtype=torch.int64)
    forced = torch.rand((n,), generator=g) < 0.02
    ladder = torch.tensor([298, 299, 300, max_steps - 3, max_steps - 2, max_steps - 1], dtype=torch.int64)
    progress = torch.where(forced, ladder[torch.arange(n) % 6], progress)
    clean = torch.empty((n, 20, 8), dtype=torch.float32)
    clean[:, :, 0] = _n(g, (n, 20), 0.15)
    clean[:, :, 1] = _u(g, (n, 20), -math.pi, math.pi)
    clean[:, :, 2:4] = _n(g, (n, 20, 2), 0.5)
    clean[:, :, 4] = _u(g, (n, 20), 4.0, 13.0)
    clean[:, :, 5] = _u(g, (n, 20), -math.pi, math.pi)
    clean[:, :, 6] = _u(g, (n, 20), -0.5, 0.5)
    delay = torch.randint(0, 5, (n,), generator=g, dtype=torch.int64)
    clean[:, :, 7] = (delay.float() / 5.0).unsqueeze(1)
    noisy = clean + _n(g, (n, 20, 8), 0.01)
    noisy[:, :, 1] = 0.0
    out = dict(
        root_tensor=base["root_tensor"], state_dof=base["state_dof"], thormang_pose=base["thormang_pose"],
        progress_buf=progress, reset_buf=base["reset_buf"], yaw_command=base["yaw_command"],
        curent_command=base["curent_command"], curent_speed=_u(g, (n,), 5.0, 20.0), steer_offsets=_u(g, (n,), -0.05, 0.05),
        curent_speed_offset=_u(g, (n,), -0.3, 0.3), curent_imu_x_offset=_u(g, (n,), -0.02, 0.02),
        steer_delay=delay, command_history=_u(g, (n, 5), -0.5, 0.5),
        buffer_obs=clean, buffer_obs_noisy=noisy, actions=_u(g, (n, 1), -1.5, 1.5),
    )
    return {k: v.to(device) for k, v in out.items()}


def gogoro_paper_noise(n: int, seed: int, device="cpu", reset_cols: int = 5) -> Dict[str, torch.Tensor]:
    """U(0,1) draws of one step of the paper variant, addressed by env id: sensors [N,6], reset [N,5] (8 columns for the
    gogoro_realistic_turning.py sibling), speed [N], yaw [N]."""
    g = _gen(seed + 1993)
    out = dict(sensor_unit=torch.rand((n, 6), generator=g), reset_unit=torch.rand((n, reset_cols), generator=g),
               speed_unit=torch.rand((n,), generator=g), yaw_unit=torch.rand((n,), generator=g))
    return {k: v.to(device) for k, v in out.items()}


def anymal_state(n: int, seed: int = 1234, device="cpu", contact_frac: float = 0.02) -> Dict[str, torch.Tensor]:
    """One post-physics snapshot for ``n`` ANYmals: 12 DOFs, 13 bodies.  ~``contact_frac`` of the envs
    carry a base / thigh contact force above the 1 N termination threshold; feet (shank) contacts are common."""
    g = _gen(seed)
    root = torch.empty((n, 13), dtype=torch.float32)
    root[:, 0:2] = _u(g, (n, 2), -20.0, 20.0)
    root[:, 2] = _u(g, (n,), 0.4, 0.7)
    root[:, 3:7] = _euler_to_quat(_n(g, (n,), 0.1), _n(g, (n,), 0.1), _u(g, (n,), -math.pi, math.pi))
    root[:, 7:10] = _n(g, (n, 3), 1.0)
    root[:, 10:13] = _n(g, (n, 3), 0.7)
    contact = torch.zeros((n, 13, 3), dtype=torch.float32)
    feet = [3, 6, 9, 12]
    contact[:, feet, 2] = torch.where(torch.rand((n, 4), generator=g) < 0.6, _u(g, (n, 4), 20.0, 200.0), torch.zeros(n, 4))
    contact[:, feet, 0:2] = _n(g, (n, 4, 2), 5.0) * (contact[:, feet, 2:3] > 0)
    bad_body = torch.tensor([0, 2, 5, 8, 11])[torch.randint(0, 5, (n,), generator=g)]
    hit = torch.rand((n,), generator=g) < contact_frac
    mag = torch.where(torch.rand((n,), generator=g) < 0.5, _u(g, (n,), 0.5, 1.5), _u(g, (n,), 1.5, 80.0))   # both sides of 1 N
    dirv = torch.randn((n, 3), generator=g)
    dirv = dirv / dirv.norm(dim=-1, keepdim=True)
    idx = torch.nonzero(hit).flatten()
    contact[idx, bad_body[idx]] = dirv[idx] * mag[idx].unsqueeze(-1)
    progress = torch.randint(0, 2500, (n,), generator=g, dtype=torch.int64)
    forced = torch.rand((n,), generator=g) < 0.02
    ladder = torch.tensor([2497, 2498, 2499, 2500], dtype=torch.int64)
    progress = torch.where(forced, ladder[torch.arange(n) % 4], progress)
    out = dict(
        root_states=root, dof_state=torch.stack([_n(g, (n * 12,), 0.5), _n(g, (n * 12,), 3.0)], dim=-1),
        torques=_n(g, (n, 12), 20.0), contact_forces=contact,
        commands=torch.stack([_u(g, (n,), -2.0, 2.0), _u(g, (n,), -1.0, 1.0), _u(g, (n,), -1.0, 1.0)], dim=-1),
        actions=_u(g, (n, 12), -1.5, 1.5), progress_buf=progress,
        reset_buf=(torch.rand((n,), generator=g) < 0.01).to(torch.int64),
    )
    return {k: v.to(device) for k, v in out.items()}


def anymal_reset_noise(n: int, seed: int, device="cpu"):
    """U(0,1) draws of Anymal.reset_idx addressed by env id: positions [N,12], velocities [N,12], commands [N,3]."""
    g = _gen(seed + 313)
    return tuple(t.to(device) for t in (torch.rand((n, 12), generator=g), torch.rand((n, 12), generator=g),
                                        torch.rand((n, 3), generator=g)))


OP3_DOF = 22
OP3_AGENTS = 2
OP3_BODIES = 25          # rigid bodies of one OP3 in the synthetic layout (the URDF is not in the reference tree)
OP3_FEET = (10, 11, 16, 17)      # the four "*ank*" links of one robot (MA_OP3.py:237-238), actor-local
OP3_GRIPPERS = (20, 24)          # the two "*gr*" links (:241-242), actor-local: [left, right]


def ma_op3_body_indices():
    """(feet [2,4], grippers [2,2]) into ``contact_forces.view(N, -1, 3)``: agent a's bodies start at a * OP3_BODIES."""
    feet = torch.tensor([[a * OP3_BODIES + b for b in OP3_FEET] for a in range(OP3_AGENTS)], dtype=torch.long)
    grip = torch.tensor([[a * OP3_BODIES + b for b in OP3_GRIPPERS] for a in range(OP3_AGENTS)], dtype=torch.long)
    return feet, grip


def ma_op3_state(n: int, seed: int = 1234, device="cpu", far_frac: float = 0.03) -> Dict[str, torch.Tensor]:
    """One post-physics snapshot for ``n`` MA_OP3 envs: two 22-DOF OP3 + one table per env,
    2 * 25 + 1 rigid bodies.  A few percent of the robots are tilted past up_proj 0.90 or further than 0.40 m from the
    table, a few tables are tilted / below 0.25 m; foot forces straddle the 0.1 / 1.1 / 450 N thresholds, gripper forces
    the 0.1 / 400 N ones; progress counters are forced onto 0 and the episode end (3012)."""
    g = _gen(seed)
    a, d, nb = OP3_AGENTS, OP3_DOF, OP3_AGENTS * OP3_BODIES + 1
    root = torch.zeros((n, 3, 13), dtype=torch.float32)
    base = torch.tensor([[-0.31, 0.0, 0.27], [0.30, 0.0, 0.27], [0.0, 0.0, 0.30]])
    origin = torch.cat([_u(g, (n, 1, 2), -2.0, 2.0), torch.zeros(n, 1, 1)], dim=-1)
    spread = torch.where(torch.rand((n, 3, 1), generator=g) < far_frac, torch.tensor(0.12), torch.tensor(0.03))
    root[:, :, 0:3] = base + origin + _n(g, (n, 3, 3)) * spread
    tilt = torch.where(torch.rand((n, 3), generator=g) < 0.02, torch.tensor(0.5), torch.tensor(0.08))
    for k in range(3):
        root[:, k, 3:7] = _euler_to_quat(_n(g, (n,)) * tilt[:, k], _n(g, (n,)) * tilt[:, k], _u(g, (n,), -math.pi, math.pi))
    root[:, :, 7:10] = _n(g, (n, 3, 3), 0.5)
    root[:, :, 10:13] = _n(g, (n, 3, 3), 1.0)
    contact = torch.zeros((n, nb, 3), dtype=torch.float32)
    feet, grip = ma_op3_body_indices()
    for i in range(a):
        fz = torch.where(torch.rand((n, 4), generator=g) < 0.5, _u(g, (n, 4), 0.0, 2.0), _u(g, (n, 4), 2.0, 60.0))
        fz = torch.where(torch.rand((n, 4), generator=g) < 0.03, _u(g, (n, 4), 430.0, 470.0), fz)
        fz = torch.where(torch.rand((n, 4), generator=g) < 0.35, torch.zeros(n, 4), fz)
        contact[:, feet[i], 2] = fz
        contact[:, feet[i], 0:2] = _n(g, (n, 4, 2), 5.0) * (fz.unsqueeze(-1) > 0)
        gx = torch.where(torch.rand((n, 2), generator=g) < 0.02, _u(g, (n, 2), 0.0, 30.0), torch.zeros(n, 2))
        gx = torch.where(torch.rand((n, 2), generator=g) < 0.005, _u(g, (n, 2), 390.0, 410.0), gx)
        contact[:, grip[i], 0] = gx
        contact[:, grip[i], 1:3] = _n(g, (n, 2, 2), 1.0) * (gx.unsqueeze(-1) > 0)
    other = torch.rand((n, nb), generator=g) < 0.05
    contact += (_n(g, (n, nb, 3), 10.0) * other.unsqueeze(-1)) * (contact.abs().sum(-1, keepdim=True) == 0)
    cmd_y = torch.where(torch.rand((n,), generator=g) < 0.1, _u(g, (n,), 0.0, 0.2), _u(g, (n,), 0.0, 10.0))
    progress = torch.randint(1, 3012, (n,), generator=g, dtype=torch.int64)
    forced = torch.rand((n,), generator=g) < 0.03
    ladder = torch.tensor([0, 3010, 3011, 3012, 3013, 1], dtype=torch.int64)
    progress = torch.where(forced, ladder[torch.arange(n) % 6], progress)
    air = torch.where(torch.rand((n, a, 4), generator=g) < 0.3, torch.zeros(n, a, 4), _u(g, (n, a, 4), 0.0, 1.2))
    # potentials of the previous step: the same closed form on slightly displaced positions (dt = 0.0166)
    goal, target = torch.tensor([0.0, -10.0, 0.0]), torch.tensor([0.0, 10.0, 0.0])
    prev_pos = root[:, 0:2, 0:3] + _n(g, (n, 2, 3), 0.004)
    to_goal = goal - prev_pos
    to_goal[..., 2] = 0.0
    pot = -to_goal.norm(dim=-1) / 0.0166
    tpot = -(target - (root[:, 2, 0:3] + _n(g, (n, 3), 0.004))).norm(dim=-1) / 0.0166
    out = dict(
        root_states=root.view(n * 3, 13), dof_state=torch.stack([_n(g, (n * a * d,), 0.5), _n(g, (n * a * d,), 3.0)], dim=-1),
        torques=_n(g, (n, a, d), 2.0), prev_torques=_n(g, (n, a, d), 2.0), contact_forces=contact,
        commands=torch.stack([torch.zeros(n), cmd_y, torch.zeros(n)], dim=-1),
        actions=_u(g, (n, a, d), -1.5, 1.5), last_actions=_u(g, (n, a, d), -1.0, 1.0),
        feet_air_time=air, last_contacts=torch.rand((n, a, 4), generator=g) < 0.5,
        potentials=pot, prev_potentials=pot + _n(g, (n, a), 0.3), table_potentials=tpot, progress_buf=progress,
        reset_buf=(torch.rand((n,), generator=g) < 0.01).to(torch.int64),
    )
    return {k: v.to(device) for k, v in out.items()}


def ma_op3_reset_noise(n: int, seed: int, device="cpu") -> torch.Tensor:
    """U(0,1) draws of MA_OP3.reset_idx addressed by env id: commands [N,3] (x, y, yaw)."""
    return torch.rand((n, 3), generator=_gen(seed + 717)).to(device)
