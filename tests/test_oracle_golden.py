"""Oracle (CPU restatement) against the golden vectors frozen from the reference's own JIT functions."""
import os

import numpy as np
import pytest
import torch

from oracle import gogoro_oracle as go
from oracle import humanoid_oracle as ho
from oracle import isaac_helpers as ih

# The golden vectors were produced by the same torch CPU operators the oracle uses, so on the
# machine that made them the match is bit-exact.  Another host CPU may pick a different SIMD
# path for atan2/sin/cos: allow 4 ulp-ish slack there, integers stay exact.
RTOL, ATOL = 2e-6, 2e-6


def _load(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def _wrap_close(a, b, atol):
    d = torch.remainder(a - b + np.pi, 2 * np.pi) - np.pi
    return bool((d.abs() <= atol).all())


def test_humanoid_observations_golden(golden_dir):
    g = _load(golden_dir, "humanoid_jit.npz")
    n, d = g["root_states"].shape[0], 21
    dof_pos = g["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = g["dof_state"].view(n, d, 2)[..., 1]
    targets = torch.tensor([1000.0, 0.0, 0.0]).repeat(n, 1)
    b0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    b1 = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    for inv, key, upk, hk in ((torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(n, 1), "obs", "up_vec", "heading_vec"),
                              (g["inv_start_rot2"], "obs2", "up_vec2", "heading_vec2")):
        obs, pot, prev, up, head = ho.humanoid_observations(
            g["root_states"], targets, g["potentials"], inv, dof_pos, dof_vel, g["dof_force"],
            g["dof_limits_lower"], g["dof_limits_upper"], 0.1, g["sensors"], g["actions"], 0.0166, 0.01, 0.25, b0, b1)
        ang = [7, 8, 9]
        rest = [c for c in range(110) if c not in ang]
        torch.testing.assert_close(obs[:, rest], g[key][:, rest], rtol=RTOL, atol=ATOL)
        assert _wrap_close(obs[:, ang], g[key][:, ang], 4e-6)
        torch.testing.assert_close(up, g[upk], rtol=RTOL, atol=ATOL)
        torch.testing.assert_close(head, g[hk], rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(pot, g["potentials_out"], rtol=RTOL, atol=0)
    assert torch.equal(prev, g["prev_potentials_out"])


def test_humanoid_reward_golden(golden_dir):
    g = _load(golden_dir, "humanoid_jit.npz")
    eff = g["motor_efforts"]
    rew, reset = ho.humanoid_reward(g["obs"], g["reset_buf"], g["progress_buf"], g["actions"], 0.1, 0.5,
                                    g["potentials_out"], g["prev_potentials_out"], 0.01, 0.05, 0.25,
                                    float(eff.max()), eff, 0.8, -1.0, 1000.0)
    assert torch.equal(reset, g["reset_out"]) and reset.dtype == torch.int64
    assert torch.equal(rew, g["rew"])
    assert set(rew.tolist()) <= {-1.0, 2.0, float(np.float32(2.0) + np.float32(0.1))}
    n = g["obs"].shape[0]
    wm = ho.humanoid_wm_reward(g["obs"].view(2, n // 2, 110), g["actions"].view(2, n // 2, 21), 0.1, 0.5, 0.01, 0.05,
                               0.25, float(eff.max()), eff, 0.8, -1.0)
    torch.testing.assert_close(wm, g["wm_reward"], rtol=0, atol=0)


def test_humanoid_scripted_matches_eager(golden_dir):
    g = _load(golden_dir, "humanoid_jit.npz")
    eff = g["motor_efforts"]
    args = (g["obs"], g["reset_buf"], g["progress_buf"], g["actions"], 0.1, 0.5, g["potentials_out"],
            g["prev_potentials_out"], 0.01, 0.05, 0.25, float(eff.max()), eff, 0.8, -1.0, 1000.0)
    r0, s0 = ho.humanoid_reward(*args)
    r1, s1 = ho.scripted("humanoid_reward")(*args)
    assert torch.equal(r0, r1) and torch.equal(s0, s1)


def test_gogoro_golden(golden_dir):
    g = _load(golden_dir, "gogoro_jit.npz")
    n = g["root_tensor"].shape[0]
    obs = go.gogoro_observations(g["root_tensor"].clone(), g["desired_yaw"].clone(), g["last_command"].clone())
    lin = [1, 2, 3, 5]
    torch.testing.assert_close(obs[:, lin], g["obs"][:, lin], rtol=RTOL, atol=ATOL)
    assert _wrap_close(obs[:, [0, 4]], g["obs"][:, [0, 4]], 4e-6)
    rew, reset = go.gogoro_reward(g["obs"].view(n, 1, 6), g["progress_buf"], g["action_history"], 1000)
    assert torch.equal(reset, g["reset"]) and reset.dtype == torch.int64
    torch.testing.assert_close(rew, g["rew"], rtol=RTOL, atol=ATOL)
    sad = go.shortest_angle_distance(g["sad_y1"], g["sad_y2"])
    torch.testing.assert_close(sad, g["sad"], rtol=RTOL, atol=ATOL)
    assert float(sad.min()) >= -np.pi - 1e-6 and float(sad.max()) <= np.pi + 1e-6


def test_helpers_against_poselib(golden_dir):
    """isaac_helpers (restated, unpinned upstream) vs the reference's in-tree rotation3d.py, fp64."""
    g = _load(golden_dir, "poselib_quat.npz")
    a, b, v = g["a"], g["b"], g["v"]
    torch.testing.assert_close(ih.quat_mul(a, b), g["quat_mul"], rtol=0, atol=1e-14)
    torch.testing.assert_close(ih.quat_conjugate(a), g["quat_conjugate"], rtol=0, atol=0)
    torch.testing.assert_close(ih.quat_rotate(a, v), g["quat_rotate"], rtol=0, atol=1e-13)
    torch.testing.assert_close(ih.quat_apply(a, v), g["quat_rotate"], rtol=0, atol=1e-13)
    torch.testing.assert_close(ih.quat_rotate_inverse(a, v), g["quat_rotate_conj"], rtol=0, atol=1e-13)


def test_helper_properties():
    g = torch.Generator().manual_seed(3)
    q = torch.randn((512, 4), generator=g); q = q / q.norm(dim=-1, keepdim=True)
    v = torch.randn((512, 3), generator=g)
    back = ih.quat_rotate_inverse(q, ih.quat_rotate(q, v))
    torch.testing.assert_close(back, v, rtol=1e-5, atol=1e-5)
    roll, pitch, yaw = ih.get_euler_xyz(q)
    for ang in (roll, pitch, yaw):
        assert float(ang.min()) >= 0.0 and float(ang.max()) <= float(np.float32(2 * np.pi))
    # Euler -> quat (reference's in-file euler_to_quaternion, gogoro_new.py:496-502) -> Euler round trip
    r = (torch.rand(512, generator=g) * 2 - 1) * 3.0
    p = (torch.rand(512, generator=g) * 2 - 1) * 1.4
    y = (torch.rand(512, generator=g) * 2 - 1) * 3.0
    r2, p2, y2 = ih.get_euler_xyz(go.euler_to_quaternion(r, p, y))
    for x, x2 in ((r, r2), (p, p2), (y, y2)):
        dd = torch.remainder(x2 - x + np.pi, 2 * np.pi) - np.pi
        assert float(dd.abs().max()) < 5e-5
    lo = -torch.rand(21, generator=g) - 0.5
    hi = torch.rand(21, generator=g) + 0.5
    # (2*lo - hi) - lo rounds twice: end point -1 holds to 1 ulp, +1 holds exactly
    torch.testing.assert_close(ih.unscale(lo, lo, hi), -torch.ones(21), rtol=0, atol=2.4e-7)
    assert torch.equal(ih.unscale(hi, lo, hi), torch.ones(21))
    x = torch.linspace(-10, 10, 1001)
    na = ih.normalize_angle(x)
    assert float(na.min()) >= -np.pi - 1e-6 and float(na.max()) <= np.pi + 1e-6
    assert ih.get_axis_params(1., 2) == [0.0, 0.0, 1.0]
    assert ih.get_axis_params(1.34, 2) == [0.0, 0.0, 1.34]


def test_ant_golden(golden_dir):
    from oracle import ant_oracle as ao
    g = _load(golden_dir, "ant_jit.npz")
    n, d = g["root_states"].shape[0], 8
    dof_pos = g["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = g["dof_state"].view(n, d, 2)[..., 1]
    targets = torch.tensor([1000.0, 0.0, 0.0]).repeat(n, 1)
    inv = torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(n, 1)
    b0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    b1 = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    obs, pot, prev, up, head = ao.ant_observations(g["root_states"], targets, g["potentials"], inv, dof_pos, dof_vel,
                                                   g["dof_limits_lower"], g["dof_limits_upper"], 0.2, g["sensors"],
                                                   g["actions"], 0.0166, 0.1, b0, b1, 2)
    ang = [7, 8, 9]
    rest = [c for c in range(60) if c not in ang]
    torch.testing.assert_close(obs[:, rest], g["obs"][:, rest], rtol=RTOL, atol=ATOL)
    assert _wrap_close(obs[:, ang], g["obs"][:, ang], 4e-6)
    torch.testing.assert_close(pot, g["potentials_out"], rtol=RTOL, atol=0)
    rew, reset = ao.ant_reward(g["obs"], g["reset_buf"], g["progress_buf"], g["actions"], 0.1, 0.5, g["potentials_out"],
                               g["prev_potentials_out"], 0.005, 0.05, 0.1, 0.31, -2.0, 1000.0)
    assert torch.equal(reset, g["reset_out"])
    torch.testing.assert_close(rew, g["rew"], rtol=RTOL, atol=ATOL)
    assert float((g["obs"][:, 12:20] > 0.99).sum()) > 0, "fixture must exercise the joints-at-limit term"


def test_gogoro_paper_golden(golden_dir):
    """8-channel x 20-frame Gogoro variant against gogoro_realistic_turning_sim_paper.py's JIT functions."""
    from oracle import gogoro_paper_oracle as gp
    g = _load(golden_dir, "gogoro_paper_jit.npz")
    obs = gp.paper_observations(g["root_tensor"].clone(), g["desired_yaw"].clone(), g["curent_command"].clone(),
                                (g["steer_delay"].clone() - 0) / (5 - 0))
    lin = [2, 3, 4, 6, 7]
    torch.testing.assert_close(obs[:, lin], g["obs"][:, lin], rtol=RTOL, atol=ATOL)
    assert _wrap_close(obs[:, [0, 1, 5]], g["obs"][:, [0, 1, 5]], 4e-6)
    rew, reset = gp.paper_reward(g["frames"][:, :, 6], g["frames"], g["progress_buf"], 3600)
    assert torch.equal(reset, g["reset"]) and reset.dtype == torch.int64
    torch.testing.assert_close(rew, g["rew"], rtol=RTOL, atol=ATOL)
    assert float(rew.min()) == -1.0 and float(rew[rew >= 0].max()) <= 0.45 + 0.1 + 0.35 + 2.0 + 0.2 + 1e-6


def test_gogoro_paper_oracle_step_semantics():
    """Method-level restatement: frame shift, reset zeroing, delay-line tap, command resampling on the hit set."""
    from oracle import gogoro_paper_oracle as gp
    from thormang_isaacgym_b200 import synthetic as syn
    n = 64
    st = syn.gogoro_paper_state(n, seed=5)
    orc = gp.GogoroPaperOracle(gp.GogoroPaperConsts(), st["root_tensor"].clone(), st["state_dof"].clone(),
                               st["thormang_pose"].clone())
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "curent_speed_offset",
                 "curent_imu_x_offset", "steer_delay", "command_history", "progress_buf", "reset_buf"):
        setattr(orc, name, st[name].clone())
    orc.buffer_obs, orc.buffer_obs_noisy = st["buffer_obs"].clone(), st["buffer_obs_noisy"].clone()
    orc.reset_buf[:] = 0
    orc.reset_buf[[3, 17]] = 1
    orc.progress_buf[:] = 10
    orc.progress_buf[5] = 299
    nz = syn.gogoro_paper_noise(n, seed=1)
    hist_before = orc.command_history.clone()
    orc.pre_physics_step(st["actions"])
    cmd = torch.clamp(st["actions"][:, 0], -1, 1) * 0.5
    assert torch.equal(orc.command_history[:, -1], cmd) and torch.equal(orc.command_history[:, :-1], hist_before[:, 1:])
    assert torch.equal(orc.dof_pos_command[:, 0], orc.command_history[:, -3])
    old_noisy = orc.buffer_obs_noisy.clone()
    ids, sp, yw = orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"][[3, 17]], nz["speed_unit"][:1], nz["yaw_unit"])
    assert ids.tolist() == [3, 17] and sp.tolist() == [5] and yw.tolist() == [5]
    keep = [i for i in range(n) if i not in (3, 17)]
    nzy = orc.buffer_obs_noisy
    assert torch.equal(nzy[keep, :-1][..., [0, 2, 3, 4, 5, 6, 7]], old_noisy[keep, 1:][..., [0, 2, 3, 4, 5, 6, 7]])   # one-frame shift
    assert float(nzy[[3, 17], :-1].abs().max()) == 0.0                                 # reset zeroes the history
    assert float(nzy[:, :, 1].abs().max()) == 0.0                                      # :544
    assert orc.obs_buf.shape == (n, 160) and orc.obs_buf.data_ptr() == nzy.data_ptr()  # flatten() view, :545
    assert torch.equal(orc.root_tensor[3, 3:7], torch.tensor([0.0, 0.0, 0.0, 1.0]))
    assert float(orc.root_tensor[3, 7]) == float(np.float32(1.3)) and float(orc.root_tensor[3, 8]) == 0.0
    assert int(orc.progress_buf[3]) == 0 and float(orc.yaw_command[3]) == 0.0
    want_speed = 5.0 + nz["speed_unit"][0] * (20.0 - 5.0)
    assert torch.equal(orc.curent_speed[5], want_speed)


def test_gogoro_turning_golden(golden_dir):
    """gogoro_realistic_turning.py (the rational-reward sibling of the paper file) against its own JIT functions."""
    from oracle import gogoro_turning_oracle as gt
    g = _load(golden_dir, "gogoro_turning_jit.npz")
    obs = gt.paper_observations(g["root_tensor"].clone(), g["desired_yaw"].clone(), g["curent_command"].clone(),
                                (g["steer_delay"].clone() - 0) / (5 - 0))
    lin = [2, 3, 4, 6, 7]
    torch.testing.assert_close(obs[:, lin], g["obs"][:, lin], rtol=RTOL, atol=ATOL)
    assert _wrap_close(obs[:, [0, 1, 5]], g["obs"][:, [0, 1, 5]], 4e-6)
    rew, reset = gt.turning_reward(g["frames"][:, :, 6], g["frames"], g["progress_buf"], 3600)
    assert torch.equal(reset, g["reset"]) and reset.dtype == torch.int64
    torch.testing.assert_close(rew, g["rew"], rtol=RTOL, atol=ATOL)
    mean_term = 10.0 * (1.0 - abs(float(g["frames"][:, :, 6].mean())))
    assert 9.0 < mean_term < 9.5, "fixture must put the all-env command mean away from zero"
    assert float(rew.min()) == -1.0 and float(rew[rew >= 0].min()) >= 10.0 + mean_term - 1e-4      # reward3 == 1 always
    assert float(rew.max()) <= 3.0 + 1.0 + 10.0 + 2.0 + mean_term + 1e-4


def test_gogoro_turning_oracle_step_semantics():
    """What the sibling file changes: steering target + offsets on the per-env tap, random spawn, no correction terms."""
    from oracle import gogoro_turning_oracle as gt
    from thormang_isaacgym_b200 import synthetic as syn
    n = 64
    st = syn.gogoro_paper_state(n, seed=6)
    orc = gt.GogoroTurningOracle(gt.turning_consts(), st["root_tensor"].clone(), st["state_dof"].clone(),
                                 st["thormang_pose"].clone())
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "curent_speed_offset",
                 "curent_imu_x_offset", "steer_delay", "command_history", "progress_buf", "reset_buf"):
        setattr(orc, name, st[name].clone())
    orc.buffer_obs, orc.buffer_obs_noisy = st["buffer_obs"].clone(), st["buffer_obs_noisy"].clone()
    orc.reset_buf[:] = 0
    orc.reset_buf[[3, 17]] = 1
    orc.progress_buf[:] = 10
    nz = syn.gogoro_paper_noise(n, seed=1, reset_cols=gt.RESET_NOISE_COLS)
    orc.pre_physics_step(st["actions"])
    tap = orc.command_history[torch.arange(n), -orc.steer_delay]
    assert torch.equal(orc.dof_pos_command[:, 0], tap + orc.steer_offsets)            # :360, :365
    ids, _, _ = orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"][[3, 17]], nz["speed_unit"][:0], nz["yaw_unit"])
    assert ids.tolist() == [3, 17]
    u = nz["reset_unit"][[3, 17]]
    tgt = (u[:, 5] * 2 - 1) * np.pi
    assert torch.equal(orc.yaw_command[[3, 17]], tgt)                                 # :573, :683 (|tgt| <= pi: no wrap)
    q = orc.root_tensor[[3, 17], 3:7]
    torch.testing.assert_close(q.norm(dim=-1), torch.ones(2), rtol=1e-6, atol=1e-6)
    rot = tgt + (-np.pi + u[:, 6] * (2 * np.pi))
    torch.testing.assert_close(orc.root_tensor[[3, 17], 7], 1.3 * torch.cos(rot), rtol=1e-6, atol=1e-6)
    nzy = orc.buffer_obs_noisy
    assert float(nzy[[3, 17], :-1].abs().max()) == 0.0 and float(nzy[:, :, 1].abs().max()) == 0.0
    # no command-change correction (the paper file's :541-542): channel 2 of the fresh noisy frame = (d_roll + noise) / 3
    keep = [i for i in range(n) if i not in (3, 17)]
    want2 = (orc.buffer_obs[keep, -1, 2] + (-0.003 + nz["sensor_unit"][keep, 2] * 0.006)) / 3.0
    torch.testing.assert_close(nzy[keep, -1, 2], want2, rtol=1e-6, atol=1e-7)


def test_anymal_golden(golden_dir):
    """Anymal against anymal.py's own JIT functions; contact norms exactly at / 1 ulp around 1 N."""
    from oracle import anymal_oracle as ao
    g = _load(golden_dir, "anymal_jit.npz")
    n, d = g["root_states"].shape[0], 12
    c = ao.AnymalConsts()
    dof_pos = g["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = g["dof_state"].view(n, d, 2)[..., 1]
    obs = ao.anymal_observations(g["root_states"], g["commands"], g["torques"], g["contact_forces"], dof_pos,
                                 g["default_dof_pos"].repeat(n, 1), dof_vel, torch.tensor([0.0, 0.0, -1.0]).repeat(n, 1),
                                 1.0, 0.05)
    torch.testing.assert_close(obs, g["obs"], rtol=RTOL, atol=ATOL)
    rew, reset = ao.anymal_reward(g["root_states"], g["commands"], g["torques"], g["contact_forces"],
                                  torch.tensor([2, 5, 8, 11]), g["progress_buf"], c.rew_scales(), 0, c.max_episode_length)
    assert torch.equal(reset.to(torch.int64), g["reset"])
    torch.testing.assert_close(rew, g["rew"], rtol=RTOL, atol=ATOL)
    # rows 0..7: base at 1 N (no reset: strict >), base 1 ulp above (reset), 1 ulp below (no), thigh above (reset),
    # thigh at 1 N (no), thigh above (reset), two foot contacts of 500 N (never a reset)
    assert g["reset"][:8].tolist() == [0, 1, 0, 1, 0, 1, 0, 0]
    assert g["reset"][8:12].tolist()[2:] == [1, 1] and c.max_episode_length == 2500


def test_ma_op3_golden(golden_dir):
    """MA_OP3 against MA_OP3.py's own JIT functions; threshold rows and the batch-global hold term."""
    from oracle import ma_op3_oracle as mo
    g = _load(golden_dir, "ma_op3_jit.npz")
    st = {k: v for k, v in g.items() if not k.startswith("out_") and k != "grip_counts"}
    o = mo.oracle_from_state(st)
    assert torch.equal(mo.grip_counts(o.contact_forces, o.gripper_indices), g["grip_counts"])
    assert bool(o._reward_gripper_contact_hold(0)) and not bool(o._reward_gripper_contact_hold(1))
    out = mo.run_jit_level(o)
    for k, v in out.items():
        w = g["out_" + k]
        if v.is_floating_point():
            assert torch.allclose(v, w, rtol=RTOL, atol=ATOL), k
        else:
            assert torch.equal(v, w), k
