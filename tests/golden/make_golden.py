"""Generate tests/golden/*.npz from the REFERENCE ITSELF (run in the build container only).

    python tests/golden/make_golden.py

* ``humanoid_jit.npz`` / ``gogoro_jit.npz`` / ``ant_jit.npz`` / ``gogoro_paper_jit.npz`` / ``gogoro_turning_jit.npz``: inputs and outputs of the reference's own
  ``@torch.jit.script`` functions, lifted by source from /root/reference
  (oracle/lift_reference.py), on seeded synthetic states + hand-built edge rows.
* ``poselib_quat.npz``: outputs of the reference's in-tree quaternion library
  ``isaacgymenvs/tasks/amp/poselib/poselib/core/rotation3d.py`` (quat_mul :36-49,
  quat_conjugate :81-86, quat_rotate :229-235) -- the only in-tree cross-check for the
  restated ``isaacgym.torch_utils`` helpers.

The fixtures store inputs AND outputs, so the tests on the GPU box need neither
/root/reference nor this script.
"""
import importlib.util
import math
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import lift_reference as L                      # noqa: E402
from thormang_isaacgym_b200 import synthetic as syn         # noqa: E402

torch.set_num_threads(1)


def _np(d):
    return {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in d.items()}


def _euler_q(roll, pitch, yaw):
    return syn._euler_to_quat(torch.tensor(roll), torch.tensor(pitch), torch.tensor(yaw))


def humanoid_golden(n_rand=96):
    ref = L.humanoid()
    asset = syn.humanoid_asset(seed=7)
    st = syn.humanoid_state(n_rand, seed=4242, asset=asset)
    root, dof_state = st["root_states"], st["dof_state"]
    d = 21
    # ---- hand-built edge rows appended after the random ones
    edge = []
    base = root[0].clone()
    for z in (0.8, float(np.nextafter(np.float32(0.8), np.float32(0))), float(np.nextafter(np.float32(0.8), np.float32(1)))):
        r = base.clone(); r[2] = z; edge.append(r)                     # height at / one ulp around termination
    for pitch in (math.pi / 2, -math.pi / 2):                         # gimbal branch |sinp| >= 1
        r = base.clone(); r[3:7] = _euler_q(0.3, pitch, -1.0); edge.append(r)
    r = base.clone(); r[0] = 1000.0; r[1] = 0.0; edge.append(r)       # to_target == 0 -> normalize eps path
    r = base.clone(); r[3:7] = torch.tensor([0.0, 0.0, 0.0, 1.0]); edge.append(r)    # identity: roll=yaw=0
    r = base.clone(); r[3:7] = torch.tensor([0.0, 0.0, 1.0, 0.0]); edge.append(r)    # yaw = pi
    r = base.clone(); r[3:7] = _euler_q(0.0, 0.0, -1e-4); edge.append(r)              # yaw just below 0 -> ~2pi after %
    r = base.clone(); r[3:7] = _euler_q(math.pi - 1e-4, 0.0, math.pi - 1e-4); edge.append(r)
    edge = torch.stack(edge)
    ne = edge.shape[0]
    n = n_rand + ne
    root = torch.cat([root, edge])
    ext = syn.humanoid_state(ne, seed=99, asset=asset)
    dof_state = torch.cat([dof_state, ext["dof_state"]])
    dof_force = torch.cat([st["dof_force"], ext["dof_force"]])
    sensors = torch.cat([st["sensors"], ext["sensors"]])
    actions = torch.cat([st["actions"], ext["actions"]])
    potentials = torch.cat([st["potentials"], ext["potentials"]])
    progress = torch.cat([st["progress_buf"], torch.tensor([996, 997, 998, 999, 1000, 0, 1, 500, 998, 999][:ne])])
    reset = torch.cat([st["reset_buf"], torch.tensor([0, 1, 0, 1, 0, 1, 0, 0, 0, 0][:ne])])
    # joints exactly at the limits and at the 0.98 knee (unscale end points)
    dof_pos = dof_state.view(n, d, 2)[..., 0]
    lo, hi = asset["dof_limits_lower"], asset["dof_limits_upper"]
    dof_pos[n_rand + 0] = lo
    dof_pos[n_rand + 1] = hi
    dof_pos[n_rand + 2] = 0.5 * (hi + lo) + 0.49 * (hi - lo)

    dt, dvs, cfs, avs = 0.0166, 0.1, 0.01, 0.25
    targets = torch.tensor([1000.0, 0.0, 0.0]).repeat(n, 1)
    inv_start_rot = torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(n, 1)
    b0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    b1 = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    dof_vel = dof_state.view(n, d, 2)[..., 1]
    obs, pot, prev_pot, up_vec, heading_vec = ref.compute_humanoid_observations(
        torch.zeros(n, 110), root, targets, potentials, inv_start_rot, dof_pos, dof_vel, dof_force,
        lo, hi, dvs, sensors, actions, dt, cfs, avs, b0, b1)
    max_eff = float(asset["motor_efforts"].max())
    rew, reset_out = ref.compute_humanoid_reward(
        obs, reset, progress, actions, 0.1, 0.5, pot, prev_pot, 0.01, 0.05, 0.25, max_eff,
        asset["motor_efforts"], 0.8, -1.0, 1000.0)
    wm = ref.compute_reward(obs.view(2, n // 2, 110), actions.view(2, n // 2, d), 0.1, 0.5, 0.01, 0.05, 0.25,
                            max_eff, asset["motor_efforts"], 0.8, -1.0)
    # a non-identity start rotation as well (inv_start_rot path of compute_heading_and_up)
    inv2 = torch.tensor([0.1, -0.2, 0.3, 0.9]); inv2 = (inv2 / inv2.norm()).repeat(n, 1)
    obs2, pot2, prev2, up2, head2 = ref.compute_humanoid_observations(
        torch.zeros(n, 110), root, targets, potentials, inv2, dof_pos, dof_vel, dof_force,
        lo, hi, dvs, sensors, actions, dt, cfs, avs, b0, b1)
    np.savez_compressed(os.path.join(HERE, "humanoid_jit.npz"), **_np(dict(
        root_states=root, dof_state=dof_state, dof_force=dof_force, sensors=sensors, actions=actions,
        potentials=potentials, progress_buf=progress, reset_buf=reset,
        dof_limits_lower=lo, dof_limits_upper=hi, motor_efforts=asset["motor_efforts"],
        obs=obs, potentials_out=pot, prev_potentials_out=prev_pot, up_vec=up_vec, heading_vec=heading_vec,
        rew=rew, reset_out=reset_out, wm_reward=wm,
        inv_start_rot2=inv2, obs2=obs2, up_vec2=up2, heading_vec2=head2,
        n_rand=n_rand)))
    print("humanoid_jit.npz", n, "rows; reward values", sorted(set(rew.tolist())))


def gogoro_golden(n_rand=96):
    ref = L.gogoro()
    st = syn.gogoro_state(n_rand, seed=4243)
    root = st["root_tensor"]
    base = root[0].clone()
    edge = []
    for roll in (0.2999, 0.3001, -0.3001, 0.0):
        r = base.clone(); r[3:7] = _euler_q(roll, 0.0, 0.5); edge.append(r)          # around max_tilt
    for yaw in (math.pi - 1e-4, -math.pi + 1e-4, 0.0, -1e-4):
        r = base.clone(); r[3:7] = _euler_q(0.05, 0.01, yaw); edge.append(r)         # yaw wrap
    for pitch in (math.pi / 2, -math.pi / 2):
        r = base.clone(); r[3:7] = _euler_q(0.1, pitch, 0.3); edge.append(r)         # gimbal
    edge = torch.stack(edge)
    ne = edge.shape[0]
    n = n_rand + ne
    root = torch.cat([root, edge])
    desired = torch.cat([st["yaw_command"], torch.tensor([0.5, 0.5, 0.5, 0.5, -math.pi + 1e-4, math.pi - 1e-4,
                                                           math.pi, 3.0, 0.3, 0.3])[:ne]])
    last_cmd = torch.cat([st["curent_command"], torch.zeros(ne)])
    hist = torch.cat([st["action_history"], syn._u(syn._gen(5), (ne, 5), -1.5, 1.5)])
    progress = torch.cat([st["progress_buf"], torch.tensor([997, 998, 999, 1000, 0, 299, 300, 301, 998, 999])[:ne]])
    obs = ref.compute_gogoro_observations(root.clone(), desired.clone(), last_cmd.clone())
    rew, reset = ref.compute_gogoro_reward(curent_obs=obs.view(n, 1, 6), progress_buf=progress,
                                           action_history=hist, max_episode_length=torch.tensor(1000))
    y1 = syn._u(syn._gen(6), (256,), -7.0, 7.0)
    y2 = syn._u(syn._gen(7), (256,), -7.0, 7.0)
    sad = ref.shortest_angle_distance(y1, y2)
    np.savez_compressed(os.path.join(HERE, "gogoro_jit.npz"), **_np(dict(
        root_tensor=root, desired_yaw=desired, last_command=last_cmd, action_history=hist, progress_buf=progress,
        obs=obs, rew=rew, reset=reset, sad_y1=y1, sad_y2=y2, sad=sad, n_rand=n_rand)))
    print("gogoro_jit.npz", n, "rows; resets", int(reset.sum()))


def ant_golden(n_rand=96):
    ref = L.ant()
    asset = syn.ant_asset()
    st = syn.ant_state(n_rand, seed=4244, asset=asset)
    root = st["root_states"]
    base = root[0].clone()
    edge = []
    for z in (0.31, float(np.nextafter(np.float32(0.31), np.float32(0))), float(np.nextafter(np.float32(0.31), np.float32(1)))):
        r = base.clone(); r[2] = z; edge.append(r)
    for pitch in (math.pi / 2, -math.pi / 2):
        r = base.clone(); r[3:7] = _euler_q(0.3, pitch, -1.0); edge.append(r)
    r = base.clone(); r[0] = 1000.0; r[1] = 0.0; edge.append(r)
    r = base.clone(); r[3:7] = torch.tensor([0.0, 0.0, 0.0, 1.0]); edge.append(r)
    r = base.clone(); r[3:7] = _euler_q(0.0, 0.0, -1e-4); edge.append(r)
    edge = torch.stack(edge)
    ne = edge.shape[0]
    n = n_rand + ne
    ext = syn.ant_state(ne, seed=98, asset=asset)
    root = torch.cat([root, edge])
    dof_state = torch.cat([st["dof_state"], ext["dof_state"]])
    sensors = torch.cat([st["sensors"], ext["sensors"]])
    actions = torch.cat([st["actions"], ext["actions"]])
    potentials = torch.cat([st["potentials"], ext["potentials"]])
    progress = torch.cat([st["progress_buf"], torch.tensor([996, 997, 998, 999, 1000, 0, 1, 500][:ne])])
    reset = torch.cat([st["reset_buf"], torch.tensor([0, 1, 0, 1, 0, 1, 0, 0][:ne])])
    d = 8
    dof_pos = dof_state.view(n, d, 2)[..., 0]
    dof_vel = dof_state.view(n, d, 2)[..., 1]
    lo, hi = asset["dof_limits_lower"], asset["dof_limits_upper"]
    dof_pos[n_rand + 0] = hi                       # scaled position exactly 1 (> 0.99 branch)
    dof_pos[n_rand + 1] = 0.5 * (hi + lo) + 0.495 * (hi - lo)
    targets = torch.tensor([1000.0, 0.0, 0.0]).repeat(n, 1)
    inv = torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(n, 1)
    b0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    b1 = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    obs, pot, prev_pot, up_vec, heading_vec = ref.compute_ant_observations(
        torch.zeros(n, 60), root, targets, potentials, inv, dof_pos, dof_vel, lo, hi, 0.2, sensors, actions, 0.0166,
        0.1, b0, b1, 2)
    rew, reset_out = ref.compute_ant_reward(obs, reset, progress, actions, 0.1, 0.5, pot, prev_pot, 0.005, 0.05, 0.1,
                                            0.31, -2.0, 1000.0)
    np.savez_compressed(os.path.join(HERE, "ant_jit.npz"), **_np(dict(
        root_states=root, dof_state=dof_state, sensors=sensors, actions=actions, potentials=potentials,
        progress_buf=progress, reset_buf=reset, dof_limits_lower=lo, dof_limits_upper=hi,
        joint_gears=asset["joint_gears"], obs=obs, potentials_out=pot, prev_potentials_out=prev_pot, up_vec=up_vec,
        heading_vec=heading_vec, rew=rew, reset_out=reset_out, n_rand=n_rand)))
    print("ant_jit.npz", n, "rows; resets", int(reset_out.sum()))


def gogoro_paper_golden(n_rand=96):
    """gogoro_realistic_turning_sim_paper.py JIT functions: observations [N,8], reward over [N,20,8]."""
    ref = L.gogoro_paper()
    st = syn.gogoro_paper_state(n_rand, seed=4245)
    root = st["root_tensor"]
    base = root[0].clone()
    edge = []
    for roll in (0.3799, 0.3801, -0.3801, 0.0):
        r = base.clone(); r[3:7] = _euler_q(roll, 0.0, 0.5); edge.append(r)          # around max_tilt = 0.38
    for yaw in (math.pi - 1e-4, -math.pi + 1e-4, 0.0, -1e-4):
        r = base.clone(); r[3:7] = _euler_q(0.05, 0.01, yaw); edge.append(r)         # yaw wrap
    for pitch in (math.pi / 2, -math.pi / 2):
        r = base.clone(); r[3:7] = _euler_q(0.1, pitch, 0.3); edge.append(r)         # gimbal
    edge = torch.stack(edge)
    ne = edge.shape[0]
    n = n_rand + ne
    root = torch.cat([root, edge])
    desired = torch.cat([st["yaw_command"], torch.tensor([0.5, 0.5, 0.5, 0.5, -math.pi + 1e-4, math.pi - 1e-4,
                                                           math.pi, 3.0, 0.3, 0.3])[:ne]])
    cmd = torch.cat([st["curent_command"], torch.tensor([0.5, -0.5, 0.0, 0.25, 0.1, 0.2, 0.3, 0.4, -0.1, -0.2])[:ne]])
    delay = torch.cat([st["steer_delay"], torch.tensor([0, 1, 2, 3, 4, 0, 1, 2, 3, 4])[:ne]])
    progress = torch.cat([st["progress_buf"], torch.tensor([3597, 3598, 3599, 3600, 0, 299, 300, 301, 3598, 3599])[:ne]])
    obs = ref.compute_gogoro_observations(root.clone(), desired.clone(), cmd.clone(), (delay.clone() - 0) / (5 - 0))
    ext = syn.gogoro_paper_state(ne, seed=97)
    frames = torch.cat([st["buffer_obs"], ext["buffer_obs"]])
    frames = torch.cat([frames[:, 1:], obs.unsqueeze(1)], dim=1)                     # the push of :502-503
    rew, reset = ref.compute_gogoro_reward(act_buff=frames[:, :, 6], curent_obs=frames, progress_buf=progress,
                                           max_episode_length=torch.tensor(3600), yaw_update_freq=torch.tensor(300))
    np.savez_compressed(os.path.join(HERE, "gogoro_paper_jit.npz"), **_np(dict(
        root_tensor=root, desired_yaw=desired, curent_command=cmd, steer_delay=delay, progress_buf=progress,
        frames=frames, obs=obs, rew=rew, reset=reset, n_rand=n_rand)))
    print("gogoro_paper_jit.npz", n, "rows; resets", int(reset.sum()), "reward range", float(rew.min()), float(rew.max()))


def gogoro_turning_golden(n_rand=96):
    """gogoro_realistic_turning.py JIT functions (SURVEY 8(f) N1 sibling): same observation function as the paper file,
    the rational reward with the all-env ``mean(act_buff)`` term over [N,20,8]."""
    ref = L.gogoro_turning()
    st = syn.gogoro_paper_state(n_rand, seed=5245)
    root = st["root_tensor"]
    base = root[0].clone()
    edge = []
    for roll in (0.3799, 0.3801, -0.3801, 0.0):
        r = base.clone(); r[3:7] = _euler_q(roll, 0.0, -0.7); edge.append(r)         # around max_tilt = 0.38
    for yaw in (math.pi - 1e-4, -math.pi + 1e-4, 0.0, -1e-4):
        r = base.clone(); r[3:7] = _euler_q(-0.05, 0.02, yaw); edge.append(r)        # yaw wrap
    for wx in (0.999, 1.001, -1.5):                                                  # dtilt clamp at +-1 (:728-729)
        r = base.clone(); r[3:7] = _euler_q(0.0, 0.0, 0.0); r[10:13] = torch.tensor([wx, 0.0, 0.1]); edge.append(r)
    edge = torch.stack(edge)
    ne = edge.shape[0]
    n = n_rand + ne
    root = torch.cat([root, edge])
    desired = torch.cat([st["yaw_command"], torch.tensor([0.5, 0.5, 0.5, 0.5, -math.pi + 1e-4, math.pi - 1e-4,
                                                           math.pi, 3.0, 0.3, 0.3, 0.0])[:ne]])
    cmd = torch.cat([st["curent_command"], torch.tensor([0.5, -0.5, 0.0, 0.25, 0.1, 0.2, 0.3, 0.4, -0.1, -0.2, 0.0])[:ne]])
    delay = torch.cat([st["steer_delay"], torch.tensor([0, 1, 2, 3, 4, 0, 1, 2, 3, 4, 0])[:ne]])
    progress = torch.cat([st["progress_buf"], torch.tensor([3597, 3598, 3599, 3600, 0, 299, 300, 301, 3598, 3599, 7])[:ne]])
    obs = ref.compute_gogoro_observations(root.clone(), desired.clone(), cmd.clone(), (delay.clone() - 0) / (5 - 0))
    ext = syn.gogoro_paper_state(ne, seed=98)
    frames = torch.cat([st["buffer_obs"], ext["buffer_obs"]])
    frames[:, :, 6] += 0.07                                                          # a command mean away from zero
    frames = torch.cat([frames[:, 1:], obs.unsqueeze(1)], dim=1)                     # the push of :495-496
    rew, reset = ref.compute_gogoro_reward(act_buff=frames[:, :, 6], curent_obs=frames, progress_buf=progress,
                                           max_episode_length=torch.tensor(3600), yaw_update_freq=torch.tensor(300))
    np.savez_compressed(os.path.join(HERE, "gogoro_turning_jit.npz"), **_np(dict(
        root_tensor=root, desired_yaw=desired, curent_command=cmd, steer_delay=delay, progress_buf=progress,
        frames=frames, obs=obs, rew=rew, reset=reset, n_rand=n_rand)))
    print("gogoro_turning_jit.npz", n, "rows; resets", int(reset.sum()), "reward range", float(rew.min()), float(rew.max()))


def anymal_golden(n_rand=96):
    """anymal.py JIT functions: observations [N,61], reward / reset incl. contact norms around 1 N."""
    from oracle import anymal_oracle as ao
    ref = L.anymal()
    st = syn.anymal_state(n_rand, seed=4246, contact_frac=0.3)
    cf = st["contact_forces"]
    one = float(np.float32(1.0))
    up, dn = float(np.nextafter(np.float32(1.0), np.float32(2))), float(np.nextafter(np.float32(1.0), np.float32(0)))
    for i, (body, val) in enumerate([(0, one), (0, up), (0, dn), (2, up), (5, one), (11, up), (3, 500.0), (12, 500.0)]):
        cf[i] = 0.0
        cf[i, body, 2] = val                                  # exactly at / one ulp around the threshold; feet do not count
    progress = st["progress_buf"].clone()
    progress[8:12] = torch.tensor([2497, 2498, 2499, 2500])
    c = ao.AnymalConsts()
    n = n_rand
    d = 12
    dof_pos = st["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = st["dof_state"].view(n, d, 2)[..., 1]
    default = torch.tensor([0.03, 0.4, -0.8, 0.03, -0.4, 0.8, -0.03, 0.4, -0.8, -0.03, -0.4, 0.8]).repeat(n, 1)
    grav = torch.tensor([0.0, 0.0, -1.0]).repeat(n, 1)
    obs = ref.compute_anymal_observations_V2(st["root_states"], st["commands"], st["torques"], cf, dof_pos, default, dof_vel,
                                             grav, st["actions"], 2.0, 0.25, 1.0, 0.05)
    knee = torch.tensor([2, 5, 8, 11])
    rew, reset = ref.compute_anymal_reward(st["root_states"], st["commands"], st["torques"], cf, knee, progress,
                                           c.rew_scales(), 0, c.max_episode_length)
    np.savez_compressed(os.path.join(HERE, "anymal_jit.npz"), **_np(dict(
        root_states=st["root_states"], dof_state=st["dof_state"], torques=st["torques"], contact_forces=cf,
        commands=st["commands"], actions=st["actions"], progress_buf=progress, default_dof_pos=default[0],
        obs=obs, rew=rew, reset=reset.to(torch.int64))))
    print("anymal_jit.npz", n, "rows; resets", int(reset.sum()), "reward max", float(rew.max()))


def ma_op3_golden(n=96):
    """MA_OP3.py JIT functions run through the method bodies of compute_observations / compute_reward:
    rows forced onto the 0.90 / 0.95 / 0.98 up-projection, 0.40 m table distance, 0.25 m table height and
    episode-length thresholds; exactly one env holds each gripper (the batch-global hold term is 1)."""
    from oracle import ma_op3_oracle as mo
    ref = L.ma_op3()
    st = syn.ma_op3_state(n, seed=5151)
    root = st["root_states"].view(n, 3, 13)
    root[0, 2, 2] = 0.25                                     # table height exactly at the threshold (not < 0.25)
    root[1, 2, 2] = float(np.nextafter(np.float32(0.25), np.float32(0)))
    root[2, 0, 0:3] = root[2, 2, 0:3] + torch.tensor([0.4, 0.0, 0.0])       # agent 0 at ~0.40 m from the table
    root[3, 1, 0:3] = root[3, 2, 0:3] + torch.tensor([0.0, 0.41, 0.0])
    st["progress_buf"][4:10] = torch.tensor([0, 1, 3011, 3012, 3013, 0])
    st["reset_buf"][9] = 1                                   # cleared by the episode_lengths == 0 line (:656)
    feet, grip = syn.ma_op3_body_indices()
    cf = st["contact_forces"]
    cf[:, grip.flatten(), 0] = 0.0
    cf[11, grip[0, 0], 0] = 0.5; cf[12, grip[0, 1], 0] = 0.7          # agent 0: exactly one env per gripper -> hold = 1
    cf[13, grip[1, 0], 0] = 0.5; cf[14, grip[1, 0], 0] = 0.5; cf[15, grip[1, 1], 0] = 0.9   # agent 1: two left holds -> 0
    o = mo.oracle_from_state(st)
    out = mo.run_jit_level(o, ref.compute_op3_observations, ref.compute_op3_reward, ref.compute_objective_reward)
    np.savez_compressed(os.path.join(HERE, "ma_op3_jit.npz"), **_np({**{k: v for k, v in st.items()},
                                                                      **{"out_" + k: v for k, v in out.items()},
                                                                      "grip_counts": mo.grip_counts(cf, grip)}))
    print("ma_op3_jit.npz", n, "rows; resets", int(out["reset"].sum()), "reward max", float(out["rew"].max()),
          "grip counts", mo.grip_counts(cf, grip).tolist())


def poselib_golden(n=256):
    path = os.path.join(L.REFERENCE_ROOT, "isaacgymenvs/tasks/amp/poselib/poselib/core/rotation3d.py")
    spec = importlib.util.spec_from_file_location("_ref_rotation3d", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["_ref_rotation3d"] = mod
    spec.loader.exec_module(mod)
    g = syn._gen(11)
    a = torch.randn((n, 4), generator=g, dtype=torch.float64); a = a / a.norm(dim=-1, keepdim=True)
    b = torch.randn((n, 4), generator=g, dtype=torch.float64); b = b / b.norm(dim=-1, keepdim=True)
    v = torch.randn((n, 3), generator=g, dtype=torch.float64)
    np.savez_compressed(os.path.join(HERE, "poselib_quat.npz"), **_np(dict(
        a=a, b=b, v=v, quat_mul=mod.quat_mul(a, b), quat_conjugate=mod.quat_conjugate(a),
        quat_rotate=mod.quat_rotate(a, v), quat_rotate_conj=mod.quat_rotate(mod.quat_conjugate(a), v))))
    print("poselib_quat.npz", n, "rows (fp64)")


if __name__ == "__main__":
    if not L.reference_available():
        sys.exit("reference tree not found: golden vectors can only be regenerated in the build container")
    humanoid_golden()
    gogoro_golden()
    ant_golden()
    gogoro_paper_golden()
    gogoro_turning_golden()
    anymal_golden()
    ma_op3_golden()
    poselib_golden()
