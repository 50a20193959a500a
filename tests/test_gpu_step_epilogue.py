"""GPU parity of the fused ``VecTask.step`` prologue / epilogue against
oracle/vec_task_oracle.py (vec_task.py:323-327, 347-353, 576-646).

Only multiplications, additions and min/max are involved and the draws are shared, so the bar is bit-exact.
"""
import ctypes as C
import math

import pytest
import torch

import parity_util as U
from oracle import vec_task_oracle as vo
from thormang_isaacgym_b200 import _native as nat
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

GAUSS = {"range": [0.01, 0.002], "range_correlated": [0.005, 0.001], "operation": "additive", "distribution": "gaussian",
         "schedule": "linear", "schedule_steps": 40000}
UNI_SCALE = {"range": [0.9, 1.1], "range_correlated": [0.95, 1.05], "operation": "scaling", "distribution": "uniform",
             "schedule": "constant", "schedule_steps": 100}
UNI_ADD = {"range": [-0.02, 0.03], "operation": "additive", "distribution": "uniform"}
GAUSS_SCALE = {"range": [1.0, 0.05], "range_correlated": [1.0, 0.02], "operation": "scaling", "distribution": "gaussian",
               "schedule": "linear", "schedule_steps": 3000}


def _lambda_of(prm, last_step):
    sp = vo.scheduled_params(prm, last_step)
    lam = nat.NoiseLambda()
    lam.distribution = 1 if prm["distribution"] == "gaussian" else 2
    lam.operation = 0 if prm["operation"] == "additive" else 1
    if "mu" in sp:
        lam.range[0], lam.range[1], lam.range_correlated[0], lam.range_correlated[1] = sp["mu"], sp["var"], sp["mu_corr"], sp["var_corr"]
    else:
        lam.range[0], lam.range[1], lam.range_correlated[0], lam.range_correlated[1] = sp["lo"], sp["hi"], sp["lo_corr"], sp["hi_corr"]
    return lam


@pytest.mark.parametrize("prm,last_step", [(GAUSS, 12345), (GAUSS, 50000), (UNI_SCALE, 50), (UNI_SCALE, 100), (UNI_ADD, 0),
                                           (GAUSS_SCALE, 1000)])
@pytest.mark.parametrize("shape,clip", [((1000, 110), math.inf), ((33, 61), 5.0), ((7, 1), 0.5), ((4096, 160), 1.0)])
def test_noise_clamp_kernel(prm, last_step, shape, clip):
    g = torch.Generator().manual_seed(hash((last_step, shape)) % 1000)
    x = torch.randn(shape, generator=g) * 3
    corr = torch.randn(shape, generator=g)
    unit = torch.randn(shape, generator=g) if prm["distribution"] == "gaussian" else torch.rand(shape, generator=g)
    want_noisy, want_clamped = vo.step_epilogue(x, clip, prm, last_step, corr, unit)
    xd, cd, ud = x.to(DEV), corr.to(DEV), unit.to(DEV)
    noisy, clamped = torch.empty_like(xd), torch.empty_like(xd)
    lam = _lambda_of(prm, last_step)
    nat.check(nat.lib().tk_step_noise_clamp(nat.ptr(xd), xd.numel(), C.byref(lam), nat.ptr(cd), nat.ptr(ud), 0, 0, 0, clip,
                                            nat.ptr(noisy), nat.ptr(clamped), nat.current_stream_ptr(DEV)))
    torch.cuda.synchronize()
    U.exact(noisy, want_noisy, "noisy tensor")
    U.exact(clamped, want_clamped, "clamped tensor")


def test_unaligned_in_place_and_clamp_only():
    n = 1001
    base = torch.randn(n + 1, device=DEV) * 4
    x = base[1:]                                     # 4-byte aligned only
    want = torch.clamp(x, -1.5, 1.5)
    nat.check(nat.lib().tk_step_noise_clamp(nat.ptr(x), n, None, 0, 0, 0, 0, 0, 1.5, 0, nat.ptr(x), nat.current_stream_ptr(DEV)))
    torch.cuda.synchronize()
    U.exact(x, want, "in-place clamp, unaligned")
    assert nat.lib().tk_step_noise_clamp(nat.ptr(x), n, None, 0, 0, 0, 0, 0, 1.0, 0, 0, 0) == -1      # no destination


def test_philox_units_and_sharding():
    n = 4096
    x = torch.randn((n, 60), device=DEV)
    lam = _lambda_of(GAUSS, 40000)
    lib, st = nat.lib(), nat.current_stream_ptr(DEV)
    unit = torch.empty_like(x)
    nat.check(lib.tk_step_noise_units(x.numel(), 1, 11, 5, 0, nat.ptr(unit), st))
    assert abs(float(unit.mean())) < 0.01 and abs(float(unit.std()) - 1.0) < 0.01
    full = torch.empty_like(x)
    nat.check(lib.tk_step_noise_clamp(nat.ptr(x), x.numel(), C.byref(lam), 0, 0, 11, 5, 0, math.inf, nat.ptr(full), 0, st))
    want, _ = vo.step_epilogue(x.cpu(), math.inf, GAUSS, 40000, torch.zeros(n, 60), unit.cpu())
    U.exact(full, want, "philox noise = materialised units")
    # two range shards with the global element offset give the same tensor
    a, b = x[:1000].contiguous(), x[1000:].contiguous()
    oa, ob = torch.empty_like(a), torch.empty_like(b)
    nat.check(lib.tk_step_noise_clamp(nat.ptr(a), a.numel(), C.byref(lam), 0, 0, 11, 5, 0, math.inf, nat.ptr(oa), 0, st))
    nat.check(lib.tk_step_noise_clamp(nat.ptr(b), b.numel(), C.byref(lam), 0, 0, 11, 5, 1000 * 60, math.inf, nat.ptr(ob), 0, st))
    torch.cuda.synchronize()
    U.exact(torch.cat([oa, ob]), full, "sharded philox noise")
    other = torch.empty_like(x)
    nat.check(lib.tk_step_noise_clamp(nat.ptr(x), x.numel(), C.byref(lam), 0, 0, 11, 6, 0, math.inf, nat.ptr(other), 0, st))
    assert not torch.equal(other, full)


def test_apply_randomizations_matches_reference_schedule():
    env, _, _ = U.make_humanoid_pair(64, seed=2)
    dr = {"frequency": 600, "observations": dict(GAUSS), "actions": dict(UNI_SCALE)}
    env.sim.frame_count = 12345
    env.apply_randomizations(dr)
    for name, prm in (("observations", GAUSS), ("actions", UNI_SCALE)):
        want = vo.scheduled_params(prm, 12345)
        got = {k: v for k, v in env.dr_randomizations[name].items() if k in want}
        assert got == want
        assert callable(env.dr_randomizations[name]["noise_lambda"])
    # not re-randomised before `frequency` frames have passed (vec_task.py:558)
    first = env.dr_randomizations["observations"]
    env.sim.frame_count = 12345 + 599
    env.apply_randomizations(dr)
    assert env.dr_randomizations["observations"] is first
    env.sim.frame_count = 12345 + 600
    env.apply_randomizations(dr)
    assert env.dr_randomizations["observations"] is not first
    # the lambda itself: tensor + (corr * var_corr + mu_corr) + z * var + mu with z from the materialised Philox draws
    x = torch.randn((64, 110), device=DEV)
    y = env.dr_randomizations["observations"]["noise_lambda"](x)
    torch.cuda.synchronize()
    p = env.dr_randomizations["observations"]
    resid = (y - x - (p["corr"] * p["var_corr"] + p["mu_corr"]) - p["mu"]).cpu()
    assert abs(float(resid.std()) - p["var"]) < 0.1 * p["var"] and abs(float(resid.mean())) < 0.2 * p["var"]


def test_step_fills_rollout_buffer_from_the_kernel():
    """Humanoid: clip inf, no obs noise -> the fused post-physics kernel stores the obs tile into the rollout slot itself."""
    n, horizon = 1000, 4
    env, orc, st = U.make_humanoid_pair(n, seed=5)
    buf = torch.full((horizon, n, 110), float("nan"), device=DEV)
    env.set_rollout_buffer(buf)
    calls0 = getattr(env, "_epilogue_calls", 0)
    for t in range(horizon + 1):
        obs_d, rew, reset, extras = env.step(torch.rand((n, 21), device=DEV) * 2 - 1)
        torch.cuda.synchronize()
        slot = t % horizon
        U.exact(buf[slot], env.obs_buf, f"rollout slot {slot}")
        assert obs_d["obs"].data_ptr() == buf[slot].data_ptr()
    assert getattr(env, "_epilogue_calls", 0) == calls0          # no extra kernel ran
    assert bool(torch.isfinite(buf).all())
    env.set_rollout_buffer(None)
    obs_d, *_ = env.step(torch.zeros((n, 21), device=DEV))
    assert obs_d["obs"].data_ptr() == env.obs_buf.data_ptr()


def test_step_with_noise_and_clamp_anymal():
    """Anymal (clipObservations 5, clipActions 1) with observation AND action noise: one fused pass each."""
    n, horizon = 512, 3
    env, orc, st = U.make_anymal_pair(n, seed=6)
    env.sim.frame_count = 77
    env.apply_randomizations({"frequency": 1000, "observations": dict(GAUSS_SCALE), "actions": dict(UNI_ADD)})
    buf = torch.zeros((horizon, n, 61), device=DEV)
    env.set_rollout_buffer(buf)
    a = torch.rand((n, 12), device=DEV) * 3 - 1.5
    call = getattr(env, "_epilogue_calls", 0)
    obs_d, rew, reset, extras = env.step(a)
    torch.cuda.synchronize()
    seed = 1234 & 0xFFFFFFFFFFFFFFFF
    # actions: noise (Philox draws of call #call+1, stream bit 1) then clamp, then the PD target write
    ua = torch.empty_like(a)
    nat.check(nat.lib().tk_step_noise_units(a.numel(), 2, env.cfg.get("seed", 0), ((call + 1) << 1) | 1, 0, nat.ptr(ua), nat.current_stream_ptr(DEV)))
    torch.cuda.synchronize()
    want_a = vo.step_prologue(a.cpu(), 1.0, UNI_ADD, 77, env.dr_randomizations["actions"]["corr"].cpu(), ua.cpu())
    U.exact(env.actions, want_a, "noisy clamped actions")
    # observations: the kernel's clean obs -> noise in place -> clamped copy in the rollout slot
    uo = torch.empty((n, 61), device=DEV)
    nat.check(nat.lib().tk_step_noise_units(uo.numel(), 1, env.cfg.get("seed", 0), ((call + 2) << 1), 0, nat.ptr(uo), nat.current_stream_ptr(DEV)))
    clean = env.obs_buf.clone()
    env.compute_observations()                                   # recompute the clean observations of the same state
    torch.cuda.synchronize()
    want_noisy, want_clamped = vo.step_epilogue(env.obs_buf.cpu(), 5.0, GAUSS_SCALE, 77,
                                                env.dr_randomizations["observations"]["corr"].cpu(), uo.cpu())
    U.exact(clean, want_noisy, "obs_buf holds the noisy observations (vec_task.py:349)")
    U.exact(buf[0], want_clamped, "rollout slot holds clamp(noisy obs)")
    assert float(buf[0].abs().max()) <= 5.0 and obs_d["obs"].data_ptr() == buf[0].data_ptr()


def test_reset_done_after_kernel_filled_slot():
    """A step that let the fused kernel fill the rollout slot must not leak into reset()/reset_done(): those return a
    clamp of obs_buf (vec_task.py:386,400) and never touch or advance the rollout buffer."""
    n, horizon = 256, 4
    env, orc, st = U.make_humanoid_pair(n, seed=9)
    buf = torch.full((horizon, n, 110), float("nan"), device=DEV)
    env.set_rollout_buffer(buf)
    env.step(torch.rand((n, 21), device=DEV) * 2 - 1)
    torch.cuda.synchronize()
    assert env._rollout_t == 1
    obs_d, done = env.reset_done()
    torch.cuda.synchronize()
    U.exact(obs_d["obs"], env.obs_buf, "reset_done returns obs_buf")
    assert env._rollout_t == 1 and bool(torch.isnan(buf[1]).all())          # slot 1 untouched, index not advanced
    obs_d = env.reset()
    U.exact(obs_d["obs"], env.obs_buf, "reset returns obs_buf")
    assert env._rollout_t == 1
    obs_d, *_ = env.step(torch.rand((n, 21), device=DEV) * 2 - 1)
    torch.cuda.synchronize()
    U.exact(buf[1], env.obs_buf, "the next step fills slot 1")
    assert obs_d["obs"].data_ptr() == buf[1].data_ptr()


def test_reset_is_clamp_only_under_observation_noise():
    """reset()/reset_done() only clamp (vec_task.py:386,400): no noise lambda, no Philox call consumed, obs_buf unchanged."""
    n = 300
    env, orc, st = U.make_anymal_pair(n, seed=10)
    env.apply_randomizations({"frequency": 1000, "observations": dict(GAUSS_SCALE)})
    env.obs_buf.copy_(torch.randn((n, 61), device=DEV) * 4)                 # some entries beyond clipObservations = 5
    before = env.obs_buf.clone()
    calls = getattr(env, "_epilogue_calls", 0)
    obs_d = env.reset()
    torch.cuda.synchronize()
    U.exact(env.obs_buf, before, "obs_buf untouched by reset()")
    U.exact(obs_d["obs"], torch.clamp(before, -5.0, 5.0), "reset() = clamp(obs_buf)")
    assert getattr(env, "_epilogue_calls", 0) == calls


def test_fake_sim_counts_frames_for_the_dr_schedule():
    """apply_randomizations reads gym.get_frame_count (vec_task.py:553): FakeSim.simulate() counts sim frames."""
    env, _, _ = U.make_humanoid_pair(64, seed=3)
    f0 = env.sim.frame_count
    for _ in range(3):
        env.step(torch.zeros((64, 21), device=DEV))
    assert env.sim.frame_count == f0 + 3 * env.control_freq_inv
    env.apply_randomizations({"frequency": 2, "observations": dict(GAUSS)})
    assert env.last_step == env.sim.frame_count


def test_non_fused_clip_path_accepts_host_and_double_actions():
    """Gogoro (finite clipActions, no fused clip): actions arriving on the CPU / as float64 are moved and cast before
    the clamp kernel sees a pointer (the reference's torch.clamp works on any device / dtype)."""
    n = 200
    env, orc, st = U.make_gogoro_pair(n, seed=12)
    a = (torch.rand((n, 1), dtype=torch.float64) * 3 - 1.5)                 # host, fp64
    env.step(a)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(env.obs_buf).all())
