"""tk_gogoro_paper_diff_sums_static -- the entry point for hosts that may not allocate device memory.

Who needs it: every other call of the 8x20 Gogoro section goes through a handle, and creating a handle cudaMallocs its
work areas.  A plug-in living inside a simulator or trainer that already owns the whole device heap, or code on a step
loop / host callback where cudaMalloc's device-wide synchronisation is not acceptable, cannot do that.  This entry point
takes the command ring and an output pointer and nothing else; its scratch is an array compiled into the library.

What is checked here, through the C ABI (ctypes) like every other GPU test: after a fused step made with the ordinary
handle-based API, the static entry point, fed the ring that step left behind, returns the same {sum, count} pair as
tk_gogoro_paper_diff_sums -- on two steps in a row (the second call finds the first call's leftovers in the shared
scratch), for both file variants, at a size that fills more than one block and one that is ragged.  Tolerance: rtol 1e-9,
the one the repo uses for fp64 sums whose addends are identical but arrive in a different order
(tests/test_gpu_humanoid.py, shard statistics).

The file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide other tests.
"""
import pytest
import torch

from thormang_isaacgym_b200 import _native as nat

DEV = "cuda:0"
PAPER, TURNING = 0, 1
MAX_ENVS = 1 << 20
ERR_INVALID = -1                                 # TK_ERR_INVALID, include/thormang_b200.h


def _static(lib, ring, n, variant, slot, out, stream):
    return lib.tk_gogoro_paper_diff_sums_static(nat.ptr(ring), n, variant, slot, nat.ptr(out), stream)


@pytest.mark.gpu
@pytest.mark.parametrize("turning", [False, True], ids=["paper", "turning"])
@pytest.mark.parametrize("n", [1000, 4096 + 17])
def test_static_scratch_entry_point_matches_the_handle_one_two_steps_in_a_row(n, turning):
    import parity_util as U
    from thormang_isaacgym_b200 import synthetic as syn
    cols = 8 if turning else 5
    st = syn.gogoro_paper_state(n, seed=91)
    if turning:
        st["buffer_obs"][:, :, 6] += 0.11            # keeps the plain sum of the ring away from zero
    env, _, _ = U.make_gogoro_paper_pair(n, seed=91, state=st, turning=turning)
    lib, stream = env._lib, env._stream()
    got = torch.full((2,), float("nan"), device=DEV, dtype=torch.float64)
    for step in range(2):                            # a fresh handle's device tick is 0: step k writes ring slot k
        nz = syn.gogoro_paper_noise(n, seed=92 + step, reset_cols=cols)
        keep = []
        dz = env._noise(nz["sensor_unit"], nz["reset_unit"], nz["speed_unit"], nz["yaw_unit"], keep)
        nat.check(lib.tk_gogoro_paper_post_physics(env._handle, dz, stream), "post_physics")
        nat.check(lib.tk_gogoro_paper_diff_sums(env._handle, nat.ptr(env.diff_sums), stream), "diff_sums")
        got.fill_(float("nan"))
        nat.check(_static(lib, env.command_ring, n, TURNING if turning else PAPER, step, got, stream), "diff_sums_static")
        nat.check(lib.tk_gogoro_paper_finalize_reward(env._handle, 0, stream), "finalize_reward")
        torch.cuda.synchronize()
        want = env.diff_sums.clone()
        assert float(want[1]) == n * (20 if turning else 19)
        assert float(got[1]) == float(want[1]), f"step {step}: count"
        assert abs(float(want[0])) > 1e-3, "degenerate input: the sum under test is ~0"
        U.close(got[:1], want[:1], rtol=1e-9, atol=0, what=f"step {step}: static vs handle sum")
    # same input again, straight after: the shared scratch holds the previous call's partials and must not leak into this one
    again = torch.zeros(2, device=DEV, dtype=torch.float64)
    nat.check(_static(lib, env.command_ring, n, TURNING if turning else PAPER, 1, again, stream), "diff_sums_static")
    torch.cuda.synchronize()
    assert torch.equal(again, got), "two calls on the same ring differ"


@pytest.mark.gpu
def test_static_scratch_entry_point_on_its_own_with_no_handle_alive():
    """Nothing but the call itself: a ring written by torch, checked against the definition in float64."""
    lib = nat.lib()
    n = 70000                                        # > 256 * 256: exercises the grid-stride loop below the block cap too
    g = torch.Generator().manual_seed(5)
    ring = (torch.rand((n, 20), generator=g) - 0.5).to(DEV)
    out = torch.zeros(2, device=DEV, dtype=torch.float64)
    stream = nat.current_stream_ptr(torch.device(DEV))
    for slot in (0, 7, 19):
        nat.check(_static(lib, ring, n, PAPER, slot, out, stream), "diff_sums_static")
        order = [(slot + 1 + t) % 20 for t in range(20)]
        hist = (ring[:, order] / 0.5)
        want = ((hist[:, 1:] - hist[:, :-1]) ** 2).double().sum()
        torch.cuda.synchronize()
        assert float(out[1]) == 19 * n
        assert abs(float(out[0]) - float(want)) <= 1e-5 * float(want)       # fp32 inside an env on both sides, other order
        nat.check(_static(lib, ring, n, TURNING, slot, out, stream), "diff_sums_static")
        torch.cuda.synchronize()
        assert float(out[1]) == 20 * n
        assert abs(float(out[0]) - float(ring.double().sum())) <= 1e-3


def test_static_scratch_entry_point_refuses_what_it_cannot_hold():
    """Argument checks come before any CUDA call, so they run on a box without a GPU too."""
    lib = nat.lib()
    fake = 256                                       # never dereferenced: every call below is refused first
    assert lib.tk_gogoro_paper_diff_sums_static(fake, MAX_ENVS + 1, PAPER, 0, fake, 0) == ERR_INVALID
    msg = lib.tk_last_error().decode()
    assert "at most 1048576" in msg and str(MAX_ENVS + 1) in msg, msg
    assert lib.tk_gogoro_paper_diff_sums_static(fake, 0, PAPER, 0, fake, 0) == ERR_INVALID
    assert lib.tk_gogoro_paper_diff_sums_static(fake, 16, PAPER, 20, fake, 0) == ERR_INVALID
    assert lib.tk_gogoro_paper_diff_sums_static(fake, 16, 2, 0, fake, 0) == ERR_INVALID
    assert lib.tk_gogoro_paper_diff_sums_static(0, 16, PAPER, 0, fake, 0) == ERR_INVALID
    assert lib.tk_gogoro_paper_diff_sums_static(fake, 16, PAPER, 0, 0, 0) == ERR_INVALID
