"""Out-of-bounds WRITE check of our own (compute-sanitizer is closed on the GPU pool): every device tensor a task owns is
moved into an arena with 4 KB canary bands on both sides, the task is re-bound to the new pointers, a few steps run at
ragged and aligned sizes, and every canary byte must be intact.  Two paddings: 4096 B keeps the 16-byte alignment (bulk-copy
/ TMA path of the kernels), 4104 B breaks it while keeping the 8-byte alignment the (pos, vel) pairs and int64 buffers need (cooperative scalar path)."""
import pytest
import torch

import parity_util as U

pytestmark = pytest.mark.gpu
CANARY = 0xA5


# tensors the C ABI REQUIRES to be 16-byte aligned (tk_gogoro_paper_bind: frames move as 128-bit chunks; tk_ma_op3_bind: obs rows leave as 128-bit stores): they keep pad 4096
ALIGN16 = {"obs_buf", "buffer_obs_ring", "command_ring"}


def _guard_all(env, pad, keep16=()):
    guards = []
    pad_in = pad
    for name, t in list(vars(env).items()):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t._base is None and t.is_contiguous() and t.numel() > 0):
            continue
        pad = 4096 if name in keep16 else pad_in
        nbytes = t.numel() * t.element_size()
        arena = torch.full((pad + nbytes + 4096,), CANARY, dtype=torch.uint8, device=t.device)
        if (arena.data_ptr() + pad) % t.element_size():
            continue                                   # 8-byte types cannot live at a 4-byte offset
        inner = arena[pad:pad + nbytes].view(t.dtype).view(t.shape)
        inner.copy_(t)
        setattr(env, name, inner)
        guards.append((name, arena, pad, nbytes))
    env.rebind()
    return guards


def _check(guards, what):
    torch.cuda.synchronize()
    for name, arena, pad, nbytes in guards:
        assert bool((arena[:pad] == CANARY).all()), f"{what}: bytes written in front of {name}"
        assert bool((arena[pad + nbytes:] == CANARY).all()), f"{what}: bytes written behind {name}"


def _actions(env, st, width):
    a = st.get("actions")
    if a is None:
        a = torch.rand(env.num_envs, width) * 2 - 1
    return a.to(env.device)


MAKERS = {
    "humanoid": (lambda n: U.make_humanoid_pair(n, seed=n), 21),
    "ant": (lambda n: U.make_ant_pair(n, seed=n), 8),
    "gogoro": (lambda n: U.make_gogoro_pair(n, seed=n), 1),
    "gogoro_paper": (lambda n: U.make_gogoro_paper_pair(n, seed=n), 1),
    "gogoro_turning": (lambda n: U.make_gogoro_paper_pair(n, seed=n, turning=True), 1),
    "anymal": (lambda n: U.make_anymal_pair(n, seed=n), 12),
    "ma_op3": (lambda n: U.make_ma_op3_pair(n, seed=n), 44),
}


@pytest.mark.parametrize("task", list(MAKERS))
@pytest.mark.parametrize("n", [33, 4096, 4103])
@pytest.mark.parametrize("pad", [4096, 4104])
def test_no_write_outside_the_bound_tensors(task, n, pad):
    make, width = MAKERS[task]
    env, _, st = make(n)
    env.reset_buf[: min(n, 40)] = 1              # resets in the first tiles, more than four per warp
    guards = _guard_all(env, pad, ALIGN16 if task in ("gogoro_paper", "gogoro_turning", "ma_op3") else ())
    assert len(guards) >= 8, "the task's tensors were not found"
    a = _actions(env, st, width)
    for _ in range(2):
        env.pre_physics_step(a)
        env.post_physics_step()                  # in-kernel Philox noise
    env.episode_statistics()
    _check(guards, f"{task} n={n} pad={pad}")
