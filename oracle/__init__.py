"""oracle/ -- TEST INFRASTRUCTURE, not product code.

CPU (torch fp32) restatement of the hot path of Yougo-robotics/thormang_isaacgym:
the per-env post-physics work and the pre-physics target write of
``isaacgymenvs/tasks/humanoid.py::Humanoid`` and
``isaacgymenvs/tasks/gogoro_new.py::Gogoro``.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import anything from here, and only as the
checker or the timed CPU baseline.  ``thormang_isaacgym_b200`` never imports
this package; the product path has no CPU fallback.

Pinning status (see DESIGN.md "Oracle"):

* ``humanoid_oracle`` / ``gogoro_oracle`` JIT-level functions are PINNED: they
  are checked bit-for-bit against the reference's own ``@torch.jit.script``
  functions, lifted by source from ``/root/reference`` with
  ``oracle/lift_reference.py`` (container only) and frozen as golden vectors in
  ``tests/golden/*.npz`` by ``tests/golden/make_golden.py``.
* ``isaac_helpers`` restates ``isaacgym.torch_utils`` (NVIDIA Isaac Gym Preview,
  un-vendored binary distribution, NO version pinned by the reference, module
  absent from ``/root/reference``).  No reference test, golden vector or source
  file pins those helpers: **parity unpinned** at that boundary.  They are
  anchored on the reference's call sites and cross-checked against the in-tree
  quaternion library ``isaacgymenvs/tasks/amp/poselib/poselib/core/rotation3d.py``.
* Method-body restatements (``pre_physics_step``, ``post_physics_step``,
  ``reset_idx``, ``compute_obs_rwd``) follow the reference line by line over
  plain tensors with the random draws passed in; the reference holds no
  fixtures for them, so they are pinned only through the
  JIT functions they call.
"""
