"""``Gogoro`` of ``isaacgymenvs/tasks/gogoro_realistic_turning.py`` -- the sibling of the paper file with the rational
reward and the random spawn heading -- on the env-step path.  Registered here as
``"GogoroTurning"`` (the reference file is not in ``isaacgym_task_map``; its class is also called ``Gogoro``).

The two reference files share the class skeleton, buffers and cfg keys; this one differs on the step path in
  * module flags as saved (:22-32): DEBUG = False, hence USE_STEER_DELAY = True (per-env delay-line tap, :360),
    PUSH_ROBOT = False, RANDOM_DAMPING = True (a ``set_actor_dof_properties`` call per reset env, :655-658: simulator
    side, left out like the seat offsets :661-672);
  * ``pre_physics_step`` :365: steering target = delayed command + ``steer_offsets``;
  * ``compute_obs_rwd`` :480-547: the paper file's command-change correction of the newest noisy frame is absent;
  * ``compute_gogoro_reward`` :713-765: ``3/(1+200 yaw^2) + 1/(1+100 tilt^2) + 10 reward3 + 2/(1+100 dtilt^2) + 10 reward5`` with
    ``reward3 = 1/(1+500 (min - min)^2)`` identically 1 (``max_`` is a second ``torch.min``, :738-740) and
    ``reward5 = 1 - |mean(act_buff)|`` a mean over EVERY env and all 20 frames (:742) -- the cross-env term, computed by
    the finalize kernel and all-reduced with ``cfg["env"]["worldReduce"]`` when range-sharded;
  * ``generate_spawn_r`` :572-588 / ``reset_idx`` :601-686: random yaw target, spawn yaw and roll; the reset draws have
    8 columns: speed (:564), steer_delay (:565), steer_offsets (:566), speed-sensor offset (:568), imu_x offset (:653),
    yaw target (:573), yaw offset (:574), roll (:575).

Everything is the CUDA extension with ``variant = TK_GOGORO_PAPER_VARIANT_TURNING`` (include/thormang_b200.h,
csrc/tk_gogoro_paper.cu); no TorchScript, no CPU fallback.
"""
from .. import _native as nat
from . import gogoro_realistic_turning_sim_paper as _paper

NUM_CH, BUFF, NUM_ACT, SENSOR_NOISE_COLS = _paper.NUM_CH, _paper.BUFF, _paper.NUM_ACT, _paper.SENSOR_NOISE_COLS
RESET_NOISE_COLS = 8


class Gogoro(_paper.Gogoro):
    VARIANT = nat.GOGORO_PAPER_VARIANT_TURNING
    RESET_NOISE_COLS = RESET_NOISE_COLS
    USE_STEER_DELAY_DEFAULT = True             # USE_STEER_DELAY = not DEBUG, :31, with DEBUG = False as saved (:22)
