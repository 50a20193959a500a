"""``Ant`` -- drop-in for ``isaacgymenvs/tasks/ant.py::Ant`` on the env-step path.

Ant is the Humanoid path with 8 DOFs, 60 observations (no DOF-force block, 24 sensor floats, RAW
Euler angles, no potential columns; ant.py:367-408) and every reward term live (ant.py:318-364).
It runs through the SAME kernel template as Humanoid (``humanoid_warp_kernel<AntSpec, FUSED>`` in
csrc/tk_humanoid_tile.cuh), selected by ``TkHumanoidParams.variant``; the method surface and the
buffers are those of the reference class (ant.py:42-117, 213-287).
"""
from .. import _native as nat
from ..synthetic import ant_asset
from .humanoid import Humanoid

NUM_OBS = 60          # ant.py:67
NUM_DOF = 8           # ant.py:68


class Ant(Humanoid):
    _NUM_OBS, _NUM_DOF, _NUM_SENSOR = NUM_OBS, NUM_DOF, 24                        # ant.py:82-83: 4 foot sensors x 6
    _VARIANT = nat.VARIANT_ANT
    _HAS_DOF_FORCE = False
    _EFFORT_KEY = "joint_gears"                                                   # ant.py:160-162,275
    _asset = staticmethod(ant_asset)

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture=False, force_render=False):
        # ant.py:48-61 reads the same cfg keys as Humanoid minus angularVelocityScale (the Ant obs row carries the raw
        # local angular velocity, ant.py:399)
        cfg["env"].setdefault("angularVelocityScale", 1.0)
        super().__init__(cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture, force_render)
        self.joint_gears = self.motor_efforts                                     # the reference's attribute name

    def reward_terms(self):
        raise NotImplementedError("every Ant reward term is live in rew_buf (ant.py:353-354)")

    def wm_reward(self, obs):
        raise NotImplementedError("humanoid.py:425-472 has no Ant counterpart")
