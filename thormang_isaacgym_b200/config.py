"""Plain-dict task configs (hydra / omegaconf are not installed here and are outside the hot path).

Values restate ``/root/reference/isaacgymenvs/cfg/task/Humanoid.yaml:7-29,42-47`` and
``cfg/task/Gogoro.yaml:5-93`` -- the constants the env-step path consumes.  The dicts have the
shape ``isaacgymenvs.make`` hands to the task constructors after ``omegaconf_to_dict``
(isaacgymenvs/__init__.py:37-41).
"""
import copy


def humanoid_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    return copy.deepcopy({
        "name": "Humanoid",
        "physics_engine": "physx",
        "seed": seed,
        "env": {
            "numEnvs": num_envs, "envSpacing": 5, "episodeLength": 1000, "enableDebugVis": False,
            "clipActions": 1.0, "powerScale": 1.0,
            "headingWeight": 0.5, "upWeight": 0.1,
            "actionsCost": 0.01, "energyCost": 0.05, "dofVelocityScale": 0.1, "angularVelocityScale": 0.25,
            "contactForceScale": 0.01, "jointsAtLimitCost": 0.25, "deathCost": -1.0, "terminationHeight": 0.8,
            "plane": {"staticFriction": 1.0, "dynamicFriction": 1.0, "restitution": 0.0},
        },
        "sim": {"dt": 0.0166, "substeps": 2, "up_axis": "z", "use_gpu_pipeline": True, "gravity": [0.0, 0.0, -9.81]},
        "task": {"randomize": False, "randomization_params": {}},
    })


def ant_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    """cfg/task/Ant.yaml:7-29,42-47."""
    return copy.deepcopy({
        "name": "Ant",
        "physics_engine": "physx",
        "seed": seed,
        "env": {
            "numEnvs": num_envs, "envSpacing": 5, "episodeLength": 1000, "enableDebugVis": False,
            "clipActions": 1.0, "powerScale": 1.0, "controlFrequencyInv": 1,
            "headingWeight": 0.5, "upWeight": 0.1,
            "actionsCost": 0.005, "energyCost": 0.05, "dofVelocityScale": 0.2, "contactForceScale": 0.1,
            "jointsAtLimitCost": 0.1, "deathCost": -2.0, "terminationHeight": 0.31,
            "plane": {"staticFriction": 1.0, "dynamicFriction": 1.0, "restitution": 0.0},
        },
        "sim": {"dt": 0.0166, "substeps": 2, "up_axis": "z", "use_gpu_pipeline": True, "gravity": [0.0, 0.0, -9.81]},
        "task": {"randomize": False, "randomization_params": {}},
    })


GOGORO_JOINTS_POS = {
    "l_arm_el_y": 0.0, "l_arm_wr_r": 0.0, "head_y": 0.0, "r_arm_grip": 0.0, "l_arm_wr_p": 0.0, "torso_y": 0.0,
    "r_arm_sh_r": -1.57, "l_arm_sh_p1": 0.0, "l_arm_sh_r": 1.57, "l_leg_an_r": 0.0, "l_leg_an_p": 0.0,
    "r_leg_hip_p": 1.4, "r_leg_an_p": 0.0, "l_arm_wr_y": 0.0, "l_leg_hip_p": -1.4, "r_leg_hip_y": 0.0,
    "l_leg_hip_r": 0.0, "l_leg_kn_p": 1.4, "r_arm_sh_p2": 0.0, "r_arm_sh_p1": 0.0, "l_leg_hip_y": 0.0,
    "r_leg_hip_r": 0.0, "l_arm_sh_p2": 0.0, "r_arm_wr_y": 0.0, "head_p": 0.0, "r_arm_wr_p": 0.0, "r_arm_wr_r": 0.0,
    "r_arm_el_y": 0.0, "l_arm_grip": 0.0, "r_leg_an_r": 0.0, "r_leg_kn_p": -1.4,
}


def gogoro_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    return copy.deepcopy({
        "name": "Gogoro",
        "physics_engine": "physx",
        "seed": seed,
        "env": {"numEnvs": num_envs, "max_steps": 1000},
        "sim": {"dt": 0.03, "up_axis": "z", "use_gpu_pipeline": True, "gravity": [0.0, 0.0, -9.81], "substeps": 3},
        "noises": {
            "seat_offset_x_range": [0, 0.02], "seat_offset_y_range": [0, 0.02], "seat_offset_z_range": [0, 0.02],
            "steering_offset": [0, 0.01], "imu_filter_noise": [0, 0.001], "imu_noise": [0, 0.001],
            "seat_offset_xr_range": [0, 0.05], "speed_sensor_offset": [-0.5, 0.5], "speed_sensor_noise": [0, 0.3],
            "steering_action_noise": [0, 0.03], "spawn_x_angle": [0, 0.05], "steering_damping_range": [100, 1000],
            "IK_noise_amplitude": [0, 3], "speed_range": [4.0, 13.0], "speed_freq_update": 300, "yaw_freq_update": 300,
        },
        "joints_pos": dict(GOGORO_JOINTS_POS),
        "task": {"randomize": False, "randomization_params": {}},
    })


def gogoro_paper_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    """cfg/task/Gogoro_paper.yaml:5-58 (+ joints_pos :62-93, same riding pose as Gogoro.yaml)."""
    return copy.deepcopy({
        "name": "Gogoro",
        "physics_engine": "physx",
        "seed": seed,
        "env": {"numEnvs": num_envs, "max_steps": 3600},
        "sim": {"dt": 0.03, "up_axis": "z", "use_gpu_pipeline": True, "gravity": [0.0, 0.0, -9.81], "substeps": 3},
        "noises": {
            "imu_filter_noise": [-0.003, 0.003], "imu_noise": [-0.003, 0.003], "speed_sensor_offset": [-0.3, 0.3],
            "speed_sensor_noise": [0, 0.3], "seat_offset_x_range": [-0.1, 0.1], "seat_offset_y_range": [-0.1, 0.1],
            "seat_offset_z_range": [-0.05, 0.05], "imu_x_offset": [-0.02, 0.02], "spawn_x_angle": [-0.02, 0.02],
            "steering_action_noise": [-0.05, 0.05], "steering_offset": [-0.05, 0.05], "command_delay": [0, 5],
            "steering_damping_range": [50, 1000], "IK_noise_amplitude": [0, 3], "speed_range": [5.0, 20.0],
            "speed_freq_update": 300, "yaw_freq_update": 300,
        },
        "joints_pos": dict(GOGORO_JOINTS_POS),
        "task": {"randomize": False, "randomization_params": {}},
    })


def gogoro_turning_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    """``tasks/gogoro_realistic_turning.py`` reads the same cfg keys as the paper file (:44-84) and has no yaml of its own:
    cfg/task/Gogoro_paper.yaml values."""
    return gogoro_paper_cfg(num_envs, seed)


ANYMAL_DOF_NAMES = ["LF_HAA", "LF_HFE", "LF_KFE", "LH_HAA", "LH_HFE", "LH_KFE", "RF_HAA", "RF_HFE", "RF_KFE",
                    "RH_HAA", "RH_HFE", "RH_KFE"]      # DOF order of anymal_c/urdf/anymal.urdf


def anymal_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    """cfg/task/Anymal.yaml:6-64,80-84."""
    return copy.deepcopy({
        "name": "Anymal",
        "physics_engine": "physx",
        "seed": seed,
        "env": {
            "numEnvs": num_envs, "envSpacing": 4.0, "clipObservations": 5.0, "clipActions": 1.0,
            "plane": {"staticFriction": 1.0, "dynamicFriction": 1.0, "restitution": 0.0},
            "baseInitState": {"pos": [0.0, 0.0, 0.62], "rot": [0.0, 0.0, 0.0, 1.0], "vLinear": [0.0, 0.0, 0.0],
                              "vAngular": [0.0, 0.0, 0.0]},
            "randomCommandVelocityRanges": {"linear_x": [-2.0, 2.0], "linear_y": [-1.0, 1.0], "yaw": [-1.0, 1.0]},
            "control": {"stiffness": 85.0, "damping": 2.0, "actionScale": 0.5, "controlFrequencyInv": 1},
            "defaultJointAngles": {"LF_HAA": 0.03, "LH_HAA": 0.03, "RF_HAA": -0.03, "RH_HAA": -0.03,
                                   "LF_HFE": 0.4, "LH_HFE": -0.4, "RF_HFE": 0.4, "RH_HFE": -0.4,
                                   "LF_KFE": -0.8, "LH_KFE": 0.8, "RF_KFE": -0.8, "RH_KFE": 0.8},
            "urdfAsset": {"collapseFixedJoints": True, "fixBaseLink": False, "defaultDofDriveMode": 4},
            "learn": {"linearVelocityXYRewardScale": 1.0, "angularVelocityZRewardScale": 0.5,
                      "torqueRewardScale": -0.000025, "linearVelocityScale": 2.0, "angularVelocityScale": 0.25,
                      "dofPositionScale": 1.0, "dofVelocityScale": 0.05, "episodeLength_s": 50},
        },
        "sim": {"dt": 0.02, "substeps": 2, "up_axis": "z", "use_gpu_pipeline": True, "gravity": [0.0, 0.0, -9.81]},
        "task": {"randomize": False, "randomization_params": {}},
    })


OP3_DOF_NAMES = ["head_pan", "head_tilt", "l_hip_yaw", "l_hip_roll", "l_hip_pitch", "l_knee", "l_ank_pitch", "l_ank_roll",
                 "l_sho_pitch", "l_sho_roll", "l_el", "l_gr", "r_hip_yaw", "r_hip_roll", "r_hip_pitch", "r_knee",
                 "r_ank_pitch", "r_ank_roll", "r_sho_pitch", "r_sho_roll", "r_el", "r_gr"]   # MA_OP3.yaml defaultJointAngles order


def ma_op3_cfg(num_envs: int = 4096, seed: int = 0) -> dict:
    """cfg/task/MA_OP3.yaml."""
    angles = {k: 0.0 for k in OP3_DOF_NAMES}
    angles.update({"l_hip_pitch": -0.95, "l_knee": 1.49, "l_ank_pitch": -0.0, "l_gr": -0.0, "r_hip_pitch": 0.95,
                   "r_knee": -1.49, "r_sho_roll": -0.0, "r_el": -0.0})
    return copy.deepcopy({
        "name": "MA_OP3",
        "physics_engine": "physx",
        "isMultiAgent": True,
        "seed": seed,
        "env": {
            "numEnvs": num_envs, "numAgents": 2, "envSpacing": 1.0, "episodeLength": 1000, "clipObservations": 5.0,
            "clipActions": 1.0, "staticFriction": 1.0, "dynamicFriction": 1.0, "restitution": 0.0, "enableDebugVis": False,
            "baseInitState": {"pos": [0.0, 0.0, 0.255], "rot": [0.0, 0.0, 0.0, 1.0], "vLinear": [0.0, 0.0, 0.0],
                              "vAngular": [0.0, 0.0, 0.0]},
            "randomCommandVelocityRanges": {"linear_x": [0.0, 0.0], "linear_y": [0.0, 10.0], "yaw": [0.0, 0.0],
                                            "heading": [-3.14, 3.14]},
            "control": {"stiffness": 1000.0, "damping": 200.0, "actionScale": 1.0, "powerScale": 0.1, "controlFrequencyInv": 1},
            "defaultJointAngles": angles,
            "urdfAsset": {"collapseFixedJoints": True, "fixBaseLink": False, "defaultDofDriveMode": 4},
            "learn": {"linearVelocityXYRewardScale": 1.0, "angularVelocityZRewardScale": 0.5, "torqueRewardScale": -0.000025,
                      "headingScale": 1, "upScale": 0.1, "actionRateRewardScale": -0.01, "hipRewardScale": -0.0,
                      "jointAccRewardScale": -0.0005, "standRewardScale": -0, "noflyRewardScale": 0.5,
                      "feetAirTimeRewardScale": 0.5, "syncronizeHipRewardScale": -0.00025, "linearVelocityScale": 3.0,
                      "angularVelocityScale": 0.25, "dofPositionScale": 1.0, "dofVelocityScale": 0.01,
                      "episodeLength_s": 50, "pushInterval_s": 15},
            "color": [[1.0, 0.2, 0.2], [0.23, 0.50, 0.92]],
        },
        "sim": {"dt": 0.0166, "substeps": 2, "up_axis": "z", "use_gpu_pipeline": True, "gravity": [0.0, 0.0, -9.81]},
        "task": {"randomize": False, "randomization_params": {}},
    })
