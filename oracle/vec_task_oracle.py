"""CPU restatement of the ``VecTask.step`` prologue / epilogue and of the non-physical domain-randomisation lambdas
(SURVEY.md 8(f) row N4).  TEST INFRASTRUCTURE.

Follows ``/root/reference/isaacgymenvs/tasks/base/vec_task.py`` :323-327 (action noise + clamp), :347-353 (observation
noise + clamp) and :576-646 (``apply_randomizations``: gaussian | uniform, additive | scaling, linear | constant schedule,
correlated component).  ``vec_task.py`` cannot be imported (``from isaacgym import gymapi``, :35) and the lambdas are
closures built inside ``apply_randomizations``, so there is nothing to lift: this is a line-by-line restatement with the
random draws passed in (``unit``: what ``torch.randn_like`` / ``torch.rand_like`` return; ``corr``: the persistent
``params['corr']`` draw).  Parity unpinned by reference fixtures (the reference has none for this path).
"""
import operator
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor


def scheduled_params(prm: dict, last_step: int) -> Dict[str, float]:
    """vec_task.py:578-634: the dict stored in ``dr_randomizations[name]`` (without the lambda)."""
    dist, op_type = prm["distribution"], prm["operation"]
    sched_type = prm["schedule"] if "schedule" in prm else None
    sched_step = prm["schedule_steps"] if "schedule" in prm else None
    if sched_type == 'linear':
        sched_scaling = 1.0 / sched_step * min(last_step, sched_step)               # :585-586
    elif sched_type == 'constant':
        sched_scaling = 0 if last_step < sched_step else 1                          # :588
    else:
        sched_scaling = 1
    if dist == 'gaussian':
        mu, var = prm["range"]                                                      # :593
        mu_corr, var_corr = prm.get("range_correlated", [0., 0.])
        if op_type == 'additive':
            mu *= sched_scaling; var *= sched_scaling; mu_corr *= sched_scaling; var_corr *= sched_scaling   # :597-600
        elif op_type == 'scaling':
            var = var * sched_scaling                                               # :602
            mu = mu * sched_scaling + 1.0 * (1.0 - sched_scaling)                   # :603-604
            var_corr = var_corr * sched_scaling                                     # :606
            mu_corr = mu_corr * sched_scaling + 1.0 * (1.0 - sched_scaling)         # :607-608
        return {'mu': mu, 'var': var, 'mu_corr': mu_corr, 'var_corr': var_corr}
    lo, hi = prm["range"]                                                           # :624
    lo_corr, hi_corr = prm.get("range_correlated", [0., 0.])
    if op_type == 'additive':
        lo *= sched_scaling; hi *= sched_scaling; lo_corr *= sched_scaling; hi_corr *= sched_scaling          # :628-631
    elif op_type == 'scaling':
        lo = lo * sched_scaling + 1.0 * (1.0 - sched_scaling)                       # :633-636
        hi = hi * sched_scaling + 1.0 * (1.0 - sched_scaling)
        lo_corr = lo_corr * sched_scaling + 1.0 * (1.0 - sched_scaling)
        hi_corr = hi_corr * sched_scaling + 1.0 * (1.0 - sched_scaling)
    return {'lo': lo, 'hi': hi, 'lo_corr': lo_corr, 'hi_corr': hi_corr}


def noise_lambda(tensor: Tensor, params: Dict[str, float], op_type: str, corr: Tensor, unit: Tensor) -> Tensor:
    """vec_task.py:612-620 (gaussian, ``unit`` = randn_like) / :635-641 (uniform, ``unit`` = rand_like)."""
    op = operator.add if op_type == 'additive' else operator.mul                    # :582
    if 'mu' in params:
        corr = corr * params['var_corr'] + params['mu_corr']                        # :617
        return op(tensor, corr + unit * params['var'] + params['mu'])               # :618-619
    corr = corr * (params['hi_corr'] - params['lo_corr']) + params['lo_corr']       # :639
    return op(tensor, corr + unit * (params['hi'] - params['lo']) + params['lo'])   # :640


def step_epilogue(obs_buf: Tensor, clip_obs: float, prm: Optional[dict] = None, last_step: int = 0,
                  corr: Optional[Tensor] = None, unit: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """vec_task.py:347-353 -> (new obs_buf, obs_dict["obs"])."""
    if prm:
        obs_buf = noise_lambda(obs_buf, scheduled_params(prm, last_step), prm["operation"], corr, unit)   # :349
    return obs_buf, torch.clamp(obs_buf, -clip_obs, clip_obs)                       # :353


def step_prologue(actions: Tensor, clip_actions: float, prm: Optional[dict] = None, last_step: int = 0,
                  corr: Optional[Tensor] = None, unit: Optional[Tensor] = None) -> Tensor:
    """vec_task.py:323-327 -> action_tensor."""
    if prm:
        actions = noise_lambda(actions, scheduled_params(prm, last_step), prm["operation"], corr, unit)   # :325
    return torch.clamp(actions, -clip_actions, clip_actions)                        # :327
