"""GAE + advantage normalisation (reference: agents/ppo/ppo.py:309-355 PPO._compute_returns,
agents/mh_ppo/mh_ppo.py:348-395 MHPPO._compute_returns, agents/ppo/ppo_mimic.py:445-495).

compute_returns(values, rewards, dones, last_values, gamma, lam) -> (returns [T,N,R], advantages [T,N,1])
runs the sm_100a scan pbhc_gae, all-reduces the three advantage statistics (count, sum, sum of
squares; fp64) over the process group when one is given (envs are sharded across GPUs),
then pbhc_adv_normalize.  No CPU fallback.
"""
import torch

from . import _lib


def compute_returns(values, rewards, dones, last_values, gamma, lam, group=None, normalize=True):
    if not values.is_cuda:
        raise _lib.PbhcError("compute_returns needs CUDA tensors (there is no CPU fallback)")
    T, N = values.shape[0], values.shape[1]
    R = values.shape[2] if values.dim() > 2 else 1
    dev = values.device
    values = values.contiguous()
    rewards = rewards.contiguous()
    last_values = last_values.contiguous()
    d8 = dones.contiguous().view(torch.uint8) if dones.dtype == torch.bool else dones.to(torch.uint8).contiguous()
    returns = torch.empty_like(values)
    adv = torch.empty(T, N, 1, device=dev, dtype=torch.float32)
    stats = torch.zeros(3, device=dev, dtype=torch.float64)
    L, st = _lib.lib(), _lib.stream_ptr(dev)
    _lib.check(L.pbhc_gae(_lib.ptr(rewards, torch.float32), _lib.ptr(values, torch.float32), d8.data_ptr(),
                          _lib.ptr(last_values, torch.float32), T, N, R, float(gamma), float(lam), _lib.ptr(returns),
                          _lib.ptr(adv), _lib.ptr(stats), st))
    if not normalize:
        return returns, adv
    from .dist import all_reduce_sum_
    all_reduce_sum_(stats, group)
    _lib.check(L.pbhc_adv_normalize(_lib.ptr(adv), T * N, _lib.ptr(stats), st))
    return returns, adv


class GaeMixin:
    """Drop-in for `_compute_returns` of the reference PPO classes: mix into (or monkey-patch onto) an
    algo object that has .gamma, .lam and ._critic_eval_step(obs) like ppo.py:309-336."""

    def _compute_returns(self, last_obs_dict, policy_state_dict):
        last_values = self.critic.evaluate(last_obs_dict["critic_obs"]).detach()
        return compute_returns(policy_state_dict["values"], policy_state_dict["rewards"], policy_state_dict["dones"],
                               last_values, self.gamma, self.lam)
