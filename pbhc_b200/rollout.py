"""RolloutStorage (reference: humanoidverse/agents/modules/data_utils.py:22-152) with the minibatch shuffle-gather
and the time-out bootstrap on the sm_100a kernels of csrc/rollout.cu.  Same class name, method
names, argument meaning and assertions as the reference, so PPO._setup_storage / _rollout_step / _training_step
(agents/ppo/ppo.py:124-135, 240-275, 357-375) run unchanged on it:

    storage = RolloutStorage(num_envs, num_steps_per_env, device)
    storage.register_key('actor_obs', shape=(dim,), dtype=torch.float)
    storage.update_key('actor_obs', obs); ...; storage.increment_step()
    storage.batch_update_data('returns', returns)
    for batch in storage.mini_batch_generator(num_mini_batches, num_epochs): ...
    storage.clear()

What changes underneath:
  * mini_batch_generator gathers EVERY stored key with one launch of pbhc_gather_keys (the reference runs one
    index_select + .contiguous() per key): indices still come from torch.randperm on the storage device, so with
    the same torch seed the batches are bit-identical to the reference's;
  * update_key is free when the producer already wrote into the slot (`slot(key)` hands the step's view to
    TrackingStep.write_obs_into, which makes the fused kernel write its observation rows there directly);
  * bootstrap_rewards(rewards, values, time_outs, gamma) = ppo.py:266-270 in one kernel, writing the stored row.
There is no CPU fallback: the buffers live on a CUDA device and a missing library raises.
"""
import ctypes as C

import torch

from . import _lib


class RolloutStorage(torch.nn.Module):
    def __init__(self, num_envs, num_transitions_per_env, device="cuda"):
        super().__init__()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.PbhcError("RolloutStorage needs a CUDA device (there is no CPU fallback)")
        self.num_transitions_per_env = num_transitions_per_env
        self.num_envs = num_envs
        self.step = 0
        self.stored_keys = list()
        self._bad = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._shuffled = {}
        self._L = _lib.lib()

    # ---- data_utils.py:39-62 ---------------------------------------------------------------------------
    def register_key(self, key, shape=(), dtype=torch.float):
        assert not hasattr(self, key), key
        assert isinstance(shape, (list, tuple)), "shape must be a list or tuple"
        buffer = torch.zeros((self.num_transitions_per_env, self.num_envs) + tuple(shape), dtype=dtype, device=self.device)
        self.register_buffer(key, buffer, persistent=False)
        self.stored_keys.append(key)

    def increment_step(self):
        self.step += 1

    def slot(self, key, step=None):
        """view of one step of a key's buffer ([num_envs, *shape]); a producer that writes here makes update_key a no-op"""
        step = self.step if step is None else step
        assert step < self.num_transitions_per_env, "Rollout buffer overflow"
        return getattr(self, key)[step]

    def update_key(self, key, data):
        assert not data.requires_grad
        assert self.step < self.num_transitions_per_env, "Rollout buffer overflow"
        dst = getattr(self, key)[self.step]
        if data.data_ptr() == dst.data_ptr() and data.shape == dst.shape and data.dtype == dst.dtype and data.is_contiguous():
            return  # written in place by the producer
        dst.copy_(data)

    def batch_update_data(self, key, data):
        assert not data.requires_grad
        buf = getattr(self, key)
        if data.data_ptr() == buf.data_ptr() and data.shape == buf.shape:
            return
        buf[:] = data

    def clear(self):
        self.step = 0

    def get_statistics(self):
        raise NotImplementedError  # data_utils.py:91-92

    def query_key(self, key):
        assert hasattr(self, key), key
        return getattr(self, key)

    # ---- ppo.py:266-270 ---------------------------------------------------------------------------------
    def bootstrap_rewards(self, rewards, values, time_outs, gamma, key="rewards"):
        """stores `rewards.unsqueeze(1) + gamma * values * time_outs.unsqueeze(1)` into this step's `key` slot and
        returns the slot (rewards [N] or [N,R], values [N,R], time_outs [N] bool)."""
        dst = self.slot(key)
        R = dst.shape[1] if dst.dim() > 1 else 1
        dst.copy_(rewards.reshape(self.num_envs, -1))
        to = time_outs.contiguous()
        to8 = to.view(torch.uint8) if to.dtype == torch.bool else to.to(torch.uint8)
        values = values.contiguous()
        _lib.check(self._L.pbhc_timeout_bootstrap(_lib.ptr(dst, torch.float32), _lib.ptr(values, torch.float32), to8.data_ptr(),
                                                  self.num_envs, R, float(gamma), _lib.stream_ptr(self.device)))
        return dst

    # ---- data_utils.py:116-152 ----------------------------------------------------------------------------
    def gather(self, indices, keys=None, out=None, check=True):
        """{key: buf.flatten(0, 1)[indices]} for every stored key in one launch.  `out` = dict of preallocated
        destinations to reuse (each [len(indices), *shape])."""
        keys = list(self.stored_keys if keys is None else keys)
        n = int(indices.numel())
        rows = self.num_envs * self.num_transitions_per_env
        idx = indices.contiguous()
        if idx.dtype != torch.int64 or idx.device != self.device:
            raise _lib.PbhcError("gather: indices must be an int64 tensor on the storage device")
        res = {}
        for lo in range(0, len(keys), _lib.MAX_GATHER_KEYS):
            part = keys[lo:lo + _lib.MAX_GATHER_KEYS]
            arr = (_lib.GatherKey * len(part))()
            for i, k in enumerate(part):
                buf = getattr(self, k)
                assert buf.is_contiguous()
                dst = out[k] if out is not None and k in out else torch.empty((n,) + tuple(buf.shape[2:]), dtype=buf.dtype, device=self.device)
                assert dst.is_contiguous() and dst.dtype == buf.dtype and tuple(dst.shape) == (n,) + tuple(buf.shape[2:]), k
                row_bytes = buf[0, 0].numel() * buf.element_size()
                arr[i].src, arr[i].dst, arr[i].row_bytes = buf.data_ptr(), dst.data_ptr(), row_bytes
                res[k] = dst
            if n:
                _lib.check(self._L.pbhc_gather_keys(C.cast(arr, C.c_void_p), len(part), idx.data_ptr(), n, rows, self._bad.data_ptr(),
                                                    _lib.stream_ptr(self.device)))
        if check and n and int(self._bad.item()):
            self._bad.zero_()
            raise IndexError(f"gather: index out of range for {rows} stored rows")
        return res

    def mini_batch_generator(self, num_mini_batches, num_epochs=8):
        batch_size = self.num_envs * self.num_transitions_per_env
        mini_batch_size = batch_size // num_mini_batches
        indices = torch.randperm(batch_size, device=self.device)
        # destinations are kept between updates: PPO calls this once per iteration with the same shapes
        reuse = {k: v for k, v in self._shuffled.items()
                 if hasattr(self, k) and tuple(v.shape) == (batch_size,) + tuple(getattr(self, k).shape[2:])}
        shuffled_data = self.gather(indices, out=reuse, check=False)
        self._shuffled = shuffled_data
        for epoch in range(num_epochs):
            for i in range(num_mini_batches):
                start = i * mini_batch_size
                end = (i + 1) * mini_batch_size
                yield {key: shuffled_data[key][start:end] for key in self.stored_keys}
