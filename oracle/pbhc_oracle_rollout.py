"""TEST INFRASTRUCTURE ONLY - numpy restatement of the reference's RolloutStorage data movement.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product package
(pbhc_b200/) never does.

Follows humanoidverse/agents/modules/data_utils.py:22-152 (RolloutStorage) and agents/ppo/ppo.py:266-270 (time-out
bootstrap of the stored rewards).  Pinned against the unmodified reference by tests/golden/ref_rollout.npz
(tools/make_golden_rollout.py runs the reference class on CPU) in tests/test_oracle_golden.py.

torch.randperm's stream is torch's own, so the permutation is an INPUT here (the product keeps calling
torch.randperm and is therefore seed-compatible with the reference).
"""
import numpy as np

_NP = {"float": np.float32, "bool": np.bool_, "long": np.int64}


class RolloutStorage:
    def __init__(self, num_envs, num_transitions_per_env):  # data_utils.py:24-37
        self.num_envs, self.num_transitions_per_env = num_envs, num_transitions_per_env
        self.step = 0
        self.stored_keys = []
        self.buf = {}

    def register_key(self, key, shape=(), dtype="float"):  # data_utils.py:39-47: zeros [T, N, *shape]
        assert key not in self.buf, key
        assert isinstance(shape, (list, tuple)), "shape must be a list or tuple"
        self.buf[key] = np.zeros((self.num_transitions_per_env, self.num_envs) + tuple(shape), _NP[dtype])
        self.stored_keys.append(key)

    def increment_step(self):  # data_utils.py:49-50
        self.step += 1

    def update_key(self, key, data):  # data_utils.py:52-56: buf[step].copy_(data) (broadcasting, dtype cast)
        assert self.step < self.num_transitions_per_env, "Rollout buffer overflow"
        self.buf[key][self.step] = np.asarray(data)

    def batch_update_data(self, key, data):  # data_utils.py:58-61
        self.buf[key][:] = np.asarray(data)

    def clear(self):  # data_utils.py:82-83
        self.step = 0

    def query_key(self, key):  # data_utils.py:94-96
        assert key in self.buf, key
        return self.buf[key]

    def mini_batch_generator(self, num_mini_batches, num_epochs, indices):
        """data_utils.py:116-152 with `indices` = the torch.randperm(T*N) the reference draws at :134.
        Every key is flattened over (T, N), gathered ONCE, then cut into num_mini_batches contiguous slices of
        batch_size // num_mini_batches rows (the remainder rows are never yielded); every epoch yields the
        same slices again."""
        batch_size = self.num_envs * self.num_transitions_per_env
        mini = batch_size // num_mini_batches
        idx = np.asarray(indices, np.int64)
        if idx.size and (idx.min() < 0 or idx.max() >= batch_size):
            raise IndexError("index out of range")
        shuffled = {k: self.buf[k].reshape((batch_size,) + self.buf[k].shape[2:])[idx] for k in self.stored_keys}
        for _ in range(num_epochs):
            for i in range(num_mini_batches):
                yield {k: shuffled[k][i * mini:(i + 1) * mini] for k in self.stored_keys}


def bootstrap_rewards(rewards, values, time_outs, gamma):
    """ppo.py:266-270: rewards_stored = rewards.clone().unsqueeze(1); rewards_stored += gamma * values * time_outs.unsqueeze(1)
    (fp32: gamma * values rounded, times the 0/1 mask, then the sum rounded)."""
    r = np.asarray(rewards, np.float32).reshape(len(values), -1).copy()
    g = np.float32(gamma) * np.asarray(values, np.float32)
    r += g * np.asarray(time_outs).astype(np.float32).reshape(-1, 1)
    return r
