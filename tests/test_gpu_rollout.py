"""GPU parity of pbhc_b200.rollout.RolloutStorage (csrc/rollout.cu through the C ABI) against the reference golden
(tests/golden/ref_rollout.npz, recorded from the reference class) and the numpy oracle.  Byte moves: bit-exact."""
import numpy as np
import pytest
import torch

from tests._util import gold
from tests.test_oracle_golden import oracle_rollout_from_golden, rollout_perm, ROLLOUT_DTYPES

pytestmark = pytest.mark.gpu
TDT = {"float": torch.float, "bool": torch.bool, "long": torch.long}


def _storage_from_golden(g, dev):
    from pbhc_b200.rollout import RolloutStorage
    T, N, MB, EPOCHS = (int(x) for x in g["meta"])
    st = RolloutStorage(N, T, device=dev)
    keys = [k[3:] for k in g if k.startswith("in_") and k != "in_returns"]
    for k in keys:
        st.register_key(k, shape=tuple(g[f"in_{k}"].shape[2:]), dtype=TDT[ROLLOUT_DTYPES[str(g[f"in_{k}"].dtype)]])
    for step in range(T):
        for k in keys:
            st.update_key(k, torch.from_numpy(g[f"in_{k}"][step]).to(dev))
        st.increment_step()
    st.register_key("returns", shape=(1,), dtype=torch.float)
    st.batch_update_data("returns", torch.from_numpy(g["in_returns"]).to(dev))
    return st, (T, N, MB, EPOCHS)


def test_reference_api_and_golden_minibatches():
    dev = torch.device("cuda:0")
    g = gold("ref_rollout.npz")
    st, (T, N, MB, EPOCHS) = _storage_from_golden(g, dev)
    with pytest.raises(AssertionError):
        st.update_key("rewards", torch.zeros(N, 1, device=dev))  # "Rollout buffer overflow" like data_utils.py:55
    with pytest.raises(AssertionError):
        st.register_key("rewards", shape=(1,))
    assert st.query_key("actor_obs").shape == (T, N, 11)
    # the reference's own permutation (recorded) through the one-launch gather
    perm = torch.from_numpy(rollout_perm(g, T * N)).to(dev)
    sh = st.gather(perm)
    mini = (T * N) // MB
    for k in st.stored_keys:
        for i in range(MB * EPOCHS):
            j = i % MB
            got = sh[k][j * mini:(j + 1) * mini].cpu().numpy()
            assert got.dtype == g[f"mb_{k}"].dtype and np.array_equal(got, g[f"mb_{k}"][i]), (k, i)
    # generator contract: MB * EPOCHS batches of `mini` rows, every epoch the same slices, rows = buffer rows of row_id
    batches = list(st.mini_batch_generator(MB, EPOCHS))
    assert len(batches) == MB * EPOCHS
    flat = {k: st.query_key(k).flatten(0, 1) for k in st.stored_keys}
    for i, b in enumerate(batches):
        assert set(b) == set(st.stored_keys)
        ids = b["row_id"].reshape(-1)
        assert ids.numel() == mini
        for k in st.stored_keys:
            assert torch.equal(b[k], flat[k][ids]), k
            assert torch.equal(b[k], batches[i % MB][k])
    seen = torch.cat([b["row_id"].reshape(-1) for b in batches[:MB]])
    assert seen.unique().numel() == seen.numel()  # a permutation: no row twice
    st.clear()
    assert st.step == 0


def test_same_seed_same_batches_as_eager_torch():
    """with the same torch seed the product's batches equal the reference expression run in eager torch on the same device"""
    dev = torch.device("cuda:0")
    from pbhc_b200.rollout import RolloutStorage
    T, N = 6, 50
    st = RolloutStorage(N, T, device=dev)
    st.register_key("obs", shape=(37,), dtype=torch.float)
    st.register_key("dones", shape=(1,), dtype=torch.bool)
    st.batch_update_data("obs", torch.randn(T, N, 37, device=dev))
    st.batch_update_data("dones", torch.rand(T, N, 1, device=dev) < 0.5)
    torch.manual_seed(5)
    mine = [{k: v.clone() for k, v in b.items()} for b in st.mini_batch_generator(4, 2)]
    torch.manual_seed(5)
    idx = torch.randperm(T * N, device=dev)
    ref = {k: getattr(st, k)[:].flatten(0, 1)[idx].contiguous() for k in st.stored_keys}
    mini = (T * N) // 4
    for i, b in enumerate(mine):
        j = i % 4
        for k in st.stored_keys:
            assert torch.equal(b[k], ref[k][j * mini:(j + 1) * mini])


@pytest.mark.parametrize("T,N", [(1, 1), (3, 33), (24, 4096), (5, 40001)])
def test_gather_all_widths_vs_oracle(T, N):
    """row widths covering the 1/4/8/16-byte vector paths, rows narrower and wider than a warp, odd strides"""
    from oracle import pbhc_oracle_rollout as OR
    from pbhc_b200.rollout import RolloutStorage
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(T * 1000 + N)
    shapes = {"w1": ((1,), "float"), "w2": ((2,), "float"), "w3": ((3,), "float"), "w4": ((4,), "float"), "w12": ((12,), "float"),
              "w31": ((31,), "float"), "w130": ((130,), "float"), "w380": ((380,), "float"), "w1540": ((1540,), "float"),
              "w33": ((33,), "float"), "w35": ((35,), "float"), "w462": ((462,), "float"), "w463": ((463,), "float"), "l17": ((17,), "long"),
              "b1": ((1,), "bool"), "b7": ((7,), "bool"), "b16": ((16,), "bool"), "l1": ((1,), "long"), "l3": ((3,), "long"),
              "m": ((3, 5), "float")}
    if N > 1000:
        shapes.pop("w1540")
    st, orc = RolloutStorage(N, T, device=dev), OR.RolloutStorage(N, T)
    for k, (shape, dt) in shapes.items():
        st.register_key(k, shape=shape, dtype=TDT[dt])
        orc.register_key(k, shape=shape, dtype=dt)
        if dt == "float":
            data = rng.standard_normal((T, N) + shape).astype(np.float32)
        elif dt == "bool":
            data = rng.random((T, N) + shape) < 0.5
        else:
            data = rng.integers(-2**40, 2**40, (T, N) + shape)
        st.batch_update_data(k, torch.from_numpy(data).to(dev))
        orc.batch_update_data(k, data)
    perm = rng.permutation(T * N)
    want = next(orc.mini_batch_generator(1, 1, perm))
    got = st.gather(torch.from_numpy(perm).to(dev))
    for k in shapes:
        assert np.array_equal(got[k].cpu().numpy(), want[k]), k
    # repeated / partial index lists and preallocated destinations
    sub = rng.integers(0, T * N, 77)
    out = {k: torch.full((77,) + tuple(getattr(st, k).shape[2:]), 1, dtype=getattr(st, k).dtype, device=dev) for k in ("w3", "b7")}
    got = st.gather(torch.from_numpy(sub).to(dev), keys=["w3", "b7"], out=out)
    assert got["w3"].data_ptr() == out["w3"].data_ptr()
    for k in ("w3", "b7"):
        assert np.array_equal(got[k].cpu().numpy(), orc.buf[k].reshape((T * N,) + orc.buf[k].shape[2:])[sub]), k
    # empty index list
    e = st.gather(torch.empty(0, dtype=torch.int64, device=dev), keys=["w3"])
    assert e["w3"].shape == (0, 3)


def test_gather_rejects_bad_indices_and_host_tensors():
    from pbhc_b200 import _lib
    from pbhc_b200.rollout import RolloutStorage
    dev = torch.device("cuda:0")
    st = RolloutStorage(8, 2, device=dev)
    st.register_key("x", shape=(5,), dtype=torch.float)
    st.batch_update_data("x", torch.arange(80, device=dev, dtype=torch.float).reshape(2, 8, 5))
    with pytest.raises(IndexError):
        st.gather(torch.tensor([0, 16], device=dev))
    with pytest.raises(IndexError):
        st.gather(torch.tensor([-1], device=dev))
    ok = st.gather(torch.tensor([15, 0], device=dev))  # the flag was cleared
    assert ok["x"][0, 0].item() == 75.0 and ok["x"][1, 4].item() == 4.0
    with pytest.raises(_lib.PbhcError):
        st.gather(torch.tensor([0]))
    with pytest.raises(_lib.PbhcError):
        RolloutStorage(8, 2, device="cpu")


@pytest.mark.parametrize("tag", ["r1", "r3"])
def test_timeout_bootstrap_matches_reference_golden(tag):
    from pbhc_b200.rollout import RolloutStorage
    dev = torch.device("cuda:0")
    g = gold("ref_rollout.npz")
    rewards, values, to = (torch.from_numpy(g[f"bs_{tag}_{n}"]).to(dev) for n in ("rewards", "values", "time_outs"))
    R = values.shape[1]
    st = RolloutStorage(rewards.shape[0], 2, device=dev)
    st.register_key("rewards", shape=(R,), dtype=torch.float)
    st.increment_step()
    slot = st.bootstrap_rewards(rewards, values, to, 0.99)
    assert slot.data_ptr() == st.rewards[1].data_ptr()
    assert np.array_equal(st.rewards[1].cpu().numpy(), g[f"bs_{tag}_out"])
    assert not st.rewards[0].any()
    # update_key on the slot itself is the in-place no-op
    st.update_key("rewards", slot)
    assert np.array_equal(st.rewards[1].cpu().numpy(), g[f"bs_{tag}_out"])


def test_timeout_bootstrap_large_vs_oracle():
    from oracle import pbhc_oracle_rollout as OR
    from pbhc_b200.rollout import RolloutStorage
    dev = torch.device("cuda:0")
    N = 70001
    gen = torch.Generator().manual_seed(3)
    rewards, values, to = torch.randn(N, generator=gen), torch.randn(N, 1, generator=gen) * 50, torch.rand(N, generator=gen) < 0.1
    st = RolloutStorage(N, 1, device=dev)
    st.register_key("rewards", shape=(1,), dtype=torch.float)
    st.bootstrap_rewards(rewards.to(dev), values.to(dev), to.to(dev), 0.99)
    assert np.array_equal(st.rewards[0].cpu().numpy(), OR.bootstrap_rewards(rewards.numpy(), values.numpy(), to.numpy(), 0.99))


def test_tracking_step_writes_observations_into_storage(tmp_path):
    """zero-copy store: the fused step writes its observation rows straight into the storage slot and update_key sees
    them in place; the rows equal those of an identical step writing its own buffers."""
    from pbhc_b200.rollout import RolloutStorage
    from benchkit.workload import Workload
    dev = torch.device("cuda:0")
    a, b = (Workload(257, dev, str(tmp_path / f"w{i}"), seed=11) for i in range(2))
    st = RolloutStorage(257, 3, device=dev)
    for gname, buf in b.ts.obs_buf_dict.items():
        st.register_key(gname, shape=(buf.shape[1],), dtype=torch.float)
    for step in range(3):
        b.ts.write_obs_into({gname: st.slot(gname) for gname in st.stored_keys})
        a.step()
        b.step()
        for gname in st.stored_keys:
            assert b.ts.obs_buf_dict[gname].data_ptr() == st.slot(gname).data_ptr()
            st.update_key(gname, b.ts.obs_buf_dict[gname])
            assert torch.equal(st.query_key(gname)[step], a.ts.obs_buf_dict[gname]), (gname, step)
        st.increment_step()
    with pytest.raises(Exception):
        b.ts.write_obs_into({st.stored_keys[0]: torch.zeros(3, 3, device=dev)})
