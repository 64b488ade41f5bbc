#!/usr/bin/env python
"""How the CPU port that bench.py times (oracle/pbhc_oracle_env.py) relates to the code it stands in for: times
`_post_physics_step` of the UNMODIFIED reference env (oracle/ref_env.py, needs /root/reference - build container only) and
the port's step on the same number of envs, the same synthetic simulator state and the same thread count, and writes
profiles/r02_cpu_port_calibration.json (bench.py copies it into cpu_baseline.calibration).

    python tools/calibrate_cpu_port.py [envs] [steps]
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_env as E, ref_loader as R  # noqa: E402
from benchkit import synth  # noqa: E402


def time_reference(n_envs, steps, warmup=2, seed=3):
    cfg = E.build_config(n_envs, R.example_motion("Horse-stance_pose"), obs="motion_tracking/main",
                         overrides={"domain_rand.push_robots": False})
    env, _ = E.make_env(cfg, env_cls="motion_tracking", seed=seed)
    sim = env.simulator
    gen = torch.Generator().manual_seed(seed)
    N, D, Jb = n_envs, env.num_dof, env.num_bodies
    env.motion_start_times[:] = torch.rand(N, generator=gen) * env.motion_len
    env.episode_length_buf[:] = torch.randint(0, 50, (N,), generator=gen)
    times = []
    for step in range(warmup + steps):
        env._pre_physics_step(synth.synth_actions(N, D, gen))
        env._physics_step()
        t = (env.episode_length_buf + 2) * env.dt + env.motion_start_times
        ref = env._motion_lib.get_motion_state(env.motion_ids, t, offset=env.env_origins)
        body, root, dof, cf = synth.synth_sim_state(ref, Jb, gen, feet=tuple(int(i) for i in env.feet_indices))
        sim.set_state(body, root, dof, cf)
        t0 = time.perf_counter()
        env._post_physics_step()
        if step >= warmup:
            times.append(time.perf_counter() - t0)
    return sum(times) / len(times) * 1e3


def time_port(n_envs, steps, warmup=2):
    sys.argv = [sys.argv[0]]
    import bench
    r = bench.cpu_reference(n_envs, steps, warmup, with_fk=False, push_robots=False)  # (like the reference env above: domain_rand.push_robots off on both sides)
    return r["ms_per_step"], r["cores"]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    os.chdir(R.REF_ROOT)
    torch.set_num_threads(os.cpu_count() or 1)
    ref_ms = time_reference(n, steps)
    os.chdir(ROOT)
    port_ms, cores = time_port(n, steps)
    out = {"what": "ms per _post_physics_step (no FK) at the same env count, synthetic simulator state and thread count, measured in the build container",
           "envs": n, "threads": cores, "timed_steps": steps, "reference_unmodified_ms": round(ref_ms, 2), "port_ms": round(port_ms, 2),
           "port_speedup_over_reference": round(ref_ms / port_ms, 3), "torch": torch.__version__,
           "note": "the port is faster than the code it stands in for, so a GPU / port ratio understates the GPU / reference ratio by this factor"}
    with open(os.path.join(ROOT, "profiles", "r02_cpu_port_calibration.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
