#!/usr/bin/env python
"""bench.py - tracked env-steps/s of the fused motion-tracking step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs E] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path over E environments per GPU: K2 forward kinematics on E poses
(`fk_batch(B=E, T=1)`), then the fused K1+K3 kernel (motion lookup + differences + termination + rewards + reset +
observations + history) and its scalar finalize kernel.  Workload = config A/B of SURVEY.md 8 (G1 23-DoF motion
tracking, obs=main, rewards=main, clip Horse-stance_pose) on synthetic simulator state, E = 65 536 envs per GPU (weak
scaling: envs are sharded, the only cross-rank exchange is the all-reduce of the 32-double statistics vector).
Rank 0 prints ONE JSON line (see README / DESIGN.md "Measurement").

`--impl reference` times the CPU oracle port of the reference's torch implementation (oracle/pbhc_oracle_env.py, pinned
bit-exact against the unmodified reference env) on the host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "tracked env-steps/s (motion lookup+FK+reward+obs)"
UNIT = "env-steps/s"
WORKLOADS = {
    "A": "G1 23-DoF motion_tracking step (obs=main, rewards=main, domain_rand=main incl. push_robots, clip Horse-stance_pose), synthetic sim state",
    "C": "G1 29-DoF general_tracking step (obs=obs_ppo_teacher, rewards=general_main, domain_rand=main incl. push_robots, 3-clip library, per-env motion ids, "
         "20-step future window), synthetic sim state",
}
WORKLOAD = WORKLOADS["A"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--envs", type=int, default=65536, help="environments per GPU")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-fk", action="store_true", help="leave the K2 launch out of the step")
    ap.add_argument("--no-push-robots", action="store_true", help="domain_rand.push_robots off (the shipped domain_rand/main.yaml has it on; the fixtures and round 1 ran without)")
    ap.add_argument("--cpu-envs", type=int, default=4096, help="envs in the CPU baseline's bounded sample")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline time budget")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-config-c", action="store_true", help="skip the config C object (general_tracking on the 1024-clip library) of the default line")
    ap.add_argument("--library-clips", type=int, default=1024, help="clips of the config C motion library (SURVEY 8d: 1024)")
    ap.add_argument("--e2e-chunks", type=int, default=4, help="env chunks of the host-buffer pipeline (HostStep)")
    ap.add_argument("--workload", default="A", choices=["A", "C"],
                    help="A: BASELINE configs[0..1] (23-DoF motion_tracking, the headline); C: configs[2] (29-DoF general_tracking)")
    a = ap.parse_args()
    global WORKLOAD
    if a.no_push_robots:  # the labels say what ran
        for k in WORKLOADS:
            WORKLOADS[k] = WORKLOADS[k].replace("domain_rand=main incl. push_robots", "domain_rand=main with push_robots off")
    WORKLOAD = WORKLOADS[a.workload]
    return a


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """polls NVML (SM clock, event reasons, power) every few ms while the measured passes run."""

    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index, period=0.004):
        super().__init__(daemon=True)
        self.period, self.samples, self.marks = period, [], {}
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001 - clocks are evidence, not a dependency of the measurement
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._halt.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, rs))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def mark(self, name):
        self.marks[name] = time.perf_counter()

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)

    def summary(self):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable: " + getattr(self, "err", "")}
        t0, t1 = self.marks.get("timed_start"), self.marks.get("timed_end")
        inside = [s for s in self.samples if t0 is not None and t0 <= s[0] <= t1]
        window = "timed region"
        if len(inside) < 3:  # a very short timed region: use every sample taken while the measured passes ran
            l0, l1 = self.marks.get("load_start", 0.0), self.marks.get("load_end", float("inf"))
            inside, window = [s for s in self.samples if l0 <= s[0] <= l1], "all measured passes (timed region < 3 samples)"
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no samples"}
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        reasons = [n for b, n in self.REASONS.items() if bits & b and n != "gpu_idle"]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(inside), "window": window}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's torch implementation, on the host cores (the one place outside tests/
# and smoke() that executes oracle/).
# ---------------------------------------------------------------------------------------------------------------
def cpu_reference(n_envs, steps, warmup, seconds=None, with_fk=True, seed=0, workload="A", push_robots=True):
    """Times `steps` env steps (or as many as fit in `seconds`, at least 3) of the CPU oracle on n_envs envs."""
    import torch
    from oracle import pbhc_oracle as O, pbhc_oracle_env as OE
    from benchkit import synth
    from benchkit.fixtures import env_config, tree_json, motion_npz

    def oracle_tree(tag):
        t = tree_json(tag)
        return O.make_tree(t["names"], t["parents"], t["offsets"], t["local_rot_wxyz"], t["num_bodies"], t["dof_axis"])

    def oracle_clip(tag, name):
        m = motion_npz(name)
        return O.load_clip(oracle_tree(tag), m["pose_aa"], m["root_trans_offset"], int(m["fps"]), m.get("contact_mask"))
    from pbhc_b200.spec import TrackSpec
    from benchkit.fixtures import CLIPS29
    from benchkit.workload import CASE, CASE_C, CLIP
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = TrackSpec.from_config(env_config(CASE_C if workload == "C" else CASE, push_robots=push_robots))
    tag, clips = ("29", list(CLIPS29)) if spec.general else ("23", [CLIP])
    tree = oracle_tree(tag)
    N, D, Jb, J = n_envs, spec.D, spec.Jb, spec.J
    tab = O.MotionTable([oracle_clip(tag, c) for c in clips], spec.Jb, motion_to_clip=[i % len(clips) for i in range(N)])
    gen = torch.Generator().manual_seed(seed)
    S = OE.new_state(spec, N)
    S["motion_ids"] = torch.arange(N)
    mlen = tab.length[S["motion_ids"]]
    S["motion_len"][:] = mlen
    S["motion_start_times"][:] = torch.rand(N, generator=gen) * mlen
    S["episode_length_buf"][:] = torch.randint(0, 50, (N,), generator=gen)
    t = (S["episode_length_buf"] + 2) * spec.dt + S["motion_start_times"]
    ref = O.get_motion_state(tab, S["motion_ids"], t, offset=S["env_origins"])
    body, root, dof, cf = synth.synth_sim_state(ref, Jb, gen, feet=tuple(spec.feet_ids))
    lim = torch.tensor(spec.torque_limit)
    inp = dict(body_state=body, root_state=root, dof_state=dof, contact_forces=cf, actions=synth.synth_actions(N, D, gen),
               torques=torch.maximum(torch.minimum(torch.randn(N, D, generator=gen) * 20.0, lim), -lim))
    pose = torch.randn(1, N, J, 3, generator=gen) * 0.3
    trans = root[:, :3].clone()[None]

    def one():
        if with_fk:
            O.fk(tree, pose, trans)
        inject = {"motion_start_times": torch.rand(N) * mlen}  # sample_time(), motion_lib_base.py:486-495
        OE.step(spec, tab, S, inp, inject=inject)

    for _ in range(warmup):
        one()
    done, t0 = 0, time.perf_counter()
    while done < steps:
        one()
        done += 1
        if seconds is not None and done >= 3 and time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    return dict(value=N * done / dt, ms_per_step=dt / done * 1e3, steps=done, cores=cores, n_envs=N,
                torch=torch.__version__)


def port_calibration():
    """the port against the unmodified reference env, timed in the build container (tools/calibrate_cpu_port.py)"""
    p = os.path.join(ROOT, "profiles", "r02_cpu_port_calibration.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return None


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # the arm runs the GPU arm's config: all a.envs envs per step (the oracle port steps 65536 envs in about a second on the
    # host cores); the time budget only guards against a very slow host
    r = cpu_reference(a.envs, a.steps, a.warmup, seconds=240.0, with_fk=not a.no_fk, workload=a.workload, push_robots=not a.no_push_robots)
    sample = (f"{r['n_envs']} envs per step (the whole {a.envs}-env workload), {r['steps']} timed steps after {a.warmup} warm-up; "
              "CPU oracle port of the reference torch path (oracle/pbhc_oracle_env.py, pinned bit-exact to the reference env), "
              f"torch {r['torch']} with {r['cores']} threads")
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": r["steps"],
            "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "envs_per_gpu": a.envs, "fk_in_step": not a.no_fk, "sample_envs": r["n_envs"]},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample,
                             "calibration": port_calibration()},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
def run_b200_arm(a):
    import torch
    import torch.distributed as dist
    from pbhc_b200 import _lib
    from benchkit.workload import Workload, step_bytes, fk_bytes
    _lib.lib()  # fails loudly when the CUDA library is missing: there is no CPU fallback
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {a.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, K, W, with_fk = a.envs, a.steps, max(a.warmup, 3), not a.no_fk
    from benchkit.workload import CASE, CASE_C
    wl = Workload(N, dev, tempfile.mkdtemp(prefix=f"pbhc_bench_r{rank}_"), seed=1000 + rank, case=CASE_C if a.workload == "C" else CASE, push_robots=not a.no_push_robots,
                  env_offset=rank * N)  # (the in-kernel random streams are keyed by the global env id)
    ts = wl.ts
    sampler = ClockSampler(torch.cuda.current_device() if os.environ.get("CUDA_VISIBLE_DEVICES") is None else local)
    sampler.start()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (eager), then capture the whole step - FK, fused kernel, reduction, all-reduce, finalize - in ONE graph
    for _ in range(W):
        wl.step(with_fk)
    torch.cuda.synchronize()
    graph, mode = None, "eager"
    if not a.no_graph:
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                wl.step(with_fk)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                wl.step(with_fk)
            mode = "cuda-graph replay"
            for _ in range(3):
                graph.replay()
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(f"[bench] CUDA-graph capture failed ({e!r}); timing eager launches", file=sys.stderr)
            graph, mode = None, "eager"
            torch.cuda.synchronize()
    run = graph.replay if graph is not None else (lambda: wl.step(with_fk))

    # ---- timed region: exactly K steps -------------------------------------------------------------------------
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.mark("load_start")
    barrier()
    sampler.mark("timed_start")
    e0.record()
    for _ in range(K):
        run()
    e1.record()
    barrier()
    sampler.mark("timed_end")
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    resets = float(ts.acc[_lib_index("A_RESET")].item()) / (N * world)

    # ---- per-kernel pass: the dominant kernel (pbhc_track_step = fused kernel + its 1-block reduction) and K2, eager,
    # CUDA events on the launching stream around each launch ----------------------------------------------------------
    Kk = min(K, 50)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(Kk)]
    torch.cuda.synchronize()
    for i in range(Kk):
        ev[i][0].record()
        if with_fk:
            wl.fk()
        ev[i][1].record()
        ts.launch()
        ev[i][2].record()
        ts.finalize()
        ev[i][3].record()
    torch.cuda.synchronize()
    fk_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / Kk
    track_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / Kk
    fin_ms = sum(e[2].elapsed_time(e[3]) for e in ev) / Kk

    track_launch = ts.launch_info()  # (what the headline workload launched - before the config C pass launches its own kernel)
    # ---- end to end: host buffers in, host results out, every step ----------------------------------------------------
    e2e = None
    if not a.no_e2e:
        e2e = e2e_pass(torch, dist, wl, min(K, 20), with_fk, world, dev, chunks=a.e2e_chunks)
    cfg_c = None
    if a.workload == "A" and world == 1 and not a.no_config_c:
        cfg_c = config_c_pass(torch, a, dev, peaks()[0])
    sampler.mark("load_end")
    sampler.stop()
    if rank != 0:
        return _finish(torch, dist, world, graph)

    peak, peak_src = peaks()
    bytes_step, reads, writes = step_bytes(wl.spec)
    achieved = bytes_step * N / (track_ms * 1e-3) / 1e9
    launches = wl.LAUNCHES_PER_STEP
    per_step = (launches["fk"] if with_fk else 0) + launches["track"] + launches["finalize"]
    line = {
        "metric": METRIC, "value": N * world * K / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "envs_per_gpu": N, "envs_total": N * world, "fk_in_step": with_fk, "launch": mode,
                   "l2": f"inputs larger than L2 ({bytes_step * N / 1e6:.0f} MB touched per step per GPU vs 126 MB L2)",
                   "parallelism": f"envs sharded over {world} GPU(s); one all-reduce of 64 doubles per step" if world > 1 else "single GPU",
                   "reset_fraction_last_step": resets, "track_launch": track_launch,
                   "exchange": (list(__import__("pbhc_b200.dist", fromlist=["peer_status"]).peer_status.values()) or ["library collective"]) if world > 1 else None},
        "roofline": {"bound": "hbm", "kernel": "pbhc_track_step: fused K1+K3 kernel (%s) + block-partials reduction" % track_launch["kernel"],
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic() if a.workload == "A" else None,
                     "peak_source": peak_src, "algorithmic_bytes_per_env_step": bytes_step, "kernel_ms": track_ms,
                     "envs_per_launch": N},
        "stages_ms": {"fk": fk_ms if with_fk else None, "track": track_ms, "finalize": fin_ms,
                      "fk_frac_of_hbm": (fk_bytes(wl.spec.J) * N / (fk_ms * 1e-3) / 1e9 / peak) if with_fk and fk_ms > 0 else None},
        "gpu_launches": per_step * K,
        "clocks": sampler.summary(),
    }
    if e2e is not None:
        line["e2e"] = e2e
    if cfg_c is not None:
        line["config_C"] = cfg_c
    if not a.no_cpu_baseline:
        r = cpu_reference(a.cpu_envs, 10 ** 9, 2, seconds=a.cpu_seconds, with_fk=with_fk, workload=a.workload, push_robots=not a.no_push_robots)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                "sample": f"{r['n_envs']} envs per step, {r['steps']} steps in ~{a.cpu_seconds:.0f} s, CPU oracle port of the "
                                          f"reference torch path (torch {r['torch']}, {r['cores']} threads), FK {'in' if with_fk else 'not in'} step",
                                "ms_per_step": r["ms_per_step"], "calibration": port_calibration()}
    print(json.dumps(line), flush=True)
    return _finish(torch, dist, world, graph)


def config_c_pass(torch, a, dev, peak):
    """BASELINE configs[2]: the G1 29-DoF general_tracking step (teacher obs, general_main rewards, 20-step future window)
    at 65 536 envs on the SURVEY 8d library - `--library-clips` 29-DoF clips of 150-600 frames with per-env motion ids, so
    that the motion table (~0.75 GB at 1024 clips) is served from HBM.  Same timing rules as the headline: CUDA-graph
    replay of the whole step, CUDA events, K steps after warm-up; roofline against SURVEY 8d's 36 165 B / env-step."""
    import time as _t
    from benchkit.workload import Workload, CASE_C, fk_bytes
    t0 = _t.perf_counter()
    wl = Workload(a.envs, dev, tempfile.mkdtemp(prefix="pbhc_bench_C_"), seed=2000, case=CASE_C, library_clips=a.library_clips, push_robots=not a.no_push_robots)
    load_s = _t.perf_counter() - t0
    for _ in range(3):
        wl.step(True)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        wl.step(True)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        wl.step(True)
    for _ in range(3):
        g.replay()
    K = min(a.steps, 50)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(10)]
    for p in ev:
        p[0].record()
        wl.ts.launch()
        p[1].record()
    torch.cuda.synchronize()
    kms = sum(p[0].elapsed_time(p[1]) for p in ev) / len(ev)
    bytes_c = 36165  # SURVEY.md 8d "Fused tracked env-step C" (table rows from HBM)
    rec = wl.lib._records
    out = {"workload": WORKLOADS["C"].replace("3-clip library", f"{wl.lib._desc.num_clips}-clip library"), "envs_per_gpu": a.envs, "steps": K,
           "value": a.envs / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "kernel_ms": kms,
           "roofline": {"bound": "hbm", "algorithmic_bytes_per_env_step": bytes_c, "achieved": bytes_c * a.envs / (kms * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "frac": bytes_c * a.envs / (kms * 1e-3) / 1e9 / peak},
           "motion_table": {"clips": int(wl.lib._desc.num_clips), "frames": int(rec.shape[0]), "mbytes": rec.numel() * 4 / 1e6,
                            "load_motions_s": load_s, "note": "built on the device: one FK launch + one pbhc_clips_prepare over all frames"},
           "track_launch": wl.ts.launch_info()}
    del g, wl
    torch.cuda.empty_cache()
    return out


def _finish(torch, dist, world, graph):
    """Leaves without tearing NCCL down: destroy_process_group() after a CUDA graph captured a collective can block
    forever (seen on the 2-GPU box); every rank drains its stream, meets at a barrier and exits."""
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)
    return 0


def _lib_index(name):
    return {"A_RESET": 25}[name]  # PBHC_A_RESET, include/pbhc_b200.h


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel from the committed ncu capture
    (profiles/track_kernel_traffic.json, written by tools/ncu_summary.py), or None."""
    p = os.path.join(ROOT, "profiles", "track_kernel_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("dram_bytes_per_launch")
    return None


def e2e_pass(torch, dist, wl, K, with_fk, world, dev, chunks=4):
    """The same step through the public host-buffer API (pbhc_b200.tracking_step.HostStep): every step copies the
    simulator state and actions from pinned host memory to the device, runs the step, and reads observations /
    rewards / reset flags back to pinned host memory (chunked so that copies and kernels overlap)."""
    from pbhc_b200.tracking_step import HostStep
    h = wl.host
    pin = lambda t: t.contiguous().pin_memory()
    host_in = {k: pin(h[k]) for k in ("body_state", "root_state", "dof_state", "contact_forces", "actions", "torques")}
    extra = None
    if with_fk:
        wl.fk()
        pos, mat, quat = wl.fk_out
        hb = wl.lib.mesh_parsers

        def fk_range(lo, hi):
            hb.fk(wl.pose_aa[lo:hi], wl.trans[lo:hi], out=(pos[lo:hi], mat[lo:hi], quat[lo:hi]))
        extra = (fk_range, [(pin(h["pose_aa"]), wl.pose_aa), (pin(h["trans"]), wl.trans)])
    hs = HostStep(wl.ts, wl.sim, wl.actions, wl.torques, wl.env_origins, wl.motion_ids, host_in, extra=extra, chunks=chunks)
    for _ in range(3):
        hs.step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        hs.step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return {"value": wl.N * world * K / (float(ms.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": hs.h2d_bytes * world,
            "d2h_bytes_per_step": hs.d2h_bytes * world, "steps": K, "ms_per_step": float(ms.item()) / K,
            "api": f"pbhc_b200.tracking_step.HostStep ({len(hs.ranges)} env chunks pipelined over 3 streams" + (", replayed as one CUDA graph)" if hs._graph is not None else ")")}


def main():
    a = parse_args()
    if a.impl == "reference":
        return run_reference_arm(a)
    return run_b200_arm(a)


if __name__ == "__main__":
    sys.exit(main())
