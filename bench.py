#!/usr/bin/env python
"""bench.py -- headline measurement of the B200 PPO rollout-and-update engine.

Metric (BASELINE.json): CartPole env-steps/sec incl. policy inference; PPO update samples/sec.
Workload at N GPUs: BASELINE.json configs[2] per GPU -- 65,536 envs x 128-step horizon, default
6-64-32-1 MLPs (weak scaling: every rank owns its own 65,536 envs; gradients all-reduced over NCCL).

One "step" = one pass of the hot path over one batch: fused rollout (K1) + returns scan (K2) [timed as the
rollout leg -> `value`, env-steps/s], then the PPO update (K3-K5, `--epochs` full-batch epochs) [timed as
the update leg -> `update.value`, sample-epochs/s].  Everything is timed on the device with CUDA events on
the engine's stream, max over ranks.  `e2e` repeats the rollout leg through the C ABI with HOST buffers:
weights H2D from pinned memory in, the whole trajectory batch D2H into pinned memory out.

Further legs inside the same JSON line: `update` (default-net PPO update on the rollout's batch), `update_wide`
(BASELINE configs[4]), `configs1` (BASELINE configs[1]: 3 env nodes x 7 episodes, N = 1 only), `c4` (BASELINE configs[3]:
1,048,576 envs over the N GPUs, strong scaling) and, at N > 1, `multi_gpu_check` (weights bit-identical across ranks,
gradients vs one GPU on the concatenated batch, peer-memory exchange vs ncclAllReduce; a violation makes the run exit 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision f64|f32|mixed]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PHYS = 12000      # 300 RK stages x 40 flop per env-step
FLOP_ACTOR = 4928      # 2 x (6*64 + 64*32 + 32*1)
FLOP_UPDATE = 28032    # per sample-epoch, default MLP (fwd 9856 + dW 9856 + dX 8320)
FLOP_UPDATE_WIDE = 801792  # per sample-epoch, 2x256 MLPs (fwd 269,312 + dW 269,312 + dX 263,168)
BYTES_STEP = 41        # obs 24 + a 4 + logp 4 + r 4 + V 4 + flag 1 written once per env-step
BYTES_UPDATE = 36      # obs 24 + a 4 + old logp 4 + R 4 read per sample-epoch


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="f64", choices=["f64", "f32", "mixed"],
                    help="f64: the parity-gated headline; mixed: double state + RK accumulators, f32 Diff (teacher-forced gate 1e-5); f32: drift reported only")
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--horizon", type=int, default=128)
    ap.add_argument("--epochs", type=int, default=4, help="full-batch PPO epochs per timed update leg")
    ap.add_argument("--chunks", type=int, default=24, help="windows per rollout in the e2e leg (copy/compute overlap): 24 for MemoryNN's 37-B "
                    "record, where the copy engine is the slower side and short windows shorten the exposed tail (6.61 vs 6.93 ms at 8, "
                    "tools/chunk_probe.py); the compact record always uses 8")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-update", action="store_true")
    ap.add_argument("--no-wide", action="store_true", help="skip the configs[4] leg (2x256 MLP update sweep on tensor cores)")
    ap.add_argument("--no-wide-whole", action="store_true", help="skip the whole-4M-batch-on-one-GPU variant of the configs[4] leg")
    ap.add_argument("--wide-samples", type=int, default=524288, help="samples per GPU of the configs[4] leg (4M over 8 GPUs)")
    ap.add_argument("--no-c4", action="store_true", help="skip the configs[3] leg (1,048,576 envs over the N GPUs)")
    ap.add_argument("--c4-envs", type=int, default=1048576)
    ap.add_argument("--no-configs1", action="store_true", help="skip the configs[1] leg (3 envs x 700 steps, K = 80; N = 1 only)")
    ap.add_argument("--no-multi-check", action="store_true", help="skip the multi-GPU parity self-check (N > 1)")
    ap.add_argument("--no-mpi-baseline", action="store_true", help="reference arm: skip the mpirun-shaped runs (b2_mpirun -n 2 / -n 4)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's OWN CPU path (oracle/_ref/ref_harness, built from the
# unmodified sources) timed on this box's host cores.  Its learner is a single serial process and the
# libtorch intra-op pool slows it down, so it is run with 1 thread and cores = 1.
# ------------------------------------------------------------------------------------------------
def run_reference(updates, K=80, threads=1, nenv=1, episodes=20):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if not os.path.exists(exe):
        return None
    t0 = time.time()
    out = subprocess.run([exe, "bench", str(nenv), str(episodes), str(K), str(threads), str(updates)],
                         capture_output=True, text=True, timeout=1800)
    if out.returncode != 0:
        return None
    r = json.loads(out.stdout.strip().splitlines()[-1])
    r["wall_s"] = time.time() - t0
    return r


def run_reference_mpi(nranks, episodes, threads=1, timeout=600):
    """The reference's OWN launch shape (README.md:82-92: mpirun -n N ./cpp-rl-training config): its unmodified
    src/main.cpp + src/ppo_nn.cpp + libtorch CPU (oracle/_ref/ref_main_mpi) as N OS processes -- rank 0 the learner, the
    others one CartPole each -- over tools/mpi_shim (shared-memory rings standing in for the MPI this image lacks), so
    every env step pays the message round trip of src/main.cpp:47-48,104-131,186,230-234.  Runs `episodes` episodes
    (>= 3 update cycles), stdout discarded; env-steps/s = 100 x episode lines in proc*_log.txt / wall."""
    import glob
    import shutil
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_main_mpi")
    launcher = os.path.join(ROOT, "tools", "bin", "b2_mpirun")
    if not (os.path.exists(exe) and os.path.exists(launcher)):
        return None
    d = tempfile.mkdtemp(prefix="b2_refmpi_")
    try:
        with open(os.path.join(d, "config"), "w") as f:   # the reference's 6-line format (config:1-6), Nepisodes shortened
            f.write(f"1 = control_vars\n6 = obs_vars\n{episodes} = Nepisodes\n100 = N_timestep\n1000000 = Max_timestep\n2000 = updateTimestep\n")
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), MKL_NUM_THREADS=str(threads))
        t0 = time.time()
        out = subprocess.run([launcher, "-n", str(nranks), "-q", exe, "./config"], cwd=d, env=env, capture_output=True,
                             text=True, timeout=timeout)
        wall = time.time() - t0
        lines = 0
        for fn in glob.glob(os.path.join(d, "proc*_log.txt")):
            lines += sum(1 for _ in open(fn))
        if lines == 0:
            return None
        nenv = nranks - 1
        per_update = 100 * nenv * -(-2000 // (100 * nenv))      # whole episodes on every env (src/main.cpp:264-307)
        return {"value": 100.0 * lines / wall, "unit": "env-steps/s", "ranks": nranks, "env_ranks": nenv,
                "libtorch_threads": threads, "episodes": lines, "env_steps": 100 * lines,
                "updates": (100 * lines) // per_update, "wall_s": wall, "rc": out.returncode,
                "how": f"tools/bin/b2_mpirun -n {nranks} -q oracle/_ref/ref_main_mpi config (unmodified src/main.cpp + "
                       f"src/ppo_nn.cpp, libtorch CPU, shared-memory MPI shim), rollout + PPO::update(K=80) in one wall clock"}
    except Exception as e:   # noqa: BLE001 -- a baseline that cannot run is reported, not fatal
        return {"unavailable": repr(e)[:200]}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def cpu_port_baseline(seconds=6.0):
    """Fallback when the reference binary is absent: the C oracle port (OpenMP over envs)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import helpers as H
    params = H.gold("init_params.f32")
    N, T = 256, 100
    t0 = time.time()
    n = 0
    while time.time() - t0 < seconds:
        H.oracle_rollout(params, N, T, seed=1)
        n += N * T
    dt = time.time() - t0
    return {"value": n / dt, "unit": "env-steps/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{n} env-steps of orc_rollout ({N} envs x {T} steps per call, OpenMP)"}


def reference_arm(args, rank):
    if rank != 0:
        return
    per_update = 2000
    ups = max(1, args.steps + args.warmup)
    r = run_reference(ups)
    if r is None:
        b = cpu_port_baseline()
        v, kind, cores, sample = b["value"], "port", b["cores"], b["sample"]
        upd = None
    else:
        v = r["env_steps"] / r["rollout_s"]
        upd = r["sample_epochs"] / r["update_s"]
        kind, cores = "reference", 1
        sample = (f"{r['updates']} cycles of (1 env x 20 episodes x 100 steps rollout + PPO::update K=80), "
                  f"in-process replay of src/main.cpp's protocol without MPI latency, stdout discarded, "
                  f"{r['torch_threads']} libtorch thread")
    # BASELINE configs[1] (results/carpole3node: 1 learner + 3 env nodes) in process, and both configs as the reference
    # launches them (N OS processes exchanging a message per env step)
    c2 = run_reference(max(2, ups // 2), nenv=3, episodes=7) if r is not None else None
    configs1 = None
    if c2:
        configs1 = {"value": c2["env_steps"] / c2["rollout_s"], "unit": "env-steps/s",
                    "update": {"value": c2["sample_epochs"] / c2["update_s"], "unit": "sample-epochs/s"},
                    "combined_env_steps_per_s": c2["env_steps"] / (c2["rollout_s"] + c2["update_s"]),
                    "sample": f"{c2['updates']} cycles of (3 envs x 7 episodes x 100 steps + PPO::update K=80 on 2100 records), in process"}
    mpi = None
    if not args.no_mpi_baseline:
        mpi = {"n2": run_reference_mpi(2, 100), "n4": run_reference_mpi(4, 105)}
    line = {"impl": "reference", "metric": "cartpole_env_steps_per_sec_incl_policy_inference", "value": v,
            "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * per_update / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "reference default config (config:1-6): 1 learner + 1 env, 2000 env-steps per update",
                       "n_envs": 1, "horizon": 2000},
            "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": kind, "sample": sample},
            "update": {"value": upd, "unit": "sample-epochs/s"},
            "configs1": configs1, "cpu_baseline_mpi": mpi,
            "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.p, self.t0 = index, [], None, None

    def start(self):
        """Started before the warm-up: nvidia-smi needs ~0.3 s before its first row, longer than a short timed region."""
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def mark(self):
        """The timed region starts here: only rows sampled from now until stop() are reported."""
        self.t0 = time.time()

    def stop(self):
        t1 = time.time()
        if self.p:
            self.p.terminate()
        t0 = self.t0 if self.t0 is not None else 0.0
        rows = [r for t, r in self.rows if t0 <= t <= t1]
        window = "timed region"
        if not rows:   # a region shorter than one sampling period: the same workload's warm-up steps just before it
            rows = [r for t, r in self.rows if t <= t1][-8:]
            window = "warm-up steps before the timed region (region shorter than one sample)"
        sm = sorted(int(float(r[0])) for r in rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        # the kernel keeps the SMs busy for most samples; take the median of the upper half as "under load"
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": (load[len(load) // 2] if load else None), "sm_max_mhz": (max(mx) if mx else None),
                "reasons": reasons, "samples": len(sm), "window": window}


def _ncu_traffic(key):
    """(bytes per launch, where it came from): dram__bytes_read.sum + dram__bytes_write.sum of the named kernel from the
    `ncu --set full` capture of THIS round, written into profiles/NCU_TRAFFIC.json by profiles/ncu_traffic.py (the
    script that reads the .ncu-rep); null when the file has no entry."""
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "NCU_TRAFFIC.json")))
        e = tj.get(key)
        if isinstance(e, dict):
            return e.get("bytes"), f"profiles/NCU_TRAFFIC.json <- {e.get('source')} ({tj.get('_captured')})"
        return e, "profiles/NCU_TRAFFIC.json"
    except Exception:
        return None, None


def _rel(a, b):
    import numpy as np
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def multi_gpu_check(pkg, capi, np, torch, dist, rank, local, world):
    """Parity of the multi-GPU path, inside the bench line (N > 1), BEFORE anything is timed.  Same shapes as
    tests/test_gpu_multi.py, which a 1-GPU test box skips.  The reference computes ONE batch mean over the merged memory
    of all env ranks (src/main.cpp:290-303, src/ppo_nn.cpp:302); sharded, that is: identical trajectories for any sharding,
    gradients equal to one GPU's on the concatenated batch, weights bit-identical on every rank.
      default nets : 515 envs (NOT divisible by N: uneven shards, global-count denominators) x 100 steps, K = 3
      wide nets    : 2 x 256, 2048 samples, K = 2, M = 2 (every minibatch sharded across the ranks)
    Both through the fused NVLink peer-memory exchange (default) and through ncclAllReduce (B2_COMM_P2P=0)."""
    from reinforcement_learning_cpp_parallel_b200 import sharding
    NG, T, K = 515, 100, 3
    res = {"world": world}

    def gather_equal(vec):
        t = torch.from_numpy(np.ascontiguousarray(vec)).cuda()
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        return all(bool(torch.equal(parts[0].view(torch.int32), q.view(torch.int32))) for q in parts[1:])

    def sharded(p2p):
        old = os.environ.get("B2_COMM_P2P")
        os.environ["B2_COMM_P2P"] = "1" if p2p else "0"       # read by b2_comm_init
        try:
            n_local, off = sharding.shard_envs(NG, rank, world)
            E = pkg.Engine(capi.default_config(n_envs=n_local, horizon=T, seed=21, env_offset=off, k_epochs=K, device=local))
            E.init_params(4)
            E.comm_init(rank, world, sharding.broadcast_comm_id(dist, pkg.Engine.comm_unique_id))
            mode = E.comm_mode()
            E.rollout(T)
            traj = E.traj_download(T, fields=("obs", "action", "reward", "flags"))
            st = E.update_resident()
            out = (E.get_grads(), E.get_params(), st.loss, mode, traj, off, n_local)
            E.close()
            return out
        finally:
            if old is None:
                os.environ.pop("B2_COMM_P2P", None)
            else:
                os.environ["B2_COMM_P2P"] = old

    g_p, w_p, loss_p, mode_p, traj, off, n_local = sharded(True)
    g_n, w_n, loss_n, mode_n, _, _, _ = sharded(False)
    res["comm_mode_default"] = mode_p
    res["weights_bit_identical_across_ranks"] = bool(gather_equal(w_p) and gather_equal(w_n))
    res["p2p_vs_nccl_relerr"] = max(_rel(w_p, w_n), _rel(g_p, g_n)) if mode_p == 2 else None
    # one GPU holding every env (every rank computes it on its own GPU: no communication, same answer)
    E = pkg.Engine(capi.default_config(n_envs=NG, horizon=T, seed=21, k_epochs=K, device=local))
    E.init_params(4)
    E.rollout(T)
    full = E.traj_download(T, fields=("obs", "action", "reward", "flags"))
    st1 = E.update_resident()
    g1, w1 = E.get_grads(), E.get_params()
    E.close()
    traj_ok = all(np.array_equal(traj[k], full[k][:, off:off + n_local]) for k in traj)
    res["trajectories_identical_to_single_gpu"] = bool(traj_ok)
    res["grad_relerr_vs_single_gpu"] = max(_rel(g_p, g1), _rel(g_n, g1))
    res["weights_relerr_vs_single_gpu"] = max(_rel(w_p, w1), _rel(w_n, w1))
    res["loss_relerr_vs_single_gpu"] = abs(loss_p - st1.loss) / max(abs(st1.loss), 1e-30)
    res["uneven_shards"] = [int(sharding.shard_envs(NG, r, world)[0]) for r in range(world)]

    # ---- wide nets: each rank owns 1/N of every minibatch
    WB, WK, WM = 2048, 2, 2
    rng = np.random.default_rng(99)
    th = 0.3 * rng.standard_normal(WB)
    obs = np.stack([rng.standard_normal(WB), rng.standard_normal(WB), rng.standard_normal(WB), th, np.cos(th), np.sin(th)],
                   1).astype(np.float32)
    eps = rng.standard_normal(WB).astype(np.float32)
    reward = (rng.random(WB) < 0.9).astype(np.float32)
    E = pkg.Engine(capi.default_config(n_envs=WB // world, horizon=1, hidden1=256, hidden2=256, k_epochs=WK, n_minibatches=WM, device=local))
    E.init_params(4)
    E.comm_init(rank, world, sharding.broadcast_comm_id(dist, pkg.Engine.comm_unique_id))
    act, lp, _ = E.act(obs, eps)
    per = WB // WM
    part = per // world
    sel = np.concatenate([np.arange(m * per + rank * part, m * per + (rank + 1) * part) for m in range(WM)])
    stw = E.update(obs[sel], act[sel], lp[sel], reward[sel], np.ones(sel.size, np.uint8))
    gw, ww = E.get_grads(), E.get_params()
    E.close()
    E = pkg.Engine(capi.default_config(n_envs=WB, horizon=1, hidden1=256, hidden2=256, k_epochs=WK, n_minibatches=WM, device=local))
    E.init_params(4)
    a1, l1, _ = E.act(obs, eps)
    st1w = E.update(obs, a1, l1, reward, np.ones(WB, np.uint8))
    g1w, w1w = E.get_grads(), E.get_params()
    E.close()
    res["wide"] = {"weights_bit_identical_across_ranks": bool(gather_equal(ww)),
                   "grad_relerr_vs_single_gpu": _rel(gw, g1w), "weights_relerr_vs_single_gpu": _rel(ww, w1w),
                   "loss_relerr_vs_single_gpu": abs(stw.loss - st1w.loss) / max(abs(st1w.loss), 1e-30)}
    ok = (res["weights_bit_identical_across_ranks"] and traj_ok and res["grad_relerr_vs_single_gpu"] < 1e-5
          and res["weights_relerr_vs_single_gpu"] < 1e-6 and res["loss_relerr_vs_single_gpu"] < 1e-5
          and (res["p2p_vs_nccl_relerr"] is None or res["p2p_vs_nccl_relerr"] < 1e-6)
          and res["wide"]["weights_bit_identical_across_ranks"] and res["wide"]["grad_relerr_vs_single_gpu"] < 2e-5
          and res["wide"]["weights_relerr_vs_single_gpu"] < 1e-6 and res["wide"]["loss_relerr_vs_single_gpu"] < 1e-4)
    # every rank must agree on the verdict (a rank-local failure fails the run everywhere)
    t = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    res["ok"] = bool(t.item() > 0.5)
    return res


def c4_leg(args, pkg, capi, np, rank, local, world, dist, barrier, max_over_ranks, peak64, peak32):
    """BASELINE configs[3]: 1,048,576 envs sharded over the N GPUs (524,288 / 262,144 / 131,072 per GPU at N = 2 / 4 / 8;
    all of them on one GPU at N = 1, the strong-scaling anchor) x 128 steps, f64 rollout + 4 full-batch epochs of the
    default-net update with the gradient exchange of DESIGN.md s5.  STRONG scaling: the global batch is fixed."""
    NG, T, K = args.c4_envs, args.horizon, args.epochs
    if NG % world:
        return {"skipped": f"{NG} envs do not divide over {world} GPUs"}
    n_local = NG // world
    E = pkg.Engine(capi.default_config(n_envs=n_local, horizon=T, precision=pkg.PREC_F64, seed=0x5EED,
                                       env_offset=rank * n_local, k_epochs=K, device=local))
    E.init_params(0)
    if world > 1:
        uid = [pkg.Engine.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        E.comm_init(rank, world, uid[0])
    E.rollout(T); E.update_resident()          # warm-up (the 1M-env buffers are touched once)
    barrier(E)
    reps = max(2, min(args.steps, 3))
    l0 = E.kernel_launches()
    rms, ums = [], []
    for _ in range(reps):
        E.timer_start(); E.rollout(T); rms.append(E.timer_stop_ms())
        E.timer_start(); st = E.update_resident(); ums.append(E.timer_stop_ms())
    barrier(E)
    launches = E.kernel_launches() - l0
    rt, ut = max_over_ranks(sum(rms)), max_over_ranks(sum(ums))
    value = NG * T * reps / (rt * 1e-3)
    upd = NG * T * K * reps / (ut * 1e-3)
    mode = E.comm_mode()
    upath = E.last_update_path()
    E.close()
    mine_r, mine_u = sum(rms) / reps, sum(ums) / reps
    ach = FLOP_PHYS * n_local * T / (mine_r * 1e-3) / 1e12
    uach = FLOP_UPDATE * n_local * T * K / (mine_u * 1e-3) / 1e12
    out = {"value": value, "unit": "env-steps/s", "ms_per_rollout": rt / reps, "scaling": "strong",
           "config": {"workload": "BASELINE configs[3]: batched CartPole, 1,048,576 envs sharded over the GPUs x 128-step horizon, "
                                  "default MLP, gradient exchange per optimizer step", "n_envs_global": NG,
                      "n_envs_per_gpu": n_local, "horizon": T, "precision": "f64", "epochs": K},
           "roofline": {"kernel": "k_rollout<double>", "bound": "fp64-fma", "achieved": ach, "peak": peak64, "unit": "TFLOP/s",
                        "frac": ach / peak64 if peak64 else None},
           "update": {"value": upd, "unit": "sample-epochs/s", "ms_per_update": ut / reps,
                      "kernel_path": {0: "k_ppo_grad (fp32-fma)", 1: "k_ppo_grad_tc (bf16x3 mma.sync)"}.get(upath),
                      "roofline": {"achieved": uach, "unit": "TFLOP/s", "fp32_fma_peak": peak32,
                                   "frac_of_fp32_fma_peak": uach / peak32 if peak32 else None,
                                   "note": "see update.roofline of the main line for the tensor denominators"}},
           "gpu_launches": int(launches), "comm_mode": mode, "last_loss": st.loss}
    # strong-scaling efficiency against the N = 1 run of the same leg ON THIS BOX (the driver runs N = 1, 2, 4, 8 back to
    # back): the N = 1 run leaves its numbers in a scratch file, later runs read it
    anchor = os.path.join(os.environ.get("TMPDIR", "/tmp"), f"b2_c4_anchor_{NG}x{T}x{K}.json")
    if rank == 0:
        try:
            if world == 1:
                json.dump({"value": value, "update": upd, "when": time.time()}, open(anchor, "w"))
            elif os.path.exists(anchor):
                a = json.load(open(anchor))
                if time.time() - a["when"] < 6 * 3600:
                    out["strong_scaling_efficiency_vs_n1"] = {"rollout": value / (world * a["value"]),
                                                              "update": upd / (world * a["update"]),
                                                              "n1_value": a["value"], "n1_update": a["update"],
                                                              "n1_age_s": time.time() - a["when"], "source": anchor}
        except Exception:
            pass
    return out


def configs1_leg(args, pkg, capi, np, local):
    """BASELINE configs[1]: the results/carpole3node shape (1 learner + 3 env nodes) on ONE GPU: 3 envs x 7 episodes x
    100 steps per update (the reference's trigger rounds 2000 up to whole episodes on every env: 2,100 records,
    src/main.cpp:264-307), K = 80 full-batch epochs.  Latency-bound on a GPU (3 lanes busy); reported beside the
    reference arm's `configs1` / `cpu_baseline_mpi.n4`, not as a throughput claim."""
    N, T, K = 3, 700, 80
    E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, k_epochs=K, seed=0x5EED, device=local))
    E.init_params(0)
    E.train_iteration(T)
    reps = 5
    rms, ums = [], []
    l0 = E.kernel_launches()
    t0 = time.perf_counter()
    for _ in range(reps):
        E.timer_start(); E.rollout(T); rms.append(E.timer_stop_ms())
        E.timer_start(); E.update_resident(); ums.append(E.timer_stop_ms())
    wall = time.perf_counter() - t0
    launches = E.kernel_launches() - l0
    E.close()
    return {"value": N * T * reps / (sum(rms) * 1e-3), "unit": "env-steps/s",
            "update": {"value": N * T * K * reps / (sum(ums) * 1e-3), "unit": "sample-epochs/s"},
            "combined_env_steps_per_s": N * T * reps / wall, "ms_per_rollout": sum(rms) / reps, "ms_per_update": sum(ums) / reps,
            "gpu_launches": int(launches),
            "config": {"workload": "BASELINE configs[1]: 1 learner + 3 env nodes (results/carpole3node) on one GPU",
                       "n_envs": N, "horizon": T, "epochs": K, "records_per_update": N * T}}


def _nvlink_kib(local):
    """NVML's NVLink data counters of GPU `local` (payload KiB sent / received over all links since the driver loaded), or
    None where NVML or the counters are unavailable.  Sampled around the timed wide updates at world > 1: the measured
    bytes per optimizer step of the peer-memory gradient exchange (DESIGN.md s5 states them as arithmetic)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        ALL = 0xFFFFFFFF      # scopeId: the sum over the GPU's links
        vals = pynvml.nvmlDeviceGetFieldValues(h, [(pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, ALL),
                                                   (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, ALL)])
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                return "nvmlReturn %d" % v.nvmlReturn
            out.append(int(v.value.ullVal))
        return tuple(out)
    except Exception as e:
        return "%s: %s" % (type(e).__name__, e)


def wide_leg(args, pkg, capi, np, rank, local, world, dist, barrier, max_over_ranks, bf16_peak, bf16_src):
    """BASELINE configs[4]: PPO update sweep, 2x256 MLPs, 4M-sample batch over 8 GPUs = 524,288 samples per GPU,
    32 minibatches x 10 epochs; weak scaling (every rank owns its own 524,288 samples, gradients all-reduced)."""
    B, T, M, K = args.wide_samples, 128, 32, 10
    N = B // T
    E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, hidden1=256, hidden2=256, k_epochs=K, n_minibatches=M,
                                       seed=0x5EED, device=local))
    E.init_params(0)
    if world > 1:
        uid = [pkg.Engine.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        E.comm_init(rank, world, uid[0])
    comm_mode = {0: "single GPU", 1: "ncclAllReduce + Adam launch", 2: "NVLink peer-memory exchange fused into reduce + Adam"}[E.comm_mode()]
    rng = np.random.default_rng(1234 + rank)
    th = 0.3 * rng.standard_normal(B)
    obs = capi.pinned_empty((B, 6), np.float32)
    obs[:] = np.stack([rng.standard_normal(B), rng.standard_normal(B), rng.standard_normal(B), th, np.cos(th), np.sin(th)], 1)
    a0, l0, _ = E.act(obs)                       # actions = mu + 0.5 eps from the initial policy
    act, lp = capi.pinned_empty((B,), np.float32), capi.pinned_empty((B,), np.float32)
    act[:], lp[:] = a0, l0
    reward = capi.pinned_empty((B,), np.float32)
    reward[:] = rng.random(B) < 0.9
    flags = capi.pinned_empty((T, N), np.uint8)
    flags[:] = 0
    flags[99] = 2                                # TIMEUP after 100 steps (config: N_timestep)
    flags[T - 1] = 4                             # horizon cut
    fl = flags.reshape(-1)
    w0 = E.get_params()
    st = E.update(obs, act, lp, reward, fl, n_envs=N, horizon=T)       # upload + warm-up 1
    for _ in range(2):
        E.set_params(w0); E.update_resident()
    nv0 = _nvlink_kib(local) if world > 1 else None   # (before the barrier: NVML start-up differs by tens of ms between ranks)
    barrier(E)
    ms = []
    l_before = E.kernel_launches()
    for _ in range(args.steps):
        E.timer_start()
        st = E.update_resident()                 # returns scan + K x M optimizer steps on the resident batch
        ms.append(E.timer_stop_ms())
    launches = E.kernel_launches() - l_before
    barrier(E)
    nv1 = _nvlink_kib(local) if world > 1 else None
    nvlink = None
    if isinstance(nv0, str) or isinstance(nv1, str):
        nvlink = {"unavailable": str(nv0 if isinstance(nv0, str) else nv1)}
    elif nv0 and nv1:
        osteps = args.steps * K * M
        nvlink = {"tx_bytes_per_optimizer_step": 1024.0 * (nv1[0] - nv0[0]) / osteps,
                  "rx_bytes_per_optimizer_step": 1024.0 * (nv1[1] - nv0[1]) / osteps, "optimizer_steps": osteps,
                  "source": "NVML NVLINK_THROUGHPUT_DATA_TX / _RX of this rank's GPU, sampled around the timed updates (rank 0)",
                  "arithmetic": {"scatter_shape": 2 * (world - 1) / world * 8 * 135682, "all_to_all": (world - 1) * 8 * 135682,
                                 "note": "payload bytes per rank and step of the {value, step} pairs, P = 135,682 parameters + 3 sums"}}
    tot = max_over_ranks(sum(ms))
    value = B * K * args.steps * world / (tot * 1e-3)
    # e2e: the ppo_nn.h-shaped call, PPO::update(MemoryNN) on HOST buffers (pinned): H2D of the batch inside the timed region
    barrier(E)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st2 = E.update(obs, act, lp, reward, fl, n_envs=N, horizon=T)
    barrier(E)
    dt = max_over_ranks(time.perf_counter() - t0)
    h2d = obs.nbytes + act.nbytes + lp.nbytes + reward.nbytes + fl.nbytes
    mine = sum(ms) / len(ms)
    ach = FLOP_UPDATE_WIDE * B * K / (mine * 1e-3) / 1e12
    E.close()
    # the same sweep with the WHOLE 4M-sample batch on one GPU (131,072-sample minibatches: 2,048 GEMM tiles per launch
    # instead of 256): shows the kernels away from the launch-latency regime of the 8-GPU share.  One GPU only.
    whole = None
    if world == 1 and not args.no_wide_whole:
        Bw = 8 * B
        Nw = Bw // T
        Ew = pkg.Engine(capi.default_config(n_envs=Nw, horizon=T, hidden1=256, hidden2=256, k_epochs=K, n_minibatches=M,
                                            seed=0x5EED, device=local))
        Ew.init_params(0)
        rep8 = lambda a: np.ascontiguousarray(np.tile(a, (8,) + (1,) * (a.ndim - 1)))
        obs_w = rep8(obs)
        a_w, l_w, _ = Ew.act(obs_w)
        fl_w = np.zeros((T, Nw), np.uint8)
        fl_w[99] = 2
        fl_w[T - 1] = 4
        Ew.update(obs_w, a_w, l_w, rep8(reward), fl_w.reshape(-1), n_envs=Nw, horizon=T)
        Ew.set_params(w0)
        msw = []
        for _ in range(max(2, min(args.steps, 3))):
            Ew.timer_start()
            Ew.update_resident()
            msw.append(Ew.timer_stop_ms())
        mw = sum(msw) / len(msw)
        aw = FLOP_UPDATE_WIDE * Bw * K / (mw * 1e-3) / 1e12
        whole = {"value": Bw * K / (mw * 1e-3), "unit": "sample-epochs/s", "ms_per_update": mw, "samples": Bw,
                 "minibatch": Bw // M, "roofline": {"bound": "tensor", "achieved": aw, "peak": bf16_peak / 3.0,
                                                    "unit": "TFLOP/s", "frac": aw / (bf16_peak / 3.0) if bf16_peak else None}}
        Ew.close()
    return {"value": value, "unit": "sample-epochs/s", "ms_per_update": tot / args.steps,
            "config": {"workload": "BASELINE configs[4]: PPO update sweep, 2x256 MLP, 4M-sample batch / 8 GPUs = 524,288 "
                                   "samples per GPU, 32 minibatches x 10 epochs", "samples_per_gpu": B, "minibatches": M,
                       "epochs": K, "mlp": "6-256-256-1 x2", "operands": "bf16x3 (hi/mid split, fp32 accumulate in TMEM)"},
            "gpu_launches": int(launches),
            "e2e": {"value": B * K * args.steps * world / dt, "unit": "sample-epochs/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": 4 * 8 * K * M, "api": "b2_update on pinned host buffers (PPO::update(MemoryNN))"},
            "roofline": {"kernel": "whole optimizer step: k_wide_fh (forward GEMM + loss head + dX GEMM), k_gemm_mn2 (dW2 on CTA pairs + dw1) on tcgen05; "
                                   "k_wide_l1r (layer 1 + the reduce and Adam of the step before; N > 1: k_wide_l1 and k_wide_reduce) on FP32; traffic = the longest kernel of the chain (k_wide_fh)",
                         "bound": "tensor", "achieved": ach, "peak": bf16_peak / 3.0,
                         "unit": "TFLOP/s", "frac": ach / (bf16_peak / 3.0) if bf16_peak else None,
                         "peak_source": bf16_src + " / 3 (every fp32 product is three bf16 MMAs)",
                         "issued_bf16_tflops": 3 * ach, "flop_per_sample_epoch": FLOP_UPDATE_WIDE,
                         "traffic": _ncu_traffic("wide_dominant_kernel")[0], "traffic_source": _ncu_traffic("wide_dominant_kernel")[1],
                         "comm_mode": comm_mode, "nvlink": nvlink},
            "whole_batch_one_gpu": whole, "last_loss": st.loss}


def _bind_to_gpu_numa_node(local):
    """Pin this process to the CPUs NVML reports as local to GPU `local`: first-touch then places the pinned host buffers
    of the e2e leg on the GPU's own NUMA node (on a two-socket box half of the ranks would otherwise write across the
    socket link).  Host-side placement only; returns a short description for the JSON line, None if NVML is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
            return f"{len(allowed)} cpus ({allowed[0]}-{allowed[-1]})"
    except Exception:
        pass
    return None


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import numpy as np
    import torch
    import __graft_entry__ as ge
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    if not os.path.exists(pkg.lib_path()):
        ge.build(oracle=False)

    torch.cuda.set_device(local)
    numa = _bind_to_gpu_numa_node(local)       # pinned host buffers are then allocated next to this rank's GPU
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    prec = {"f64": pkg.PREC_F64, "f32": pkg.PREC_F32, "mixed": pkg.PREC_MIXED}[args.precision]
    N, T = args.envs, args.horizon

    # ---------------- multi-GPU parity self-check (N > 1), before anything is timed ----------------
    mcheck = None
    if world > 1 and not args.no_multi_check:
        mcheck = multi_gpu_check(pkg, capi, np, torch, dist, rank, local, world)

    def engine(k_epochs):
        return pkg.Engine(capi.default_config(n_envs=N, horizon=T, precision=prec, seed=0x5EED, env_offset=rank * N,
                                              k_epochs=k_epochs, device=local))

    E = engine(args.epochs)
    # identical synthetic weights on every rank: Linear-style init from the Philox host stream (seed 0)
    E.init_params(0)
    if world > 1:
        uid = [pkg.Engine.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        E.comm_init(rank, world, uid[0])

    def barrier(eng=None):
        (eng or E).sync()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # roofline denominators measured live (MEASURED_PEAKS.json only has HBM and bf16-tensor)
    peak32 = E.measure_fma_peak(False)
    peak64 = E.measure_fma_peak(True)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    hbm_src = "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    bf16_peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1590.0))
    bf16_src = ("MEASURED_PEAKS.json bf16_tflops_sustained" if "bf16_tflops_sustained" in peaks else
                "fallback 1590 TFLOP/s (B200_PROFILING.md)")

    cs = ClockSampler(local)
    if rank == 0:
        cs.start()
    # ---------------- warm-up ----------------
    for _ in range(max(args.warmup, 3)):
        E.rollout(T)
        if not args.no_update:
            E.update_resident()
    barrier()

    # ---------------- timed: rollout leg and update leg, K steps ----------------
    # inputs: the 344 MB trajectory buffer written by each rollout is larger than the 126 MB L2, and the
    # update streams the whole buffer every epoch -> no explicit flush needed ("inputs larger than L2").
    l0 = E.kernel_launches()
    roll_ms, upd_ms = [], []
    barrier()
    cs.mark()
    t_wall0 = time.time()
    for _ in range(args.steps):
        E.timer_start()
        E.rollout(T)
        roll_ms.append(E.timer_stop_ms())
        if not args.no_update:
            E.timer_start()
            st = E.update_resident()
            upd_ms.append(E.timer_stop_ms())
    barrier()
    t_wall = time.time() - t_wall0
    launches = E.kernel_launches() - l0
    upd_path = E.last_update_path() if upd_ms else None
    clocks = cs.stop() if rank == 0 else None
    roll_total = max_over_ranks(sum(roll_ms))
    upd_total = max_over_ranks(sum(upd_ms)) if upd_ms else 0.0
    steps_all = N * T * args.steps * world
    value = steps_all / (roll_total * 1e-3)
    upd_value = (N * T * args.epochs * args.steps * world) / (upd_total * 1e-3) if upd_ms else None

    # ---------------- e2e: C ABI with host buffers (pinned), copies inside the timed region ----------------
    e2e = None
    if not args.no_e2e:
        hp = capi.pinned_empty((E.P,), np.float32)
        hp[:] = E.get_params()
        out = {"obs": capi.pinned_empty((T, N, 6), np.float32), "action": capi.pinned_empty((T, N), np.float32),
               "logp": capi.pinned_empty((T, N), np.float32), "reward": capi.pinned_empty((T, N), np.float32),
               "flags": capi.pinned_empty((T, N), np.uint8)}
        for _ in range(2):
            E.set_params(hp); E.rollout_host(T, out, n_chunks=args.chunks)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            E.set_params(hp)                                  # H2D: the learner's current weights
            E.rollout_host(T, out, n_chunks=args.chunks)      # K1 in chunks; D2H of chunk i overlaps chunk i+1
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        d2h = sum(a.nbytes for a in out.values())
        e2e = {"value": steps_all / dt, "unit": "env-steps/s", "h2d_bytes_per_step": int(hp.nbytes),
               "d2h_bytes_per_step": int(d2h), "host_numa_binding": numa, "api": "b2_set_params + b2_rollout_host (pinned host buffers; chunked launches with the copy engine moving chunk i "
                      "while chunk i+1 is simulated; B2_ROLLOUT_HOST_DIRECT=1 selects direct host stores from the kernel)"}
        # the same through the opt-in compact record (25 instead of 37 B per env-step: obs4 + reward bit in the flag byte)
        if True:
            del out
            outc = {"obs4": capi.pinned_empty((T, N, 4), np.float32), "action": capi.pinned_empty((T, N), np.float32),
                    "logp": capi.pinned_empty((T, N), np.float32), "rflags": capi.pinned_empty((T, N), np.uint8)}
            for _ in range(2):
                E.set_params(hp); E.rollout_host_compact(T, outc, n_chunks=8)
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                E.set_params(hp)
                E.rollout_host_compact(T, outc, n_chunks=8)
            barrier()
            dtc = max_over_ranks(time.perf_counter() - t0)
            e2e["compact_record"] = {"value": steps_all / dtc, "unit": "env-steps/s",
                                     "d2h_bytes_per_step": int(sum(a.nbytes for a in outc.values())),
                                     "api": "b2_set_params + b2_rollout_host_compact (opt-in; the headline e2e above is MemoryNN's 37-B layout)"}
            del outc

    update_wide = None
    if not args.no_wide:
        update_wide = wide_leg(args, pkg, capi, np, rank, local, world, dist, barrier, max_over_ranks, bf16_peak, bf16_src)

    # ---------------- roofline denominators: measured FMA-issue peak AND the arithmetic one; frac uses the larger ----------
    sm_mhz = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    n_sm = torch.cuda.get_device_properties(local).multi_processor_count
    arith32 = n_sm * 128 * 2 * sm_mhz * 1e6 / 1e12     # 128 FP32 lanes per SM x 2 flop per FMA
    arith64 = n_sm * 64 * 2 * sm_mhz * 1e6 / 1e12      # 64 FP64 lanes per SM
    peak32_meas, peak64_meas = peak32, peak64
    peak32, peak64 = max(peak32_meas, arith32), max(peak64_meas, arith64)
    if dist:   # every rank needs rank 0's denominators for its own legs below
        t = torch.tensor([peak32, peak64], dtype=torch.float64, device="cuda")
        dist.broadcast(t, src=0)
        peak32, peak64 = float(t[0]), float(t[1])
    E.close()

    c4 = None
    if not args.no_c4 and args.precision == "f64":
        c4 = c4_leg(args, pkg, capi, np, rank, local, world, dist, barrier, max_over_ranks, peak64, peak32)

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        if mcheck is not None and not mcheck["ok"]:
            raise SystemExit(1)
        return
    configs1 = configs1_leg(args, pkg, capi, np, local) if (world == 1 and not args.no_configs1) else None

    # ---------------- roofline of the dominant kernel (k_rollout) ----------------
    ms_launch = sum(roll_ms) / len(roll_ms)           # one k_rollout launch per rollout leg (rank 0's own)
    steps_launch = N * T
    if args.precision == "f64":
        bound, peak = "fp64-fma", peak64
        achieved = FLOP_PHYS * steps_launch / (ms_launch * 1e-3) / 1e12
    else:
        bound, peak = "fp32-fma", peak32
        achieved = (FLOP_PHYS + FLOP_ACTOR) * steps_launch / (ms_launch * 1e-3) / 1e12
    # dram__bytes_read.sum + dram__bytes_write.sum of one k_rollout launch at this exact shape, from the committed
    # `ncu --set full` capture (profiles/NCU_TRAFFIC.json; algorithmic bytes are 41 B x N x T = 344 MB)
    traffic, traffic_src = _ncu_traffic(f"k_rollout_{args.precision}_{N}x{T}")
    roofline = {"kernel": "k_rollout", "bound": bound, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak if peak else None, "traffic": traffic, "traffic_source": traffic_src,
                "peak_measured": peak64_meas if args.precision == "f64" else peak32_meas,
                "peak_arithmetic": arith64 if args.precision == "f64" else arith32,
                "peak_source": "max(b2_measure_fma_peak: live FMA-issue microbenchmark on this GPU in this run, arithmetic: "
                               f"SMs x lanes x 2 x {sm_mhz:.0f} MHz sampled under load)",
                "flop_per_env_step": FLOP_PHYS if args.precision == "f64" else FLOP_PHYS + FLOP_ACTOR,
                "hbm": {"achieved_gbs": BYTES_STEP * steps_launch / (ms_launch * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                        "peak_source": hbm_src}}
    update = None
    if upd_ms:
        ums = sum(upd_ms) / len(upd_ms)
        ua = FLOP_UPDATE * N * T * args.epochs / (ums * 1e-3) / 1e12
        if upd_path == 1:
            # k_ppo_grad_tc: warp-level bf16x3 MMAs (legacy mma.sync path, HMMA).  Three denominators, all stated: the
            # measured tcgen05 bf16 peak / 3 (what a tcgen05 formulation could reach; the binding one for "tensor"), the
            # legacy-HMMA rate this kernel actually issues at (1,024 MAC/clk/SM measured by ncu: profiles/r2/) / 3, and the
            # FP32-FMA peak that bounded the kernel it replaces (a fraction > 1 means: past the old formulation's roofline)
            hmma = n_sm * 1024 * 2 * sm_mhz * 1e6 / 1e12
            uroof = {"kernel": "k_ppo_grad_tc", "bound": "tensor", "achieved": ua, "peak": bf16_peak / 3.0, "unit": "TFLOP/s",
                     "frac": ua / (bf16_peak / 3.0) if bf16_peak else None,
                     "peak_source": bf16_src + " / 3 (bf16x3: every fp32 product is three bf16 MMAs)",
                     "legacy_hmma_peak_div3": hmma / 3.0, "frac_of_legacy_hmma_div3": ua / (hmma / 3.0),
                     "fp32_fma_peak": peak32, "frac_of_fp32_fma_peak": ua / peak32 if peak32 else None,
                     "hbm_achieved_gbs": 76 * N * T * args.epochs / (ums * 1e-3) / 1e9}
        else:
            uroof = {"kernel": "k_ppo_grad", "bound": "fp32-fma", "achieved": ua, "peak": peak32,
                     "unit": "TFLOP/s", "frac": ua / peak32 if peak32 else None,
                     "peak_measured": peak32_meas, "peak_arithmetic": arith32,
                     "hbm_achieved_gbs": BYTES_UPDATE * N * T * args.epochs / (ums * 1e-3) / 1e9}
        update = {"value": upd_value, "unit": "sample-epochs/s", "ms_per_update": ums, "epochs": args.epochs,
                  "minibatches": 1, "samples": N * T, "kernel_path": {0: "fp32-fma", 1: "bf16x3 tensor cores (mma.sync)"}.get(upd_path),
                  "roofline": uroof, "last_loss": st.loss}

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        r = run_reference(10)
        if r:
            cpu_baseline = {"value": r["env_steps"] / r["rollout_s"], "unit": "env-steps/s", "cores": 1,
                            "kind": "reference",
                            "sample": f"{r['env_steps']} env-steps + {r['updates']} PPO::update(K=80) of the unmodified "
                                      f"reference sources via oracle/_ref/ref_harness (no MPI latency), {r['wall_s']:.1f} s wall",
                            "update_sample_epochs_per_s": r["sample_epochs"] / r["update_s"]}
        else:
            cpu_baseline = cpu_port_baseline()

    line = {"metric": "cartpole_env_steps_per_sec_incl_policy_inference", "value": value, "unit": "env-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": roll_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "BASELINE configs[2]: batched CartPole 65,536 envs/GPU x 128-step horizon, default MLP",
                       "n_envs_per_gpu": N, "horizon": T, "episode_len": 100, "mlp": "6-64-32-1 x2", "precision": args.precision,
                       "l2": "inputs larger than L2 (344 MB trajectory buffer rewritten per step)",
                       "parallelism": f"env-sharded x{world}, NCCL grad all-reduce"},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "update": update,
            "update_wide": update_wide, "c4": c4, "configs1": configs1, "multi_gpu_check": mcheck,
            "cpu_baseline": cpu_baseline,
            "fma_peaks_tflops": {"fp32_measured": peak32_meas, "fp64_measured": peak64_meas, "fp32_arithmetic": arith32,
                                 "fp64_arithmetic": arith64, "sm_mhz": sm_mhz},
            "wall_s_timed_region": t_wall}
    print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()
    if mcheck is not None and not mcheck["ok"]:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
