#!/usr/bin/env python
"""bench.py -- batched NMPC SQP solves/sec (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config 1|2|3|4|5]
                    [--batch B | --global-batch G] [--qp-warm-start 0|1|2] [--no-gather] [--labels]

A "step" is one pass of the hot path (QuadrupedAcadosSolver.init + solve,
/root/reference/mpc_controller/utils/solver.py:356-429) over one batch of synthetic OCPs.  Default:
BASELINE config 2 -- Go2 trot, 1024 perturbed initial states per GPU (the reference's own null-space
sampler, workloads.py), SQP_RTI -- started from the same first guess every step, so every step does
identical work.  The solver options are the reference's defaults (config_abstract.py:55-64: warm_start_nlp,
warm_start_qp = True -> HPIPM level 1, max_qp_iter 6, qp_tol 1e-2, nlp_tol 1e-1); with acados'
`warm_start_first_qp = False` the first QP of every solve is cold, so re-priming the stored QP iterate
between steps is a no-op for the one-QP configs 2, 4, 5.

Sharding: problems are independent.  `--batch B` gives every rank its own B problems (weak scaling, the
default); `--global-batch G` splits a fixed set of G problems into contiguous blocks (strong scaling, configs 3
and 5 at their BASELINE sizes).  With more than one rank the step ends with the path's one collective
(Behavior_Cloning/utils/Rollout_MPC.py:275,298): the packed solution + status rows
[B/G, 1868] are all-gathered over NCCL from the solver's device buffer (nmpc_pack_solution) inside the timed
region, so that every rank -- the dataset writer in particular -- holds the batch in global order.

  value    whole-job solves/s with the batch resident in HBM, CUDA events on the solver's stream, max over ranks.
  e2e      the same through the C ABI with HOST (pinned) buffers: H2D of every input of
           `QuadrupedAcadosSolver.init`, solve, (gather,) D2H of the solution (nmpc_set_output: wave by wave, behind the
           solve) and of the status inside the timed region.
  roofline dominant kernel against the HBM roof with SURVEY.md 8(d)'s numerator: 21.8k doubles per stage per IPM
           iteration x 26 stages x 8 B = 4.53 MB per IPM iteration per OCP; launch time from CUDA events around each
           launch (nmpc_set_profiling); what limits the kernel (`bound`) and the ncu counters that show it.
  cpu_baseline  the CPU oracle (a port of the algorithm, NOT acados) on the host cores, bounded sample.

`--impl reference` times that CPU oracle alone with all host threads (the reference's own
acados/HPIPM stack is not installable here: DESIGN.md, "Oracle").
`--config 1` is BASELINE config 1: ONE closed-loop-cadence rollout, `LocomotionMPC.open_loop` for 10 s = 250
warm-started sequential solves at B = 1 through the unbatched API (ms per tick against the 40 ms budget).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "batched NMPC SQP solves/sec (Go2 trot)"
UNIT = "solves/s"
N_NODES = 25
SOL_ROW = 26 * 42 + 25 * 31 + 1          # x | u | dt per node (parse_sol, solver.py:403-429: 1867) | status

# bytes if the stage blocks stream from HBM once per IPM iteration -- read BAbt + RSQrq + DCt
# (11.2k doubles per stage), write and re-read the factor L (5.3k doubles per stage, twice).
ALGO_DOUBLES_PER_STAGE_PER_IPM_ITER = 11_200 + 2 * 5_300          # 21 800
ALGO_BYTES_PER_IPM_ITER = 8 * ALGO_DOUBLES_PER_STAGE_PER_IPM_ITER * (N_NODES + 1)      # 4.53 MB per OCP
ALGO_BYTES_PER_LINEARISATION = 8 * 11_200 * (N_NODES + 1)         # 2.33 MB per OCP (8d: "linearise writes")
ALGO_IO_BYTES_PER_SOLVE = 55_000
# 8(d): ~18 MFLOP per IPM iteration of the dense algorithm (Riccati 11.7 + constraint Hessian 5.2 + back-solves 1)
ALGO_FLOP_PER_IPM_ITER = 18.0e6
FP64_PEAK_TFLOPS = 37.0                  # B200 nominal FP64 (vector = tensor path), no measured figure on this pool


def launch_plan(Bl: int) -> str:
    """How run_qp_fused (csrc/nmpc_abi.cu) sends one QP of a chunk of Bl OCPs to the GPU: full waves of 4 CTAs per SM, then
    the rest with the residency capped so that it spreads over all SMs.  `launch_ms` is the mean over these launches."""
    if os.environ.get("NMPC_WAVE_PLAN", "1") == "0":
        return f"1 launch of {Bl} CTAs"
    import torch
    n_sm = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
    full, rest = divmod(Bl, 4 * n_sm)
    if rest == 0 and full > 0:
        full, rest = full - 1, 4 * n_sm
    parts = ([f"{full} x {4 * n_sm} CTAs at 4 per SM"] if full else []) + [f"{rest} CTAs at {-(-rest // n_sm)} per SM"]
    return " + ".join(parts) + " (one launch each; launch_ms and algorithmic_bytes_per_launch are means over these launches)"


def ocps_per_launch(B: int) -> float:
    """The QP workspace is chunked (NMPC_CHUNK, default 4096 OCPs): a launch covers one chunk."""
    chunk = int(os.environ.get("NMPC_CHUNK", "4096"))
    return B / max(1, -(-B // chunk))


def workload_opts(cfg: int, qp_warm: int):
    base = dict(max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-1, warm_start_qp=qp_warm, warm_start_nlp=1)
    if cfg == 2:
        return dict(base, real_time_it=1, max_iter=1), ("trot",)
    if cfg == 3:
        return dict(base, real_time_it=0, max_iter=15), ("trot", "pace", "bound")
    if cfg == 4:          # + enable_time_opt: nu = 31, dt in [0.02, 0.07] (opt_dt_scale = (0.5, 1.75)), weight time_opt = 1e4
        return dict(base, real_time_it=0, max_iter=1, enable_time_opt=1, dt_min=0.02, dt_max=0.07,
                    time_opt_weight=1e4), ("trot",)
    if cfg == 5:          # default mpc_opt.py: SQP (not RTI), max_iter = 1 per replanning tick (mpc_opt.py:26-38)
        return dict(base, real_time_it=0, max_iter=1), ("trot",)
    raise SystemExit("--config must be 1, 2, 3, 4 or 5")


def workload_kwargs(cfg: int, kin) -> dict:
    if cfg == 4:
        return dict(restrict_cnt=True, kin=kin)
    if cfg == 5:
        return dict(push=True)
    return {}


def workload_name(cfg: int, B: int, qp_warm: int) -> str:
    tail = f", N=25, max_qp_iter=6, qp_warm_start={qp_warm} (first QP of a solve cold), reference null-space perturbation sampler"
    if cfg == 2:
        return f"config2: Go2 trot, {B} perturbed initial states per GPU, SQP_RTI" + tail
    if cfg == 3:
        return f"config3: Go2 trot/pace/bound mixed, {B} OCPs per GPU, full SQP max_iter=15 nlp_tol=1e-1" + tail
    if cfg == 4:
        return (f"config4: Go2 trot, {B} OCPs per GPU, enable_time_opt (nu=31, dt in [0.02, 0.07]), "
                "cnt_patch_restriction with Raibert footholds + opt_peak, SQP max_iter=1" + tail)
    return f"config5: DAgger sweep tick, Go2 trot, {B} push-perturbed OCPs per GPU, SQP max_iter=1" + tail


# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
def time_cpu_oracle(batch, opts: dict, n_problems: int, threads: int, repeats: int = 1) -> dict:
    """Bounded sample of the same workload on the host cores with the CPU oracle."""
    from oracle import oracle as O
    O.build()
    n = min(n_problems, len(batch["qv"]))
    sub = {k: (v[:n] if isinstance(v, np.ndarray) and v.shape[:1] == (len(batch["qv"]),) else v)
           for k, v in batch.items()}
    _, _, hg = O.kinematics(sub["qv"])
    s = O.OracleSolver(n, O.default_opts(**opts))
    if not opts.get("enable_time_opt"):
        s.set_fast(True)      # structured variant (oracle/nmpc_oracle_fast.h): 3x the dense restatement, same iterates
    from iterative_learning_nmpc_b200 import workloads as WL
    dt = 0.0
    for _ in range(repeats):                   # identical work every repeat, like the GPU steps
        WL.load_into(s, sub, hg)
        t0 = time.perf_counter()
        s.solve(threads)
        dt += time.perf_counter() - t0
    v = n * repeats / dt
    return {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
            "ms_per_solve_per_core": 1e3 * threads / v,
            "sample": f"{repeats} x {n} OCPs of the same batch, one solver per thread (OpenMP dynamic), {dt:.2f} s wall; "
                      "structured CPU restatement of the algorithm (classical Riccati on the 30x30 input block, structural "
                      "zeros skipped, gcc -O3), not acados/HPIPM"}


def run_reference(args) -> None:
    """`--impl reference`: the reference's CPU path is not installable (acados/HPIPM/CasADi/contact_tamp
    absent); the CPU oracle port stands in, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.config == 1:
        run_config1(args, reference=True)
        return
    from iterative_learning_nmpc_b200 import workloads as WL
    opts, gaits = workload_opts(args.config, args.qp_warm_start)
    threads = os.cpu_count() or 1
    n = max(threads * 64, 512)                # ~1 s of CPU work per step on 16 cores
    from oracle import oracle as O
    O.build()
    batch = WL.make_batch(n, seed=args.config - 1, gaits=gaits, **workload_kwargs(args.config, O.kinematics))
    vals = []
    for i in range(args.warmup + args.steps):
        r = time_cpu_oracle(batch, opts, n, threads)
        if i >= args.warmup:
            vals.append(r["value"])
    v = float(np.mean(vals))
    cb = dict(r, value=v, ms_per_solve_per_core=1e3 * threads / v)
    B = args.batch if not args.global_batch else args.global_batch // max(args.gpus, 1)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * n / v, "higher_is_better": True,
        "scaling": "strong" if args.global_batch else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args.config, B, args.qp_warm_start), "sample_per_step": n},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---------------------------------------------------------------------------------------------
def run_config1(args, reference: bool = False) -> None:
    """BASELINE config 1 (Behavior_Cloning/examples/test_mpc_rollout.py:12-22): one rollout, `LocomotionMPC.open_loop`
    for 10 s at v_des = [0.15, 0, 0] -- 250 sequential replanning ticks, each shifted-warm-started from the last,
    the first with 15 SQP iterations at tol/10 (mpc.py:416-473) -- through the UNBATCHED API (B = 1).  A "step" is one
    10 s rollout; the line reports ms per tick against the 40 ms replanning budget (mpc_opt.py:15)."""
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    from iterative_learning_nmpc_b200.mpc_controller.utils.profiling import time_fn  # noqa: F401
    from iterative_learning_nmpc_b200 import problem as P, workloads as WL
    from iterative_learning_nmpc_b200.mpc_controller.utils.transform import ypr_to_quat_state
    feet = WL.FEET
    T = 10.0

    def rollout(make_mpc, eps=0.0):
        mpc = make_mpc()
        mpc.set_command(np.array([0.15, 0.0, 0.0]), 0.0)
        qv = P.nominal_state(WL.STAND_HEIGHT)
        qv[6] += eps
        q_mj, v_mj = mpc.solver.dyn.convert_to_mujoco(qv[:18], qv[18:])
        ticks = []
        inner = mpc.optimize

        def timed_optimize(q, v):
            t0 = time.perf_counter()
            out = inner(q, v)
            ticks.append((time.perf_counter() - t0) * 1e3)
            return out
        mpc.optimize = timed_optimize
        t0 = time.perf_counter()
        traj = mpc.open_loop(q_mj, v_mj, T)
        wall = time.perf_counter() - t0
        return traj, np.array(ticks), wall

    def gpu_mpc():
        return LocomotionMPC(None, feet, "go2", "trot", joint_ref=P.GO2_JOINT_HOME, print_info=False,
                             compute_timings=False, solve_async=False)

    def cpu_mpc():
        # the same front-end on the CPU oracle through the test-only adapter (tests/oracle_backend.py): this is the
        # cpu_baseline / reference arm, the one place bench.py may execute oracle/
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from oracle_backend import OracleHandle
        from iterative_learning_nmpc_b200.mpc_controller.utils.acados_helper import AcadosSolverHelper as H
        orig = H._make_handle
        H._make_handle = lambda self: OracleHandle(self.batch, self._opts, self.device)
        try:
            return gpu_mpc()
        finally:
            H._make_handle = orig

    res = {}
    if not reference:
        import torch
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
        clocks = ClockSampler(0)
        for _ in range(1):
            rollout(gpu_mpc)
        clocks.start()
        runs = [rollout(gpu_mpc) for _ in range(max(1, args.steps))]
        clk = clocks.stop()
        traj = runs[0][0]
        ticks = np.concatenate([r[1][1:] for r in runs])
        first = float(np.median([r[1][0] for r in runs]))
        wall = float(np.mean([r[2] for r in runs]))
        n_ticks = len(runs[0][1])
        res = {"metric": METRIC, "value": n_ticks / wall, "unit": UNIT, "n_gpus": 1, "steps": len(runs), "warmup": 1,
               "ms_per_step": wall * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f64", "data": "synthetic",
               "config": {"workload": "config1: Go2 trot, LocomotionMPC.open_loop 10 s at v_des=[0.15,0,0], default mpc_opt.py, "
                                      f"{n_ticks} sequential warm-started solves at B=1 (unbatched API), first tick 15 SQP "
                                      "iterations at tol/10"},
               "tick_ms": {"median": float(np.median(ticks)), "p95": float(np.percentile(ticks, 95)),
                           "max": float(ticks.max()), "first_tick": first, "budget_ms": 40.0, "ticks_per_rollout": n_ticks},
               "e2e": {"value": n_ticks / wall, "unit": UNIT, "h2d_bytes_per_step": n_ticks * (42 + 26 * 44 + 25 * 76 + 52) * 8,
                       "d2h_bytes_per_step": n_ticks * 1867 * 8,
                       "note": "config 1 IS the host API: every tick copies its inputs in and the solution out"},
               "gpu_launches": None, "clocks": clk}
    if reference or not args.no_cpu_baseline:
        traj_c, ticks_c, wall_c = rollout(cpu_mpc)
        cb = {"value": len(ticks_c) / wall_c, "unit": UNIT, "cores": 1, "kind": "port",
              "ms_per_solve_per_core": float(np.median(ticks_c[1:])), "first_tick_ms": float(ticks_c[0]),
              "sample": f"the same 10 s rollout on the CPU oracle, one solver on one core, {wall_c:.1f} s wall; "
                        "CPU restatement of the algorithm, not acados/HPIPM"}
        if reference:
            v = cb["value"]
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                              "steps": 1, "warmup": 0, "ms_per_step": wall_c * 1e3, "higher_is_better": True,
                              "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                              "config": {"workload": "config1 (see the native arm)"}, "cpu_baseline": cb,
                              "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
            return
        res["cpu_baseline"] = cb
        n = min(len(traj), len(traj_c))
        # the sim-free loop feeds every plan back as the next state and the 6-iteration IPM never converges (status
        # MAXITER each tick, which the reference ignores): the map is not contractive.  For scale, the oracle against
        # itself with the first joint angle moved by 1e-9 rad is reported next to GPU vs oracle.
        traj_p, _, _ = rollout(cpu_mpc, eps=1e-9)
        res["trajectory_deviation_vs_oracle"] = {
            "max_abs_q": float(np.abs(traj[:n] - traj_c[:n]).max()),
            "max_abs_q_first_second": float(np.abs(traj[:1000] - traj_c[:1000]).max()),
            "oracle_vs_oracle_perturbed_1e-9": {"max_abs_q": float(np.abs(traj_p[:n] - traj_c[:n]).max()),
                                                "max_abs_q_first_second": float(np.abs(traj_p[:1000] - traj_c[:1000]).max())},
            "final_base_xyz_gpu": traj[n - 1, :3].tolist(), "final_base_xyz_oracle": traj_c[n - 1, :3].tolist()}
    print(json.dumps(res), flush=True)


# ---------------------------------------------------------------------------------------------
def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--batch", type=int, default=1024, help="OCPs per GPU (weak scaling)")
    ap.add_argument("--global-batch", type=int, default=0, help="total OCPs, split over the ranks (strong scaling)")
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--qp-warm-start", type=int, default=1, choices=[0, 1, 2],
                    help="HPIPM warm_start level; the reference's warm_start_qp=True is 1, 2 is the hot start")
    ap.add_argument("--no-gather", action="store_true", help="leave the NCCL gather of the solutions out of the step")
    ap.add_argument("--labels", action="store_true",
                    help="also compute and gather the DAgger labels of the first replanning interval (nmpc_plan_torques)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return
    if args.config == 1:
        if int(os.environ.get("RANK", "0")) == 0:
            run_config1(args)
        return

    import torch
    import torch.distributed as dist

    from iterative_learning_nmpc_b200 import workloads as WL
    from iterative_learning_nmpc_b200._lib import default_opts
    from iterative_learning_nmpc_b200.core import NmpcSolver

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    out_stream = sys.stdout
    if world > 1:
        # rank 0 prints ONE JSON line on stdout: NCCL's version banner (and anything else a library writes to
        # file descriptor 1) is sent to stderr for the whole run; the JSON line goes to the saved descriptor
        sys.stdout.flush()
        out_stream = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    opts, gaits = workload_opts(args.config, args.qp_warm_start)
    if args.global_batch:
        # strong scaling: ONE fixed problem set (seeded by the config alone), rank g owns the contiguous block
        # [g*B, (g+1)*B) of it; the last block is padded with repeats of its last problem if G % world != 0
        G = args.global_batch
        B = -(-G // world)
        lo, hi = rank * B, min((rank + 1) * B, G)
        n_real = max(hi - lo, 0)
    else:
        B = args.batch
        G = B * world
        n_real = B
    solver = NmpcSolver(B, default_opts(**opts), device=local)
    if args.global_batch:
        # every rank draws the whole set from the same seed and keeps its block (generation is outside the timed region)
        full = WL.make_batch(G, seed=args.config - 1, gaits=gaits, **workload_kwargs(args.config, solver.kinematics))
        sel = np.minimum(np.arange(lo, lo + B), G - 1)
        batch = WL.take(full, sel)
        del full
    else:
        batch = WL.make_batch(B, seed=1000 * rank + args.config - 1, gaits=gaits,
                              **workload_kwargs(args.config, solver.kinematics))
    stream = torch.cuda.ExternalStream(solver.stream_ptr, device=local)
    # compact planner outputs, as `QuadrupedAcadosSolver.init` receives them (the expansion into per-node
    # parameters / references, h0, stance-foot pinning and the first guess are device kernels)
    cp = batch["compact"]
    tensors = {"W": batch["W"], "W_e": batch["W_e"], "qv": batch["qv"], "i_node": batch["i_node"],
               "base_ref": cp["base_ref"], "base_ref_e": cp["base_ref_e"], "step_height": cp["step_height"],
               "cnt_sequence": cp["cnt_sequence"], "swing_peak": cp["swing_peak"]}
    if cp["cnt_locations"] is not None:
        tensors["cnt_locations"] = cp["cnt_locations"]
    host = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in tensors.items()}
    dev = {k: v.cuda(non_blocking=False) for k, v in host.items()}
    out_x = torch.empty(B, N_NODES + 1, 42, dtype=torch.float64).pin_memory()
    out_u = torch.empty(B, N_NODES, 30, dtype=torch.float64).pin_memory()
    out_st = torch.empty(B, dtype=torch.int32).pin_memory()
    lib, hnd = solver.lib, solver._h
    from iterative_learning_nmpc_b200._lib import check, ptr
    joint_ref = np.ascontiguousarray(cp["joint_ref"], dtype=np.float64)
    p_gain = np.ascontiguousarray(cp["p_gain"], dtype=np.float64)

    gather = world > 1 and not args.no_gather
    n_lab = 40                                     # sim steps of one replanning interval (25 Hz replanning, 1 kHz sim)
    send = torch.empty(B, SOL_ROW, dtype=torch.float64, device="cuda")
    recv = torch.empty(world * B, SOL_ROW, dtype=torch.float64, device="cuda") if gather else None
    lab_send = torch.empty(B, n_lab, 12, dtype=torch.float64, device="cuda") if args.labels else None
    lab_recv = torch.empty(world * B, n_lab, 12, dtype=torch.float64, device="cuda") if (gather and args.labels) else None
    g_ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def step(src, sync_out: bool, time_gather: bool = False):
        # QuadrupedAcadosSolver.init: weights, references + contact plan (expanded on the device), x0, stance feet, guess
        solver.set_weights(src["W"], src["W_e"])
        check(lib.nmpc_init_tick(hnd, ptr(src["base_ref"]), ptr(src["base_ref_e"]), ptr(joint_ref), 0,
                                 ptr(src["step_height"]), 1, ptr(src["cnt_sequence"]), ptr(src["swing_peak"]),
                                 ptr(src.get("cnt_locations")), ptr(p_gain), float(cp["cnt_radius"]),
                                 float(cp["height_offset"]), int(cp["restrict_cnt"])), "nmpc_init_tick")
        solver.set_x0(src["qv"])
        solver.cold_guess()
        solver.pin_stance_feet(src["i_node"])
        solver.reset_duals()
        solver.solve(status=None, sync=False)
        if gather:
            # the path's one collective: packed [B, 1868] rows (x | u | dt | status) straight from the solver's device
            # buffers, all-gathered on the solver's stream (the contiguous partition makes `recv` globally ordered)
            solver.pack_solution(send)
            if args.labels:
                solver.plan_torques_into(1000, 0, n_lab, action=lab_send)
            with torch.cuda.stream(stream):
                if time_gather:
                    g_ev[0].record(stream)
                dist.all_gather_into_tensor(recv, send)
                if args.labels:
                    dist.all_gather_into_tensor(lab_recv, lab_send)
                if time_gather:
                    g_ev[1].record(stream)
        if sync_out:     # QuadrupedAcadosSolver.solve + parse_sol: status and solution back on the host
            # x / u reach the pinned host buffers through nmpc_set_output (below): every wave's rows are copied while the
            # next wave is being solved; reading the status synchronises the solve stream, which waits for those copies
            solver.get_status_into(out_st)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(src, sync_out: bool, steps: int):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            step(src, sync_out)
        e1.record(stream)
        solver.synchronize()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms, wall], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    # ---- device-resident: value -------------------------------------------------------------
    for _ in range(args.warmup):
        step(dev, False)
    solver.synchronize()
    clocks = ClockSampler(local)
    clocks.start()
    l0 = solver.launch_count
    ms_dev, _ = timed(dev, False, args.steps)
    launches = solver.launch_count - l0
    # ---- per-kernel timing for the roofline (separate pass so the events do not perturb `value`)
    solver.set_profiling(True)
    barrier()
    gather_ms = []
    for _ in range(args.steps):
        step(dev, False, time_gather=True)
        if gather:
            solver.synchronize()
            torch.cuda.synchronize()
            gather_ms.append(g_ev[0].elapsed_time(g_ev[1]))
    kt = solver.kernel_times()
    solver.set_profiling(False)
    stats = solver.get_stats()
    st_dev = solver.solve_status()
    # ---- host buffers through the C ABI: e2e ---------------------------------------------------
    solver.set_output(out_x, out_u)
    for _ in range(2):
        step(host, True)
    ms_e2e, wall_e2e = timed(host, True, args.steps)
    solver.set_output()
    xs_chk, us_chk, _, _ = solver.get_solution(duals=False)      # the streamed copies are the solver's final iterate
    assert np.array_equal(out_x.numpy(), xs_chk) and np.array_equal(out_u.numpy(), us_chk), "streamed output differs"
    clk = clocks.stop()
    status_counts = {int(k): int(v) for k, v in zip(*np.unique(st_dev[:max(n_real, 1)], return_counts=True))}
    sqp_hist = np.bincount(stats["sqp_iter"][:max(n_real, 1)], minlength=2).tolist()
    if gather:      # the gathered buffer is the global batch in order: check the own block against the solver's buffers
        solver.synchronize()
        torch.cuda.synchronize()
        x_own = recv[rank * B:(rank + 1) * B, :26 * 42].cpu().numpy().reshape(B, 26, 42)
        xs, _, _, _ = solver.get_solution(duals=False)
        assert np.array_equal(x_own, xs), "gathered block differs from the solver's own solution"
        assert np.array_equal(recv[rank * B:(rank + 1) * B, SOL_ROW - 1].cpu().numpy().astype(np.int32), st_dev)

    h2d = sum(host[k].numel() * host[k].element_size() for k in host)
    d2h = out_x.numel() * 8 + out_u.numel() * 8 + out_st.numel() * 4
    total = G
    value = total * args.steps / (ms_dev * 1e-3)
    e2e = total * args.steps / (max(ms_e2e, wall_e2e) * 1e-3)

    n_ipm = float(np.mean(stats["qp_iter"]))
    n_sqp = float(np.mean(stats["sqp_iter"]))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    kinds = {k: {"ms_total": v[0], "launches": v[1]} for k, v in kt.items() if v[1]}
    for k in kinds:
        kinds[k]["launch_ms"] = kinds[k]["ms_total"] / kinds[k]["launches"]
    dom = max(("qp_fused", "linearize"), key=lambda k: kinds.get(k, {}).get("ms_total", 0.0))
    Bl = ocps_per_launch(B)
    # SURVEY 8(d) numerator per launch of the dominant kernel.  One launch of the fused kernel = one QP of every problem
    # of the chunk that is still active; with RTI / max_iter = 1 that is n_ipm IPM iterations of all of them.
    qp_launches_per_step = kinds.get("qp_fused", {}).get("launches", args.steps) / max(args.steps, 1)
    if dom == "qp_fused":
        # all QPs of the step (sum over problems of their SQP iterations x IPM iterations) over the launches that ran them
        algo = ALGO_BYTES_PER_IPM_ITER * n_ipm * n_sqp * B / max(qp_launches_per_step, 1e-9)
        flop = ALGO_FLOP_PER_IPM_ITER * n_ipm * n_sqp * B / max(qp_launches_per_step, 1e-9)
    else:
        algo = ALGO_BYTES_PER_LINEARISATION * Bl
        flop = None
    ms_l = kinds[dom]["launch_ms"]
    achieved = algo / (ms_l * 1e-3) / 1e9
    counters, traffic = None, None
    try:      # ncu counters of the same kernel from the committed capture (profiles/fused_counters.json: round, commit, batch)
        counters = json.load(open(os.path.join(ROOT, "profiles", "fused_counters.json")))
        if dom == "qp_fused" and args.config == int(counters.get("config", 2)) and counters.get("dram_bytes_per_launch"):
            # measured bytes per OCP-QP of the capture x the OCP-QPs one launch of this run covers on average (a chunk goes
            # out in several launches, one per wave): per launch, like `achieved`
            traffic = (float(counters["dram_bytes_per_launch"]) / float(counters["batch"])
                       * B * n_sqp / max(qp_launches_per_step, 1e-9))
    except (OSError, ValueError):
        pass
    frac = achieved / peak
    roofline = {"bound": "hbm" if frac >= 0.6 else "latency", "roof": "hbm",
                "kernel": {"qp_fused": "nmpc_qp_fused_kernel", "linearize": "nmpc_model_kernel+nmpc_linearize_sr_kernel"}[dom],
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": frac, "traffic": traffic,
                "algorithmic_bytes": "SURVEY.md 8(d): (11.2k read + 2 x 5.3k factor) doubles per stage per IPM iteration x 26 stages "
                                     "x 8 B = 4.53 MB per IPM iteration per OCP, x mean IPM iterations x OCPs per launch",
                "algorithmic_bytes_per_launch": algo, "launch_ms": ms_l,
                "launch_plan": launch_plan(int(round(Bl))),
                "fp64_flop_frac": (flop / (ms_l * 1e-3) / (FP64_PEAK_TFLOPS * 1e12)) if flop else None,
                "fp64_peak_tflops_nominal": FP64_PEAK_TFLOPS,
                "whole_solve_algorithmic_bytes": n_sqp * (ALGO_BYTES_PER_LINEARISATION + n_ipm * ALGO_BYTES_PER_IPM_ITER) + ALGO_IO_BYTES_PER_SOLVE,
                "whole_step_frac": (n_sqp * (ALGO_BYTES_PER_LINEARISATION + n_ipm * ALGO_BYTES_PER_IPM_ITER) + ALGO_IO_BYTES_PER_SOLVE)
                                   * B / (ms_dev / args.steps * 1e-3) / 1e9 / peak,
                "ncu_counters": counters,
                "traffic_source": "static ncu capture (profiles/fused_counters.json), scaled to this batch; not re-measured in this run",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                "per_kind": kinds, "profiled_steps": args.steps, "mean_ipm_iter": n_ipm, "mean_sqp_iter": n_sqp}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        # ~0.9k solves/s on 16 cores: 8192 OCP solves (the 1024-OCP batch, eight times) is about 10 s of CPU work
        cpu = time_cpu_oracle(batch, opts, min(B, 1024), threads, repeats=8 if opts["max_iter"] == 1 else 1)

    if rank == 0:
        comm = (world * B * SOL_ROW * 8 + (world * B * n_lab * 12 * 8 if args.labels else 0)) if gather else 0
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.global_batch else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.config, B, args.qp_warm_start), "batch_per_gpu": B, "global_batch": total,
                       "parallelism": (f"problem-sharded x{world} (contiguous blocks); one NCCL all-gather of the packed solution "
                                       f"+ status rows [B,{SOL_ROW}] per step inside the timed region" if gather else
                                       f"problem-sharded x{world}, no data-path collective"),
                       "l2": "per-step working set (stage records 1692 + factor records + stage vectors, > 6k doubles per "
                             f"stage) is > {Bl * 26 * 6000 * 8 / 1e9:.2f} GB per launch, larger than the 126 MB L2"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": max(ms_e2e, wall_e2e) / args.steps},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
            "status_counts": status_counts, "sqp_iter_histogram": sqp_hist,
            "comm_bytes_per_step": comm, "gather_ms": (float(np.median(gather_ms)) if gather_ms else None),
        }), file=out_stream, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
