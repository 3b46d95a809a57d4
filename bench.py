#!/usr/bin/env python
"""bench.py -- batched NMPC SQP solves/sec (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--batch B] [--config 2|3|4|5]

A "step" is one pass of the hot path (QuadrupedAcadosSolver.init + solve,
/root/reference/mpc_controller/utils/solver.py:356-429) over one batch of synthetic OCPs:
BASELINE config 2 -- Go2 trot, 1024 perturbed initial states per GPU, SQP_RTI -- cold-started from
the same initial guess every step (so every step does identical work).  Weak scaling: every rank
owns its own 1024 problems; problems are independent, there is no collective on the data path.

  value    whole-job solves/s with the batch resident in HBM (device copies of the guess + solve),
           CUDA events on the solver's stream, max over ranks.
  e2e      the same through the C ABI with HOST (pinned) buffers: H2D of every input of
           `QuadrupedAcadosSolver.init` (weights, base references, contact masks / peaks / locations, x0,
           node index), solve, D2H of the solution and status inside the timed region.
  roofline dominant kernel (by summed device time): algorithmic bytes per launch / mean launch time
           (CUDA events around each launch, nmpc_set_profiling) over the measured HBM peak.
  cpu_baseline  the CPU oracle (a port of the algorithm, NOT acados) on the host cores, bounded sample.

`--impl reference` times that CPU oracle alone with all host threads (the reference's own
acados/HPIPM stack is not installable here: DESIGN.md, "Oracle").
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

# Algorithmic HBM traffic per OCP per launch of the heavy kernels, in doubles per stage (DESIGN.md
# "Roofline"): every dense stage block a kernel needs is counted once per sweep that needs it.
_G, _M, _C, _D, _LP, _LPU = 3108, 5330, 3024, 168, 2880, 1830
ALGO_DOUBLES_PER_STAGE = {
    # K1: read C, M (H + gradient), G, bounds/masks, iterate vectors; write Ht and residual vectors
    "qp_stage": _C + _M + _G + _D + 500 + _LP + 400,
    # K3: backward: read Ht, G, (qt, rb), write L; forward: read the u panels of L, G, rb, write dz
    "qp_factor": (_LP + _G + 114 + _LP) + (_LPU + _G + 42 + 72),
    # K7: backward: read L, G, (qt, rb), write l, p; forward: read L, G, (rb, p, l), write dz, dpi
    "qp_corr": (_LP + _G + 114 + 114) + (_LP + _G + 156 + 114),
    # model evaluation + assembly: read x, u, p, yref; write G, M, C, D (record written + read once)
    "linearize": _G + _M + _C + _D + 2 * 735 + 250,
}
# The fused kernel (nmpc_qp_fused.cuh) runs a whole QP: max_qp_iter x (stage + factor + corrector) plus the final
# residual pass.  Its ALGORITHMIC bytes are those of the dense algorithm above (the invariant SURVEY.md 8d
# prescribes: structure exploitation must not shrink the numerator); the bytes it really moves are `traffic`.
ALGO_DOUBLES_PER_STAGE_PER_ITER = (ALGO_DOUBLES_PER_STAGE["qp_stage"] + ALGO_DOUBLES_PER_STAGE["qp_factor"]
                                   + ALGO_DOUBLES_PER_STAGE["qp_corr"])


def ocps_per_launch(B: int) -> float:
    """The QP workspace is chunked (NMPC_CHUNK, default 4096 OCPs): a launch covers one chunk."""
    chunk = int(os.environ.get("NMPC_CHUNK", "4096"))
    return B / max(1, -(-B // chunk))


def algo_bytes_per_launch(kind: str, B: int, N: int = N_NODES, n_ipm: float = 6.0) -> float:
    B = ocps_per_launch(B)
    if kind == "qp_fused":
        per_stage = n_ipm * ALGO_DOUBLES_PER_STAGE_PER_ITER + ALGO_DOUBLES_PER_STAGE["qp_stage"]
        return 8.0 * per_stage * (N + 1) * B
    return 8.0 * ALGO_DOUBLES_PER_STAGE[kind] * (N + 1) * B


def workload_opts(cfg: int):
    if cfg == 2:
        return dict(real_time_it=1, max_iter=1, max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-1, warm_start_qp=0), ("trot",)
    if cfg == 3:
        return dict(real_time_it=0, max_iter=15, max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-1, warm_start_qp=0), \
            ("trot", "pace", "bound")
    if cfg in (4, 5):     # default mpc_opt.py: SQP (not RTI), max_iter = 1 per replanning tick (mpc_opt.py:26-38)
        return dict(real_time_it=0, max_iter=1, max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-1, warm_start_qp=0), ("trot",)
    raise SystemExit("--config must be 2, 3, 4 or 5")


def workload_kwargs(cfg: int, kin) -> dict:
    if cfg == 4:
        return dict(restrict_cnt=True, kin=kin)
    if cfg == 5:
        return dict(push=True)
    return {}


def workload_name(cfg: int, B: int) -> str:
    if cfg == 2:
        return f"config2: Go2 trot, {B} perturbed initial states per GPU, SQP_RTI, N=25, max_qp_iter=6"
    if cfg == 3:
        return f"config3: Go2 trot/pace/bound mixed, {B} OCPs per GPU, full SQP max_iter=15, N=25, max_qp_iter=6"
    if cfg == 4:
        return (f"config4 (without enable_time_opt, which is not built): Go2 trot, {B} OCPs per GPU, cnt_patch_restriction "
                "with Raibert footholds + opt_peak, SQP max_iter=1, N=25, max_qp_iter=6")
    return f"config5: DAgger sweep tick, Go2 trot, {B} push-perturbed OCPs per GPU, SQP max_iter=1, N=25, max_qp_iter=6"


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
    from iterative_learning_nmpc_b200 import workloads as WL
    dt = 0.0
    for _ in range(repeats):                   # identical cold-started work every repeat, like the GPU steps
        WL.load_into(s, sub, hg)
        t0 = time.perf_counter()
        s.solve(threads)
        dt += time.perf_counter() - t0
    return {"value": n * repeats / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{repeats} x {n} OCPs of the same batch, one solver per thread (OpenMP dynamic), {dt:.2f} s wall; "
                      "CPU restatement of the algorithm, not acados/HPIPM"}


def run_reference(args) -> None:
    """`--impl reference`: the reference's CPU path is not installable (acados/HPIPM/CasADi/contact_tamp
    absent); the CPU oracle port stands in, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from iterative_learning_nmpc_b200 import workloads as WL
    opts, gaits = workload_opts(args.config)
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
    cb = dict(r, value=v)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * n / v, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args.config, args.batch), "sample_per_step": n},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---------------------------------------------------------------------------------------------
def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--batch", type=int, default=1024, help="OCPs per GPU")
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup

    if args.impl == "reference":
        run_reference(args)
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

    B = args.batch
    opts, gaits = workload_opts(args.config)
    solver = NmpcSolver(B, default_opts(**opts), device=local)
    batch = WL.make_batch(B, seed=1000 * rank + args.config - 1, gaits=gaits, **workload_kwargs(args.config, solver.kinematics))
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

    def step(src, sync_out: bool):
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
        if sync_out:     # QuadrupedAcadosSolver.solve + parse_sol: status and solution back on the host
            solver.solve(status=out_st)
            solver.get_solution_into(x=out_x, u=out_u)
        else:
            solver.solve(status=None, sync=False)

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
    for _ in range(args.steps):
        step(dev, False)
    kt = solver.kernel_times()
    solver.set_profiling(False)
    stats = solver.get_stats()
    # ---- host buffers through the C ABI: e2e ---------------------------------------------------
    for _ in range(2):
        step(host, True)
    ms_e2e, wall_e2e = timed(host, True, args.steps)
    clk = clocks.stop()
    status_counts = {int(k): int(v) for k, v in zip(*np.unique(out_st.numpy(), return_counts=True))}

    h2d = sum(host[k].numel() * host[k].element_size() for k in host)
    d2h = out_x.numel() * 8 + out_u.numel() * 8 + out_st.numel() * 4
    total = B * world
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
    kinds = {k: {"ms_total": v[0], "launches": v[1]} for k, v in kt.items()}
    heavy = [k for k in list(ALGO_DOUBLES_PER_STAGE) + ["qp_fused"] if kinds.get(k, {}).get("launches")]
    for k in heavy:
        ms_l = kinds[k]["ms_total"] / kinds[k]["launches"]
        kinds[k]["launch_ms"] = ms_l
        kinds[k]["achieved_gbs"] = algo_bytes_per_launch(k, B, n_ipm=n_ipm) / (ms_l * 1e-3) / 1e9
        kinds[k]["frac_of_hbm_peak"] = kinds[k]["achieved_gbs"] / peak
    dom = max(heavy, key=lambda k: kinds[k]["ms_total"])
    traffic = None
    try:                                   # DRAM bytes per launch from the committed ncu capture, scaled to this batch
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if dom in tr and args.config == int(tr.get("config", 2)):
            traffic = float(tr[dom]) * ocps_per_launch(B) / float(tr["batch"])
    except (OSError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": {"qp_stage": "nmpc_qp_stage_kernel", "qp_factor": "nmpc_qp_factor_kernel",
                                           "qp_corr": "nmpc_qp_corr_solve_kernel", "qp_fused": "nmpc_qp_fused_kernel",
                                           "linearize": "nmpc_model_kernel+nmpc_linearize_sr_kernel"}[dom],
                "achieved": kinds[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                "frac": kinds[dom]["frac_of_hbm_peak"], "traffic": traffic,
                "algorithmic_bytes": "dense-algorithm bytes (SURVEY.md 8d invariant); `traffic` = measured DRAM bytes "
                                     "of the structured implementation (ncu, profiles/traffic.json)",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                "algorithmic_bytes_per_launch": algo_bytes_per_launch(dom, B, n_ipm=n_ipm), "launch_ms": kinds[dom]["launch_ms"],
                "per_kind": kinds, "profiled_steps": args.steps,
                "mean_ipm_iter": n_ipm, "mean_sqp_iter": n_sqp}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        # ~0.85k solves/s on 16 cores: 8192 OCP solves (the 1024-OCP batch, eight times) is about 10 s of CPU work
        cpu = time_cpu_oracle(batch, opts, B, threads, repeats=8 if args.config == 2 else 1)

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.config, B), "batch_per_gpu": B, "global_batch": total,
                       "parallelism": f"problem-sharded x{world}, no data-path collective",
                       "l2": "per-step working set (stage records 1692 + factor records 3680 + stage vectors 752 doubles per "
                             f"stage) is {ocps_per_launch(B) * 26 * (1692 + 3680 + 752) * 8 / 1e9:.2f} GB per launch, larger than the 126 MB L2"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": max(ms_e2e, wall_e2e) / args.steps},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
            "status_counts": status_counts,
        }), file=out_stream, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
