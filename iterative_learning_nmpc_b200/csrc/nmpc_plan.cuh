// nmpc_plan.cuh -- batched plan -> torque step, straight off the resident solution.
//
// Reference (one problem at a time, on the host):
//   LocomotionMPC.interpolate_state_trajectory / interpolate_trajectory_with_derivatives  mpc.py:371-414
//       cubic Hermite of (q, v) and of (v, a') on the node grid, a' = [a_0, a_0, a_1, ..., a_{N-1}],
//       sampled at linspace(0, T, n_interp + 1)[1:]
//   LocomotionMPC._adopt_plan                                                             mpc.py:535-545
//       a_plan = a_sol[id_repeat], f_plan = f_sol[id_repeat], id_repeat = int32(linspace(0,1,n) (N-1))
//   QuadrupedDynamics.id_torques                                                          dynamics.py:136-163
//   LocomotionMPC._compute_pd_torques                                                     mpc.py:592-599
//   DAgger label  action = (tau + kd v) / kp + q                                          Rollout_MPC.py:298
//
// One thread per (problem, plan step): the generated inverse dynamics dominates, everything else is
// a handful of flops on data the thread gathers from its own OCP's solution.
#pragma once
#include "nmpc_common.cuh"

struct PlanArgs {
    const double* x;        // [B][(N+1)*42]
    const double* u;        // [B][N*30]
    const double* qv_meas;  // [B][n_steps][36] or nullptr (the plan itself is the measured state)
    double* q_plan; double* v_plan; double* tau_ff; double* tau; double* action;   // any may be nullptr
    int B, N, n_interp, step0, n_steps;
    double dt, kp, kd;
    const double* dtn;      // [B][N] node durations of the solution (enable_time_opt) or nullptr
};

// Sample `ip` of the interpolated plan (sample 0 of linspace(0, T, n + 1) is dropped): cubic Hermite of (q, v) and of
// (v, a') on the node grid.  dtn != nullptr: per-node durations (enable_time_opt: time_traj = cumsum(dt_node_sol)).
__device__ __forceinline__ void nmpc_plan_sample(const double* xb, const double* ub, const double* dtn, double dt, int N,
                                                 int n, int ip, double* q, double* v)
{
    // node grid by running sum (numpy.cumsum order), sample time as numpy.linspace computes it
    double T = 0.0;
    for (int k = 0; k < N; k++) T += dtn ? dtn[k] : dt;
    const double step = T / n;
    const int j = ip + 1;
    const double t = (j >= n) ? T : j * step;
    int k = 0;
    double tk = 0.0, tk1 = dtn ? dtn[0] : dt;
    while (k < N - 1 && t >= tk1) { k++; tk = tk1; tk1 += dtn ? dtn[k] : dt; }
    const double hh = tk1 - tk, tau_ = (t - tk) / hh, t2 = tau_ * tau_, t3 = t2 * tau_;
    const double h00 = 2.0 * t3 - 3.0 * t2 + 1.0, h10 = t3 - 2.0 * t2 + tau_, h01 = -2.0 * t3 + 3.0 * t2, h11 = t3 - t2;
    const double* x0 = xb + (size_t)k * NX;
    const double* x1 = x0 + NX;
    const double* a0 = ub + (size_t)(k > 0 ? k - 1 : 0) * NU;     // slope of v at knot k
    const double* a1 = ub + (size_t)k * NU;                      // slope of v at knot k + 1
#pragma unroll
    for (int c = 0; c < NQ; c++) {
        q[c] = h00 * x0[c] + h10 * hh * x0[NQ + c] + h01 * x1[c] + h11 * hh * x1[NQ + c];
        v[c] = h00 * x0[NQ + c] + h10 * hh * a0[c] + h01 * x1[NQ + c] + h11 * hh * a1[c];
    }
}

// zero-order hold of the inputs: node index of plan step ip (id_repeat, mpc.py:110)
__device__ __forceinline__ int nmpc_plan_hold_node(int N, int n, int ip)
{
    const double frac = (ip >= n - 1) ? 1.0 : ip * (1.0 / (n - 1));
    return (int)(frac * (N - 1));
}

__global__ void __launch_bounds__(64, 4) nmpc_plan_torque_kernel(PlanArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.B * a.n_steps) return;
    const int b = i / a.n_steps, s = i % a.n_steps, N = a.N, n = a.n_interp;
    const int ip = a.step0 + s;                       // index into the plan arrays (sample 0 is dropped)
    const double* xb = a.x + (size_t)b * (N + 1) * NX;
    const double* ub = a.u + (size_t)b * N * NU;
    const double* dtn = a.dtn ? a.dtn + (size_t)b * N : nullptr;

    double q[NQ], v[NQ], qr[NQ], vr[NQ], acc[NQ], ff[12], qm[NQ], vm[NQ], tq[NQ - 6];
    nmpc_plan_sample(xb, ub, dtn, a.dt, N, n, ip, q, v);
    // The PD term tracks the NEXT plan sample: `_compute_torques_ff` uses a_plan / f_plan[plan_step], then `_step()`
    // increments plan_step, and `_compute_pd_torques` reads q_plan / v_plan[plan_step] (mpc.py:575-599)
    nmpc_plan_sample(xb, ub, dtn, a.dt, N, n, ip + 1 < n ? ip + 1 : n - 1, qr, vr);
    const double* ur = ub + (size_t)nmpc_plan_hold_node(N, n, ip) * NU;
#pragma unroll
    for (int c = 0; c < NQ; c++) acc[c] = ur[c];
#pragma unroll
    for (int c = 0; c < 12; c++) ff[c] = ur[NQ + c];
    if (a.qv_meas) {
        const double* m = a.qv_meas + (size_t)i * 36;
#pragma unroll
        for (int c = 0; c < NQ; c++) { qm[c] = m[c]; vm[c] = m[NQ + c]; }
    } else {
#pragma unroll
        for (int c = 0; c < NQ; c++) { qm[c] = q[c]; vm[c] = v[c]; }
    }
    nmpc_model_rnea(qm, vm, acc, ff, tq);
    if (a.q_plan) for (int c = 0; c < NQ; c++) a.q_plan[(size_t)i * NQ + c] = q[c];
    if (a.v_plan) for (int c = 0; c < NQ; c++) a.v_plan[(size_t)i * NQ + c] = v[c];
    for (int c = 0; c < NQ - 6; c++) {
        const double tpd = tq[c] + a.kp * (qr[6 + c] - qm[6 + c]) + a.kd * (vr[6 + c] - vm[6 + c]);
        if (a.tau_ff) a.tau_ff[(size_t)i * 12 + c] = tq[c];
        if (a.tau) a.tau[(size_t)i * 12 + c] = tpd;
        if (a.action) a.action[(size_t)i * 12 + c] = (tpd + a.kd * vm[6 + c]) / a.kp + qm[6 + c];
    }
}
