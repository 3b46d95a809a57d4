// nmpc_init.cuh -- QuadrupedAcadosSolver.init on the device (8f rank 2).
//
// In the reference, `init` expands a handful of per-problem planner outputs -- base references, the
// contact mask, swing peaks, optional contact locations -- into per-node dictionaries and pushes them
// into acados node by node (`update_solver`, solver.py:344-353: ~40 KB per solve through Python).
// Here the compact planner outputs cross PCIe (0.5 KB per problem instead of 41 KB) and the expansion
// into the node-major parameter / reference records is done by the kernels below:
//   setup_reference           solver.py:153-177   -> yref [B,N,76], yref_e [B,52]
//   init_contacts_parameters  solver.py:212-225 } -> p [B,N+1,44]
//   setup_cnt_status          solver.py:227-252 }
//   setup_contact_loc         solver.py:254-275   -> plane_point in p, foot targets in yref / yref_e
// plus the cold first guess the batched front-end uses (x0 held over the horizon, the robot's weight
// shared by the stance feet of each node).
#pragma once
#include "nmpc_common.cuh"

// y layout (include/nmpc_b200.h): base 12 | joint 24 | acc 12 | swing 4 | f_reg 12 | foot pos 12
#define Y_BASE 0
#define Y_JOINT 12
#define Y_SWING 48
#define Y_POS 64
#define YE_BASE 0
#define YE_JOINT 12
#define YE_SWING 36
#define YE_POS 40

struct TickArgs {
    const double* base_ref;      // [B,12]
    const double* base_ref_e;    // [B,12]
    const double* joint_ref;     // [B,12] or [12]
    const double* step_height;   // [B] or [1]
    const signed char* cnt;      // [B,4,N+1]
    const signed char* peak;     // [B,4,N+1] or nullptr (opt_peak off)
    const double* loc;           // [B,4,N+1,3] or nullptr
    double p_gain[NF];
    double radius, height_offset;
    int restrict_cnt, joint_per_problem, height_per_problem;
    double* p; double* yref; double* yref_e;
    int B, N;
};

__global__ void __launch_bounds__(128) nmpc_init_tick_kernel(TickArgs a)
{
    const int b = blockIdx.x, tid = threadIdx.x, N = a.N;
    const double* br = a.base_ref + (size_t)b * 12;
    const double* bre = a.base_ref_e + (size_t)b * 12;
    const double* jr = a.joint_ref + (a.joint_per_problem ? (size_t)b * 12 : 0);
    const double sh = a.step_height[a.height_per_problem ? b : 0];
    const signed char* cnt = a.cnt + (size_t)b * NF * (N + 1);
    const signed char* peak = a.peak ? a.peak + (size_t)b * NF * (N + 1) : nullptr;
    const double* loc = a.loc ? a.loc + (size_t)b * NF * (N + 1) * 3 : nullptr;
    const bool use_loc = a.restrict_cnt && loc;

    // ---- running references ----
    double* yr = a.yref + (size_t)b * N * NY;
    for (int e = tid; e < N * NY; e += 128) {
        const int k = e / NY, c = e % NY;
        double v = 0.0;
        if (c < Y_JOINT) v = br[c];
        else if (c < Y_JOINT + 12) v = jr[c - Y_JOINT];
        else if (c >= Y_SWING && c < Y_SWING + NF) v = sh;
        else if (c >= Y_POS && loc) { const int f = (c - Y_POS) / 3, i = (c - Y_POS) % 3; v = loc[((size_t)f * (N + 1) + k + 1) * 3 + i]; }
        yr[e] = v;
    }
    // ---- terminal references ----
    double* ye = a.yref_e + (size_t)b * NYE;
    for (int c = tid; c < NYE; c += 128) {
        double v = 0.0;
        if (c < YE_JOINT) v = bre[c];
        else if (c < YE_JOINT + 12) v = jr[c - YE_JOINT];
        else if (c >= YE_SWING && c < YE_SWING + NF) v = sh;
        else if (c >= YE_POS && loc) { const int f = (c - YE_POS) / 3, i = (c - YE_POS) % 3; v = loc[((size_t)f * (N + 1) + N) * 3 + i]; }
        ye[c] = v;
    }
    // ---- per-node, per-foot parameters ----
    double* pb = a.p + (size_t)b * (N + 1) * NPAR;
    for (int e = tid; e < (N + 1) * NF; e += 128) {
        const int k = e / NF, f = e % NF;
        double* o = pb + (size_t)k * NPAR + f * 11;
        const int act = cnt[f * (N + 1) + k];
        o[P_ACTIVE] = (double)act;
        o[P_POINT] = use_loc ? loc[((size_t)f * (N + 1) + k) * 3] : 0.0;
        o[P_POINT + 1] = use_loc ? loc[((size_t)f * (N + 1) + k) * 3 + 1] : 0.0;
        o[P_POINT + 2] = use_loc ? loc[((size_t)f * (N + 1) + k) * 3 + 2] : a.height_offset;
        o[P_NORMAL] = 0.0; o[P_NORMAL + 1] = 0.0; o[P_NORMAL + 2] = 1.0;
        o[P_GAIN] = a.p_gain[f];
        int made = 0;                                                   // 1 on the node where the foot touches down
        if (a.restrict_cnt && k > 0) { made = act - (int)cnt[f * (N + 1) + k - 1]; if (made < 0) made = 0; }
        o[P_RESTRICT] = (double)made;
        o[P_RADIUS] = a.restrict_cnt ? a.radius : 1.0e10;
        o[P_PEAK] = peak ? (double)peak[f * (N + 1) + k] : 0.0;
    }
}

// cold first guess: x_k = x0, a_k = 0, vertical forces = weight / number of stance feet of node k
__global__ void __launch_bounds__(128) nmpc_cold_guess_kernel(int N, const double* x0, const double* p, double* x, double* u)
{
    const int b = blockIdx.x, tid = threadIdx.x;
    const double* x0b = x0 + (size_t)b * NX;
    const double* pb = p + (size_t)b * (N + 1) * NPAR;
    double* xb = x + (size_t)b * (N + 1) * NX;
    double* ub = u + (size_t)b * N * NU;
    for (int e = tid; e < (N + 1) * NX; e += 128) xb[e] = x0b[e % NX];
    for (int e = tid; e < N * NU; e += 128) {
        const int k = e / NU, c = e % NU;
        double v = 0.0;
        if (c >= IFRC && (c - IFRC) % 3 == 2) {
            const double* pk = pb + (size_t)k * NPAR;
            const double n_act = pk[P_ACTIVE] + pk[11 + P_ACTIVE] + pk[22 + P_ACTIVE] + pk[33 + P_ACTIVE];
            v = pk[((c - IFRC) / 3) * 11 + P_ACTIVE] * (NMPC_MASS * 9.81) / (n_act > 1.0 ? n_act : 1.0);
        }
        ub[e] = v;
    }
}

// ---------------------------------------------------------------------------------------------
// Planning on the device: what LocomotionMPC.optimize computes on the host before `init`
// for the cyclic contact planner --
//   GaitPlanner.get_contacts / get_peaks   contact_planner.py:121-149   window of the gait cycle at i_node
//   compute_base_ref_vel_tracking          mpc.py:210-272               base references from q, v_des, w_yaw
// -- into the compact arrays nmpc_init_tick_kernel expands.  One CTA per OCP.
// ---------------------------------------------------------------------------------------------
struct PlannerArgs {
    const int32_t* i_node;           // [B] optimisation node of every problem
    const signed char* gait_seq;     // [4][npc] stance mask of one gait cycle
    const signed char* peak_seq;     // [4][npc] swing mask of one gait cycle
    const double* q;                 // [B][18] Euler state
    const double* v_des;             // [B][3]  velocity command, base frame
    const double* w_yaw;             // [B]
    double* tracking;                // [B][12] integrated reference pose (base_ref_vel_tracking), advanced in place
    int advance_steps; double sim_dt;// increment_base_ref_position (mpc.py:204-208) for the sim steps since the last tick
    double nom_height, height_offset, t_horizon;
    int npc, N, B;
    double* base_ref; double* base_ref_e;       // [B][12]
    signed char* cnt; signed char* peak;        // [B][4][N+1]
};

// numpy.round(x, d) = rint(x * 10^d) / 10^d; Python's round(x, 1) is the correctly rounded decimal, which differs
// from that when the product x * 10 rounds onto a tie: the fused multiply-add recovers the exact product
__device__ __forceinline__ double np_round(double x, double p10) { return rint(x * p10) / p10; }
__device__ __forceinline__ double py_round1(double x)
{
    const double p = x * 10.0, err = fma(x, 10.0, -p);
    double r = rint(p);
    if (fabs(p - trunc(p)) == 0.5 && err != 0.0) r = (err > 0.0) ? ceil(p) : floor(p);
    return r / 10.0;
}
__device__ __forceinline__ double clip(double v, double lo, double hi) { return fmin(fmax(v, lo), hi); }

__global__ void __launch_bounds__(128) nmpc_plan_refs_kernel(PlannerArgs a)
{
    const int b = blockIdx.x, tid = threadIdx.x, N1 = a.N + 1;
    const int node = a.i_node[b];
    for (int e = tid; e < NF * N1; e += 128) {
        const int f = e / N1, j = e % N1;
        const int idx = (node % a.npc + j) % a.npc;
        a.cnt[(size_t)b * NF * N1 + e] = a.gait_seq[f * a.npc + idx];
        a.peak[(size_t)b * NF * N1 + e] = a.peak_seq[f * a.npc + idx];
    }
    if (tid != 127) return;
    const double* q = a.q + (size_t)b * 18;
    double* tr = a.tracking + (size_t)b * 12;
    const double* vd = a.v_des + (size_t)b * 3;
    const double wz = a.w_yaw[b], th = a.t_horizon;
    if (a.advance_steps > 0) {
        // increment_base_ref_position, once per elapsed sim step, in the reference's order of operations
        const double p = tr[4], ro = tr[5];
        const double cr = cos(ro), sr = sin(ro), cp = cos(p), sp = sin(p);
        double px = tr[0], py = tr[1], yaw = tr[3];
        for (int s = 0; s < a.advance_steps; s++) {
            const double cy = cos(yaw), sy = sin(yaw);
            const double vx = np_round(((cy * cp) * vd[0] + (cy * sp * sr - sy * cr) * vd[1]) + (cy * sp * cr + sy * sr) * vd[2], 10.0);
            const double vy = np_round(((sy * cp) * vd[0] + (sy * sp * sr + cy * cr) * vd[1]) + (sy * sp * cr - cy * sr) * vd[2], 10.0);
            px += vx * a.sim_dt; py += vy * a.sim_dt;
            yaw += wz * a.sim_dt;
        }
        tr[0] = px; tr[1] = py; tr[3] = yaw;
    }
    double r[12], e[12];
    for (int i = 0; i < 12; i++) r[i] = 0.0;
    r[0] = np_round(q[0], 100.0); r[1] = np_round(q[1], 100.0);
    r[2] = a.nom_height + a.height_offset;
    r[3] = py_round1(q[3]);
    // R_WB = rpyToMatrix(roll, pitch, yaw) of the tracked pose (stored yaw, pitch, roll)
    {
        const double y = tr[3], p = tr[4], ro = tr[5];
        const double cr = cos(ro), sr = sin(ro), cp = cos(p), sp = sin(p), cy = cos(y), sy = sin(y);
        const double R[3][3] = {{cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr},
                                {sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr},
                                {-sp, cp * sr, cp * cr}};
        for (int i = 0; i < 3; i++) r[6 + i] = np_round((R[i][0] * vd[0] + R[i][1] * vd[1]) + R[i][2] * vd[2], 10.0);
    }
    r[9] = wz; r[10] = 0.0; r[11] = 0.0;             // w_des[::-1] with w_des = (0, 0, w_yaw)
    for (int i = 0; i < 12; i++) e[i] = r[i];
    {   // velocity at the end of the horizon: rpyToMatrix(w_des * t_horizon) = Rz(w_yaw t)
        const double cy = cos(wz * th), sy = sin(wz * th);
        e[6] = (cy * r[6] + -sy * r[7]) + 0.0 * r[8];
        e[7] = (sy * r[6] + cy * r[7]) + 0.0 * r[8];
        e[8] = (-0.0 * r[6] + 0.0 * r[7]) + 1.0 * r[8];
    }
    for (int i = 0; i < 2; i++) {
        e[i] = tr[i] + r[6 + i] * th;
        e[i] = clip(e[i], -r[i] + r[6 + i] * th * 1.2, r[i] + r[6 + i] * th * 1.2);
    }
    e[3] = tr[3] + wz * th;
    e[3] = clip(e[3], -tr[3] + wz * th * 1.5, tr[3] + wz * th * 1.5);
    r[0] += (e[0] - r[0]) * 0.75; r[1] += (e[1] - r[1]) * 0.75;
    r[3] += (e[3] - r[3]) * 0.75;
    e[8] = 0.0; e[4] = 0.0; e[5] = 0.0; r[4] = 0.0; r[5] = 0.0; e[10] = 0.0; e[11] = 0.0;
    for (int i = 0; i < 12; i++) { a.base_ref[(size_t)b * 12 + i] = r[i]; a.base_ref_e[(size_t)b * 12 + i] = e[i]; }
}

// ---------------------------------------------------------------------------------------------
// Raibert contact planner on the device (BASELINE config 4): what LocomotionMPC.optimize does on
// the host with contact_planner="raibert" --
//   RaiberContactPlanner.get_locations     contact_planner.py:265-322   one foothold per touch-down of the window,
//                                                                       held until the foot's next touch-down
//   compute_base_ref_cnt_restricted        mpc.py:274-315               base references centred on the first / last
//                                                                       most complete set of planned contacts
// The reference builds the planner with cache_cnt=False (mpc.py:89): the footholds are a function of the current state
// only, nothing is carried between ticks.  One CTA per OCP: the windows by all threads, one foot per thread for the
// footholds, one thread for the references.
// ---------------------------------------------------------------------------------------------
struct RaibertParams {
    double offset_hip_b[NF * 3];     // hip offsets in the base frame (x / y offsets already applied)
    double stance_ratio[NF];
    double nominal_period, foot_size, height_offset, gravity;
};

struct RaibertArgs {
    const int32_t* i_node;           // [B]
    const signed char* gait_seq;     // [4][npc]
    const signed char* peak_seq;     // [4][npc]
    const signed char* make_seq;     // [4][npc] 1 on the node of the cycle where the foot touches down
    const double* qv;                // [B][36] Euler state
    const double* com;               // [B][3]  centre of mass (nmpc_kin0_kernel)
    const double* v_des;             // [B][3]
    const double* w_yaw;             // [B]
    const double* t_cnt;             // [N+1]   round(j * dt_nodes, 3), as the host computes it
    RaibertParams rp;
    double nom_height, height_offset;
    int npc, N, B;
    double* base_ref; double* base_ref_e;       // [B][12]
    signed char* cnt; signed char* peak;        // [B][4][N+1]
    double* loc;                                // [B][4][N+1][3]
};

// lexicographic order of two [4][3] slabs of contact locations (numpy.unique(axis=1) sorts the nodes this way)
__device__ __forceinline__ int raibert_slab_cmp(const double* loc, int N1, int ja, int jb)
{
    for (int f = 0; f < NF; f++)
        for (int c = 0; c < 3; c++) {
            const double x = loc[((size_t)f * N1 + ja) * 3 + c], y = loc[((size_t)f * N1 + jb) * 3 + c];
            if (x < y) return -1;
            if (x > y) return 1;
        }
    return 0;
}

__global__ void __launch_bounds__(128) nmpc_plan_raibert_kernel(RaibertArgs a)
{
    const int b = blockIdx.x, tid = threadIdx.x, N1 = a.N + 1;
    const int node = a.i_node[b];
    for (int e = tid; e < NF * N1; e += 128) {
        const int f = e / N1, j = e % N1;
        const int idx = (node % a.npc + j) % a.npc;
        a.cnt[(size_t)b * NF * N1 + e] = a.gait_seq[f * a.npc + idx];
        a.peak[(size_t)b * NF * N1 + e] = a.peak_seq[f * a.npc + idx];
    }
    double* loc = a.loc + (size_t)b * NF * N1 * 3;
    const double* vd = a.v_des + (size_t)b * 3;
    if (tid < NF) {
        const int f = tid;
        const double* q = a.qv + (size_t)b * 36;
        const double* com = a.com + (size_t)b * 3;
        const double wz = a.w_yaw[b];
        const double yaw = q[3], cy = cos(yaw), sy = sin(yaw);
        const double* off = a.rp.offset_hip_b + 3 * f;
        // R_horizontal = rpyToMatrix(0, 0, yaw): rows x, y applied to the hip offset
        const double hx = (cy * off[0] + -sy * off[1]) + 0.0 * off[2];
        const double hy = (sy * off[0] + cy * off[1]) + 0.0 * off[2];
        const double ratio = a.rp.stance_ratio[f], t_stance = a.rp.nominal_period * ratio;
        const double com_z = com[2] - a.rp.height_offset;
        const double k_ang = 0.5 * sqrt(com_z / a.rp.gravity);
        // cross(k v, (0, 0, w)) = (k v_y w, -k v_x w)
        const double ax = (k_ang * vd[1]) * wz, ay = -((k_ang * vd[0]) * wz);
        double px = 0.0, py = 0.0, pz = 0.0;
        for (int j = 0; j < N1; j++) {
            const int idx = (node % a.npc + j) % a.npc;
            if (a.make_seq[f * a.npc + idx] == 1) {
                const double t = a.t_cnt[j];
                if (t >= 0.0) {
                    const double gx = (com[0] + hx) + (vd[0] * t) * (1.0 + ratio);
                    const double gy = (com[1] + hy) + (vd[1] * t) * (1.0 + ratio);
                    px = ((gx + 0.1 * (vd[0] - q[18])) + (0.5 * vd[0]) * t_stance) + ax;
                    py = ((gy + 0.1 * (vd[1] - q[19])) + (0.5 * vd[1]) * t_stance) + ay;
                    pz = a.rp.foot_size;
                }
            }
            double* o = loc + ((size_t)f * N1 + j) * 3;
            o[0] = px; o[1] = py; o[2] = pz;
        }
    }
    __syncthreads();
    if (tid != 127) return;
    // compute_base_ref_cnt_restricted: nodes with the most feet placed; among them the smallest and the largest slab
    int best = 0, jmin = -1, jmax = -1;
    for (int j = 0; j < N1; j++) {
        int c = 0;
        for (int f = 0; f < NF; f++) {
            const double* o = loc + ((size_t)f * N1 + j) * 3;
            c += (o[0] != 0.0 && o[1] != 0.0 && o[2] != 0.0);
        }
        if (c == 0) continue;
        if (c > best) { best = c; jmin = jmax = j; }
        else if (c == best) {
            if (raibert_slab_cmp(loc, N1, j, jmin) < 0) jmin = j;
            if (raibert_slab_cmp(loc, N1, j, jmax) > 0) jmax = j;
        }
    }
    if (jmin < 0) { jmin = 0; jmax = N1 - 1; }       // nothing planned: first / last node as they are
    double cf[2], cl[2];
    for (int c = 0; c < 2; c++) {
        cf[c] = (((loc[((size_t)0 * N1 + jmin) * 3 + c] + loc[((size_t)1 * N1 + jmin) * 3 + c]) + loc[((size_t)2 * N1 + jmin) * 3 + c]) + loc[((size_t)3 * N1 + jmin) * 3 + c]) / 4.0;
        cl[c] = (((loc[((size_t)0 * N1 + jmax) * 3 + c] + loc[((size_t)1 * N1 + jmax) * 3 + c]) + loc[((size_t)2 * N1 + jmax) * 3 + c]) + loc[((size_t)3 * N1 + jmax) * 3 + c]) / 4.0;
    }
    double* r = a.base_ref + (size_t)b * 12;
    double* e = a.base_ref_e + (size_t)b * 12;
    for (int i = 0; i < 12; i++) { r[i] = 0.0; e[i] = 0.0; }
    const double alpha = 0.35;
    r[0] = alpha * cf[0] + (1.0 - alpha) * cl[0];
    r[1] = alpha * cf[1] + (1.0 - alpha) * cl[1];
    e[0] = cl[0]; e[1] = cl[1];
    r[2] = a.nom_height + a.height_offset;
    e[2] = a.nom_height + a.height_offset;
}
