/* nmpc_b200.h -- C ABI of the B200-native batched quadruped NMPC solver (libnmpc_b200.so).
 *
 * Drop-in boundary for the one hot path of Atarilab/iterative_learning_NMPC:
 *   QuadrupedAcadosSolver.init()/solve()   (mpc_controller/utils/solver.py:356-429)
 * which, in the reference, goes through the missing `contact_tamp` class AcadosSolverHelper
 * (solver.py:7-8,15) -> acados_template.AcadosOcpSolver -> ctypes -> generated C solver
 * (acados SQP / SQP_RTI + HPIPM + BLASFEO + CasADi-generated model code).
 *
 * Every entry point names the reference interface it replaces.  All arrays are fp64, C order,
 * batch-first; pointers may be HOST or DEVICE (cudaMemcpyDefault); no torch types cross this boundary.
 *
 * Streams.  A handle owns two non-blocking CUDA streams: the solve stream (nmpc_stream) that every
 * entry point uses except nmpc_kinematics / nmpc_inverse_dynamics, which run on a private utility
 * stream with their own staging buffer.  Nothing is ordered against the caller's streams: a DEVICE
 * input must be complete before the call (the Python binding awaits torch's current stream), and
 * every getter synchronises its stream before returning, so outputs are complete on return.
 *
 * Threads.  The reference solves on a worker thread while the control thread evaluates id_torques
 * at 1 kHz on the same solver object (mpc.py:164,516,544-545).  The handle supports exactly that:
 * the solve path and the utility path are serialised by two independent locks, so concurrent calls
 * are safe and a utility call never waits for a queued solve.  Distinct handles are independent.
 * Return value: 0 on success, a negative NMPC_E* code for API misuse, a positive cudaError_t
 * for CUDA failures.  Per-problem solver outcome is reported in status[B] with acados numbering
 * (0 success, 1 NaN, 2 max-iter, 3 min-step, 4 QP failure).
 *
 * Fixed problem dimensions of this build (Go2 topology):
 *   nx = 42  x = [q(18) | v(18) | h(6)]                    (solver.py:185-187)
 *   nu = 30  u = [a(18) | f_FL(3) f_FR(3) f_RL(3) f_RR(3)] (solver.py:311-336,411-421)
 *   np = 44  per node, per foot: active, plane_point(3), plane_normal(3), p_gain, restrict,
 *            range_radius, peak                            (solver.py:212-252)
 *   ny = 76  y   = [base(12) | joint(24) | acc(12) | swing(4) | f_reg(4x3) | pos_cost(4x3)]
 *   nye= 52  y_e = [base(12) | joint(24) | swing(4) | pos_cost(4x3)]   (solver.py:108-141)
 *   ng = 42  g   = [friction 4x5 | contact 4x3 | patch 4x1 | centroidal consistency 6]
 */
#ifndef NMPC_B200_H
#define NMPC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NMPC_NX 42
#define NMPC_NU 30
#define NMPC_NZ 72
#define NMPC_NP 44
#define NMPC_NY 76
#define NMPC_NYE 52
#define NMPC_NG 42
#define NMPC_NP_FOOT 11
#define NMPC_MAX_NODES 64

/* error codes (negative); positive returns are cudaError_t */
#define NMPC_OK 0
#define NMPC_EINVAL (-1)   /* bad argument / null pointer */
#define NMPC_ENOSYS (-2)   /* option not available in this build */
#define NMPC_ENODEV (-3)   /* no CUDA device / wrong architecture */
#define NMPC_ENOMEM (-4)
#define NMPC_E2BIG (-5)    /* request beyond a fixed capacity (nmpc_kinematics_static: n > NMPC_STATIC_MAX_N) */

/* per-problem solver status, acados numbering */
#define NMPC_STATUS_SUCCESS 0
#define NMPC_STATUS_NAN 1
#define NMPC_STATUS_MAXITER 2
#define NMPC_STATUS_MINSTEP 3
#define NMPC_STATUS_QP_FAILURE 4

/* HPIPM_MODE as used by MPCOptConfig.hpipm_mode (config/config_abstract.py:5,64) */
#define NMPC_HPIPM_SPEED_ABS 0
#define NMPC_HPIPM_SPEED 1
#define NMPC_HPIPM_BALANCE 2
#define NMPC_HPIPM_ROBUST 3

typedef struct nmpc_handle nmpc_handle;

/* Solver options.  Replaces AcadosSolverHelper.setup(recompile, use_cython, real_time_it,
 * max_qp_iter, hpipm_mode) and set_max_iter / set_warm_start_inner_qp / set_warm_start_nlp /
 * set_qp_tol / set_nlp_tol (solver.py:68-79; also mpc.py:464-473). */
typedef struct nmpc_opts {
    int32_t n_nodes;          /* N, shooting intervals (MPCOptConfig.n_nodes = 25)          */
    int32_t max_iter;         /* SQP iterations (1; 15 on the first tick, mpc.py:465-467)   */
    int32_t max_qp_iter;      /* IPM iterations per QP (6, mpc_opt.py:27)                   */
    int32_t real_time_it;     /* 1: SQP_RTI (one linearise + one QP + full step)            */
    int32_t warm_start_qp;    /* acados qp_solver_warm_start = HPIPM warm_start level: 0 cold, 1 primal
                                 iterate of the previous QP (pi, lam, t re-initialised), 2 primal and
                                 dual ("hot": lam, t kept, clipped at 0.1 from below).  The reference's
                                 `warm_start_qp = True` (config_abstract.py:62) maps to 1             */
    int32_t warm_start_nlp;   /* kept for API parity; the iterate always lives in the handle */
    int32_t hpipm_mode;       /* NMPC_HPIPM_*; this build implements SPEED                  */
    int32_t enable_time_opt;  /* 1: the node durations dt_k are optimised as well (nu = 31: cost weight
                                 time_opt_weight on dt - dt_nodes, bounds dt_min <= dt_k <= dt_max)   */
    int32_t warm_start_first_qp; /* acados `warm_start_first_qp` (default 0): the first QP of every
                                 nmpc_solve call is cold-started whatever warm_start_qp says       */
    int32_t cond_pred_corr;   /* HPIPM `cond_pred_corr` (1 in every HPIPM mode): fall back to the
                                 centring direction when the corrector predicts mu > 2 mu_aff      */
    double dt_nodes;          /* MPCOptConfig.get_dt_nodes()                                */
    double dt_min, dt_max;    /* MPCOptConfig.get_dt_bounds(); used when enable_time_opt    */
    double nlp_tol;           /* SQP exit: all four residual inf-norms below (1e-1)         */
    double qp_tol;            /* IPM exit: all four residual inf-norms below (1e-2)         */
    double reg_eps, reg_eps_e;/* Levenberg-Marquardt terms (MPCCostConfig.reg_eps[_e])      */
    double mu_contact;        /* friction coefficient (0.8 hard-coded at solver.py:38)      */
    double time_opt_weight;   /* MPCCostConfig.time_opt (1e4, mpc_cost.py:69): W["dt"], solver.py:124-125 */
} nmpc_opts;

/* Per-problem statistics of the last nmpc_solve (acados print_stats / get_stats equivalent). */
typedef struct nmpc_stats {
    int32_t* sqp_iter;   /* [B]   SQP iterations performed                          */
    int32_t* qp_iter;    /* [B]   IPM iterations of the last QP                     */
    int32_t* qp_status;  /* [B]   HPIPM numbering: 0 ok, 1 max-iter, 2 min-step, 3 NaN */
    double* res;         /* [B,4] NLP residual inf-norms stat, eq, ineq, comp at the last linearisation */
    double* qp_res;      /* [B,4] QP residual inf-norms at IPM exit                 */
    double* cost;        /* [B]   NLS cost at the last linearisation point          */
    int32_t* qp_ncond;   /* [B]   IPM iterations of the last QP in which the conditional predictor-corrector
                                  replaced the corrector direction by the centring direction */
} nmpc_stats;

/* ---- lifetime ------------------------------------------------------------------------- */
/* AcadosSolverHelper.__init__ + setup() (solver.py:49-55,68-72).  NULL on failure; *err (may
 * be NULL) receives the code. */
nmpc_handle* nmpc_create(const nmpc_opts* opts, int32_t batch, int32_t device, int32_t* err);
void nmpc_destroy(nmpc_handle* h);
int nmpc_set_opts(nmpc_handle* h, const nmpc_opts* opts);        /* set_max_iter ... set_nlp_tol */
int nmpc_get_dims(const nmpc_handle* h, int32_t* batch, int32_t* n_nodes);
int nmpc_stream(const nmpc_handle* h, void** cuda_stream);        /* the handle's own stream */

/* ---- problem data (QuadrupedAcadosSolver.init, solver.py:356-394) --------------------- */
/* setup_initial_state + set_initial_state (solver.py:179-192).  qv = [B,36] (q_euler | v);
 * the centroidal momentum h0 = hg(q,v) (dynamics.py:48-50, pin.computeCentroidalMomentum) is
 * evaluated on the device.  Also overwrites node 0 of the iterate with x0. */
int nmpc_set_x0(nmpc_handle* h, const double* qv);
/* same, but the caller supplies the full state [B,42] including h */
int nmpc_set_x0_full(nmpc_handle* h, const double* x0);
/* update_ref + set_ref_terminal (solver.py:352-353): yref [B,N,76], yref_e [B,52] */
int nmpc_set_yref(nmpc_handle* h, const double* yref, const double* yref_e);
/* set_cost_weight_constant / set_cost_weight_terminal / update_cost_weights (solver.py:85,
 * 140-141): diagonal weights; per_problem=0 -> W [76], W_e [52] broadcast; 1 -> [B,76],[B,52] */
int nmpc_set_weights(nmpc_handle* h, const double* W, const double* W_e, int32_t per_problem);
/* update_parameters (solver.py:351): p [B,N+1,44] */
int nmpc_set_params(nmpc_handle* h, const double* p);
/* QuadrupedAcadosSolver.init with the expansion done on the device (replaces the host loops of
 * setup_reference solver.py:153-177, init_contacts_parameters :212-225, setup_cnt_status :227-252,
 * setup_contact_loc :254-275 and the node-by-node update_solver :344-353).  Compact planner outputs:
 * base_ref, base_ref_e [B,12] (base_ref_e NULL = base_ref); joint_ref [B,12] or [12]; step_height [B] or
 * [1]; cnt_sequence int8 [B,4,N+1]; swing_peak int8 [B,4,N+1] or NULL (opt_peak off); cnt_locations
 * [B,4,N+1,3] or NULL; p_gain [4] = W_foot_pos_constr_stab; cnt_radius, height_offset; restrict_cnt as
 * set_contact_restriction (NMPC_EINVAL if set without locations).  Writes p, yref, yref_e of the handle. */
int nmpc_init_tick(nmpc_handle* h, const double* base_ref, const double* base_ref_e, const double* joint_ref,
                   int32_t joint_per_problem, const double* step_height, int32_t height_per_problem,
                   const int8_t* cnt_sequence, const int8_t* swing_peak, const double* cnt_locations,
                   const double* p_gain, double cnt_radius, double height_offset, int32_t restrict_cnt);
/* The planning step of LocomotionMPC.optimize followed by QuadrupedAcadosSolver.init, on the device, for the cyclic
 * contact planner: GaitPlanner.get_contacts / get_peaks (contact_planner.py:121-149: the window
 * of the gait cycle at i_node[b]; gait_seq / peak_seq int8 [4, nodes_per_cycle]) and compute_base_ref_vel_tracking
 * (mpc.py:210-272: q [B,18] Euler state, v_des [B,3], w_yaw [B]), then the expansion of nmpc_init_tick without
 * contact restriction.  The tracked reference pose base_ref_vel_tracking [B,12] is resident: `tracking` != NULL
 * replaces it, and it is first advanced by `advance_steps` calls of increment_base_ref_position (mpc.py:204-208,
 * one per sim step of length sim_dt since the last tick).  0.2 KB per problem cross PCIe. */
int nmpc_plan_tick(nmpc_handle* h, const int32_t* i_node, const int8_t* gait_seq, const int8_t* peak_seq,
                   int32_t nodes_per_cycle, const double* q, const double* v_des, const double* w_yaw,
                   const double* tracking, int32_t advance_steps, double sim_dt, double nom_height,
                   double height_offset, double t_horizon, const double* joint_ref, double step_height,
                   const double* p_gain, int32_t opt_peak);
/* Parameters of RaiberContactPlanner (contact_planner.py:190-233) and the contact-patch radius of the cost config. */
typedef struct nmpc_raibert_params {
    double offset_hip_b[12];   /* [4][3] hip offsets in the base frame, x_offset / y_offset applied, z = 0 */
    double stance_ratio[4];    /* GaitConfig.stance_ratio */
    double nominal_period;     /* GaitConfig.nominal_period [s] */
    double foot_size;          /* z of a planned contact */
    double height_offset;      /* subtracted from the CoM height */
    double gravity;            /* RaiberContactPlanner.GRAVITY = 9.81 */
    double cnt_radius;         /* cost config cnt_radius: half-width of the contact patch (solver.py:227-252) */
} nmpc_raibert_params;
/* LocomotionMPC.optimize with contact_planner="raibert", on the device (BASELINE config 4):
 * update_pin's centre of mass, GaitPlanner.get_contacts / get_peaks, RaiberContactPlanner.get_locations with
 * cache_cnt=False as LocomotionMPC builds it (contact_planner.py:265-322; mpc.py:76-90),
 * compute_base_ref_cnt_restricted (mpc.py:274-315), then the expansion of nmpc_init_tick with restrict_cnt = 1.
 * make_seq int8 [4, nodes_per_cycle] is 1 on the cycle node where a foot touches down (GaitPlanner.switch_cnt == 1);
 * qv [B,36] Euler state; t_cnt [N+1] = round(j * dt_nodes, 3) as the host computes it; cnt_locations, if not NULL,
 * receives the planned footholds [B,4,N+1,3] (the call then synchronises). */
int nmpc_plan_tick_raibert(nmpc_handle* h, const int32_t* i_node, const int8_t* gait_seq, const int8_t* peak_seq,
                           const int8_t* make_seq, int32_t nodes_per_cycle, const double* qv, const double* v_des,
                           const double* w_yaw, const double* t_cnt, const nmpc_raibert_params* rp, double nom_height,
                           double height_offset, const double* joint_ref, double step_height, const double* p_gain,
                           int32_t opt_peak, double* cnt_locations);
/* base_ref_vel_tracking [B,12] as held (and advanced) on the device */
int nmpc_get_tracking(nmpc_handle* h, double* tracking);
/* cold first guess of the batched front-end: x0 held over the horizon, zero accelerations, the robot's
 * weight shared by the stance feet of every node (from x0 and p on the device) */
int nmpc_cold_guess(nmpc_handle* h);
/* problem data as held on the device: p [B,N+1,44], yref [B,N,76], yref_e [B,52], x0 [B,42]; NULLs skipped */
int nmpc_get_problem_data(nmpc_handle* h, double* p, double* yref, double* yref_e, double* x0);
/* setup_initial_feet_pos (solver.py:194-210) done on the device: for every foot whose node-0
 * `active` is 1 (forced to 1 where i_node[b]==0), plane_point[:, :next_swing] = FK foot position
 * of the current x0.  i_node [B] int32 (host or device). Call after set_params and set_x0. */
int nmpc_pin_stance_feet(nmpc_handle* h, const int32_t* i_node);
/* update_states / update_inputs (solver.py:349-350): overwrite the iterate.  x [B,N+1,42] and/or
 * u [B,N,30]; either may be NULL. */
int nmpc_set_guess(nmpc_handle* h, const double* x, const double* u);
/* enable_time_opt: inputs["dt"] (solver.py:189-190, 325-336), the node durations of the iterate, dt [B,N]; NULL
 * resets them to dt_nodes (nmpc_set_guess with inputs and nmpc_cold_guess do the same) */
int nmpc_set_dt(nmpc_handle* h, const double* dt);
/* warm_start_solver + warm_start_multipliers (solver.py:290-342): shift the stored iterate of
 * problem b by start_node[b] nodes (states 1.., inputs, duals), zero the force tail (or repeat
 * the last node when repeat_last). start_node [B] int32. */
int nmpc_shift_warm_start(nmpc_handle* h, const int32_t* start_node, int32_t repeat_last);
/* forget duals and the stored QP solution (AcadosSolverHelper.reset-like) */
int nmpc_reset_duals(nmpc_handle* h);

/* ---- checkpoint / resume ------------------------------------------ */
/* The state a warm-started solve depends on -- iterate (x, u, pi, lam, dt), the stored QP solution, problem data --
 * as one host blob of nmpc_state_bytes(h) bytes; import needs a handle of the same batch, horizon and nu. */
int64_t nmpc_state_bytes(nmpc_handle* h);
int nmpc_export_state(nmpc_handle* h, void* blob);
int nmpc_import_state(nmpc_handle* h, const void* blob);

/* ---- batched rigid-body rollout ---------------------------------- */
/* Closes the loop around the resident plans without MuJoCo: what Rollout_MPC.run / Rollout_combined_controller do one
 * rollout at a time (Behavior_Cloning/utils/Rollout_MPC.py:465, DAgger/utils/Rollout_combined_controller.py:484-499)
 * -- every sim step LocomotionMPC.compute_torques_dof (mpc.py:503-599: feed-forward inverse dynamics on the measured
 * state + PD on the plan) and one physics step -- for the whole batch on the device.  Physics: articulated-body
 * forward dynamics of the URDF model in the solver's Euler coordinates, compliant point-foot contact with the plane
 * z = ground_z (unilateral normal spring-damper, viscous tangential force inside the friction cone), semi-implicit
 * Euler.  It is NOT MuJoCo's contact solver. */
typedef struct nmpc_sim_params {
    double sim_dt;          /* simulation step [s] (1e-3)                                     */
    double kp, kd;          /* PD gains (MPCOptConfig.Kp, Kd)                                 */
    double k_n, d_n;        /* contact: normal stiffness [N/m], normal damping [N s/m]        */
    double d_t;             /* contact: tangential viscous gain [N s/m]                       */
    double mu;              /* friction coefficient (cone cap of the tangential force)        */
    double ground_z;        /* ground plane height                                            */
    double tau_max;         /* joint torque limit, 0 = none                                   */
} nmpc_sim_params;
/* simulation state of every rollout, qv [B,36] = [q_euler 18 | v 18] */
int nmpc_sim_set_state(nmpc_handle* h, const double* qv);
int nmpc_sim_get_state(nmpc_handle* h, double* qv, int32_t* diverged /* [B] or NULL */);
/* n_steps simulation steps on the plan of the last nmpc_solve, plan samples step0 .. step0+n_steps-1 of the
 * n_interp-sample plan.  Optional outputs (host or device, NULL to skip): out_qv [B,n_steps,36] the state the
 * controller saw at every step, out_tau [B,n_steps,12] the applied torques, out_fc [B,n_steps,12] contact forces. */
int nmpc_sim_steps(nmpc_handle* h, const nmpc_sim_params* p, int32_t n_interp, int32_t step0, int32_t n_steps,
                   double* out_qv, double* out_tau, double* out_fc);

/* ---- solve (AcadosSolverHelper.solve, called at solver.py:402) ------------------------ */
/* Enqueues the whole SQP / SQP_RTI solve on the handle's stream and, if status != NULL, copies
 * the per-problem status [B] back and synchronises.  status may be host or device. */
int nmpc_solve(nmpc_handle* h, int32_t* status);
int nmpc_synchronize(nmpc_handle* h);

/* ---- results (parse_sol, solver.py:403-429) ------------------------------------------- */
/* x [B,N+1,42], u [B,N,30], pi [B,N,42] (multiplier of x_{k+1} = phi(x_k,u_k)),
 * lam [B,N+1,84] (lower | upper multipliers of g).  Any pointer may be NULL. */
int nmpc_get_solution(nmpc_handle* h, double* x, double* u, double* pi, double* lam);
/* parse_sol hidden behind the solve (solver.py:397-429: `solve` = solver call + parse_sol): destinations x [B,N+1,42] and
 * u [B,N,30] (either may be NULL; NULL, NULL switches it off) that every following nmpc_solve fills with the final
 * iterate.  The QP kernel is launched wave by wave; as soon as a wave's last SQP step is applied its rows are copied,
 * on a separate copy stream, while the next wave is still being solved.  The copies are complete when the solve stream
 * is: after nmpc_solve with status != NULL, nmpc_get_status, nmpc_synchronize.  For host destinations use pinned memory
 * (pageable memory works but the copies then block the host).  The buffers must stay valid until nmpc_set_output is
 * called again or the handle is destroyed. */
int nmpc_set_output(nmpc_handle* h, double* x, double* u);
/* dt_node_sol (solver.py:423-427): dt [B,N] -- dt_nodes everywhere unless enable_time_opt -- and the multipliers of
 * dt_min <= dt <= dt_max, lam_dt [B,N,2] (lower | upper).  Either pointer may be NULL. */
int nmpc_get_dt(nmpc_handle* h, double* dt, double* lam_dt);
int nmpc_get_stats(nmpc_handle* h, nmpc_stats* stats);
/* per-problem status [B] int32 of the last nmpc_solve (host or device pointer) */
int nmpc_get_status(nmpc_handle* h, int32_t* status);
/* Packed result rows for the one collective of the use case (Rollout_MPC.py:275,298: every rollout's
 * plan ends up with the dataset writer): out [B, (N+1)*42 + N*30 + N + 1] = per problem x | u | dt_node | status,
 * i.e. the 1867 numbers `parse_sol` returns at N = 25 (solver.py:403-429) plus the status, written into a DEVICE
 * buffer by a kernel on the solve stream (no synchronisation, no host round trip): hand it to ncclAllGather. */
int nmpc_pack_solution(nmpc_handle* h, double* out_device);
/* QP step of the last SQP iteration: dz [B,N+1,72] ([du|dx] per node; du of node N is 0) */
int nmpc_get_qp_step(nmpc_handle* h, double* dz);

/* ---- kinematics helper (QuadrupedDynamics.update_pin/get_feet_position_w, dynamics.py:48-50,
 * 101-107; pin.centerOfMass at mpc.py:339) ---------------------------------------------- */
/* qv [n,36] -> feet [n,4,3], com [n,3], hg [n,6]; outputs may be NULL. */
int nmpc_kinematics(nmpc_handle* h, const double* qv, int32_t n, double* feet, double* com, double* hg);

/* Allocation-free variant of nmpc_kinematics.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  No handle: it runs on the CURRENT device, on `stream` (a cudaStream_t; NULL = the
 * default stream), and calls no cudaMalloc / cudaFree / allocating library routine.  Its staging buffer is a file-scope
 * device array of NMPC_STATIC_MAX_N instances x 57 doubles (1.78 MiB) that every GPU which loads the library holds
 * until the process exits; n > NMPC_STATIC_MAX_N is rejected with NMPC_E2BIG and a message on stderr (split the batch,
 * or use nmpc_kinematics).  The first call on a device resolves the array's device address (cudaGetSymbolAddress):
 * make that call outside stream capture.  Same arguments and results as nmpc_kinematics: qv [n,36] -> feet [n,4,3],
 * com [n,3], hg [n,6]; outputs may be NULL; pointers host or device.  Outputs are complete on return (the stream is
 * synchronised) unless `stream` is being captured.
 * SHARED STATE: every call on a device uses the same array and finds the previous call's contents in it (nothing has
 * to start at zero: the kernel overwrites all it reads back).  Calls of one process are serialised by a lock up to
 * their launch, but two calls whose GPU work overlaps (different streams, or captured graphs replayed concurrently)
 * RACE on the array: order them. */
#define NMPC_STATIC_MAX_N 4096
int nmpc_kinematics_static(const double* qv, int32_t n, double* feet, double* com, double* hg, void* stream);

/* QuadrupedDynamics.id_torques (dynamics.py:136-163): tau [n,12] = rnea(q, v, a)[joints] -
 * sum_feet J_foot(q)^T f.  qva [n,54] = (q_euler | v | a), f [n,12] world-frame contact forces. */
int nmpc_inverse_dynamics(nmpc_handle* h, const double* qva, const double* f, int32_t n, double* tau);

/* ---- plan -> torque, batched.  Replaces, for every problem of the batch at once:
 * LocomotionMPC.interpolate_state_trajectory (mpc.py:371-414: cubic Hermite of (q,v) and (v,a) on the node
 * grid, sampled at linspace(0, T, n_interp+1)[1:]), the zero-order hold a_plan/f_plan = sol[id_repeat]
 * (mpc.py:110,544-545), QuadrupedDynamics.id_torques (dynamics.py:136-163), _compute_pd_torques
 * (mpc.py:592-599) and the DAgger label action = (tau + kd v)/kp + q (Behavior_Cloning/utils/Rollout_MPC.py:298).
 * Works on the solution resident in the handle (after nmpc_solve).  Plan steps step0 .. step0+n_steps-1 of the
 * n_interp-sample plan.  qv_meas [B,n_steps,36] = measured (q_euler | v) at each step, or NULL to evaluate on
 * the plan itself (open-loop labels).  Outputs (host or device, any may be NULL): q_plan, v_plan
 * [B,n_steps,18]; tau_ff, tau, action [B,n_steps,12]. */
int nmpc_plan_torques(nmpc_handle* h, const double* qv_meas, int32_t n_interp, int32_t step0, int32_t n_steps,
                      double kp, double kd, double* q_plan, double* v_plan, double* tau_ff, double* tau,
                      double* action);

/* ---- test / diagnostics hooks (parity tests compare these with the oracle) ------------ */
/* Linearise at the current iterate without solving and export the dense OCP-QP:
 * BAbt [B,N,73,42] (rows: B^T(30) | A^T(42) | b^T), RSQrq [B,N+1,73,72] (lower part valid; last
 * row = gradient), DCt [B,N+1,72,42], lg/ug [B,N+1,42], mask [B,N+1,84].  NULLs skipped. */
int nmpc_export_qp(nmpc_handle* h, double* BAbt, double* RSQrq, double* DCt,
                   double* lg, double* ug, double* mask);
/* Solve externally supplied OCP-QPs (same layouts as nmpc_export_qp) with the device IPM;
 * dx0 [B,42]; outputs dz [B,N+1,72], pi [B,N,42], lam [B,N+1,84], t [B,N+1,84],
 * qp_iter/qp_status [B] int32.  Cold start (warm_start 0).  The six stage-block arrays must be HOST
 * pointers (they are repacked on the host; NMPC_EINVAL otherwise).  NMPC_ENOSYS if the QP lacks the
 * structure of this OCP family (see nmpc_last_qp_path). */
int nmpc_solve_qp(nmpc_handle* h, const double* BAbt, const double* RSQrq, const double* DCt,
                  const double* lg, const double* ug, const double* mask, const double* dx0,
                  double* dz, double* pi, double* lam, double* t,
                  int32_t* qp_iter, int32_t* qp_status);

/* outcome of the structure check of the last nmpc_solve_qp call: 1 = solved by the fused structured kernel
 * (the product path); -1 = refused with NMPC_ENOSYS because the supplied QP does not have the structure of
 * this OCP family (this build has no generic dense QP solver) */
int nmpc_last_qp_path(const nmpc_handle* h);

/* Per-kernel device timing with CUDA events recorded on the handle's stream around the launches,
 * grouped by kind.  Replaces acados' print_time / the reference's @time_fn("solve")
 * (solver.py:396) at kernel granularity.  ms / count have NMPC_PROF_KINDS entries. */
#define NMPC_PROF_LINEARIZE 0   /* model evaluation + stage assembly                    */
#define NMPC_PROF_QP_STAGE 1    /* unused (split-kernel IPM removed)                    */
#define NMPC_PROF_RESIDUAL 2    /* NLP residual norms (full SQP only)                   */
#define NMPC_PROF_QP_FACTOR 3   /* unused                                               */
#define NMPC_PROF_QP_CORR 4     /* unused                                               */
#define NMPC_PROF_QP_OTHER 5    /* SQP full step                                        */
#define NMPC_PROF_QP_FUSED 6    /* the fused per-OCP interior-point kernel (product path) */
#define NMPC_PROF_KINDS 7
int nmpc_set_profiling(nmpc_handle* h, int32_t on);
int nmpc_get_kernel_times(nmpc_handle* h, double* ms /*[NMPC_PROF_KINDS]*/, int64_t* count /*[NMPC_PROF_KINDS]*/);

/* number of kernels launched by the handle since creation (bench.py's gpu_launches claim) */
int64_t nmpc_launch_count(const nmpc_handle* h);
const char* nmpc_version(void);
/* total mass of the compiled robot model (kg): pin_model mass, used for force guesses */
double nmpc_model_mass(void);

#ifdef __cplusplus
}
#endif
#endif /* NMPC_B200_H */
