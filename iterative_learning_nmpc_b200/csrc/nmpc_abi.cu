// nmpc_abi.cu -- host side of libnmpc_b200.so: the C ABI declared in include/nmpc_b200.h.
//
// Owns the device buffers and one CUDA stream per handle; every entry point only enqueues
// copies / kernels on that stream (cudaMemcpyDefault lets callers pass host or device pointers).
// There is NO CPU fallback: creation fails with NMPC_ENODEV when no sm_100 device is present.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "nmpc_common.cuh"
#include "nmpc_linearize.cuh"
#include "nmpc_qp_layout.cuh"
#include "nmpc_qp_fused.cuh"
#include "nmpc_sqp.cuh"
#include "nmpc_linearize_sr.cuh"
#include "nmpc_plan.cuh"
#include "nmpc_init.cuh"
#include "nmpc_gen_sim.h"
#include "nmpc_rollout.cuh"

#define CK(call)                                                     \
    do {                                                             \
        cudaError_t e_ = (call);                                     \
        if (e_ != cudaSuccess) return (int)e_;                       \
    } while (0)

struct nmpc_handle {
    nmpc_opts o;
    int B, N, device, Bc;        // Bc: problems per chunk of QP workspace
    int n_sm, wave_plan;         // SMs of the device; 1: the fused QP kernel is launched wave by wave (run_qp_fused)
    // nmpc_set_output: host (or device) destinations the final x / u of every wave are copied to, on `cstream`, while the
    // next wave is still being solved
    double *out_x, *out_u;
    cudaStream_t cstream;
    cudaEvent_t ev_wave, ev_copied;
    int out_pending;             // copies enqueued on cstream that the solve stream has not been made to wait for yet
    cudaStream_t stream;
    long long launches;      // kernels launched on `stream` (guarded by mtx)
    long long ulaunches;     // kernels launched on `ustream` (guarded by umtx)
    // iterate and problem data [B]
    double *x, *u, *pi, *lam, *x0, *p, *yref, *yref_e, *W, *We;
    double *qz, *qpi, *qlam, *qt;
    double *sim_qv; int32_t* sim_div;    // batched rollout: simulation state [B][36], divergence flags [B]
    double *tracking;            // [B][12] base_ref_vel_tracking of the batched MPC, resident (nmpc_plan_tick)
    double *dtn, *lamt, *qx;     // enable_time_opt: node durations [B][N], multipliers of their bounds [B][N][2], QP iterate [B][N][QX_SZ]
    double *res, *qp_res, *cost;
    unsigned long long *res_acc;
    int32_t *status, *done, *sqp_iter, *qp_iter, *qp_status, *qp_ncond, *itmp;
    // QP workspace [Bc]
    double *vec, *cost_stage, *rec;
    double *SR, *FR;             // structured stage / factor records of the fused QP path
    double *G, *M, *C, *D;       // dense stage blocks: diagnostics only (nmpc_export_qp / nmpc_solve_qp), allocated on first use
    int32_t* sr_flag;
    int last_qp_path;            // nmpc_solve_qp: 1 fused (the only solver), -1 refused (QP without the OCP's structure)
    double* scratch;             // staging of the solve path (init tick, x0, dx0, plan -> torque)
    size_t scratch_bytes;
    // Utility calls that do not touch the solver state (nmpc_kinematics, nmpc_inverse_dynamics) have their own
    // stream, staging buffer and lock: LocomotionMPC's control thread evaluates id_torques at 1 kHz while the
    // worker thread's solve is queued on `stream` (mpc.py:164,516,544-545); neither waits for the other.
    cudaStream_t ustream;
    double* uscratch;
    size_t uscratch_bytes;
    std::mutex* mtx;             // solve path: every entry point that uses `stream` / `scratch` / the iterate
    std::mutex* umtx;            // utility path
    // optional per-kernel timing with CUDA events on the handle's stream (bench.py roofline)
    int profiling;
    std::vector<cudaEvent_t>* ev;    // start/stop pairs
    std::vector<int>* ev_kind;       // NMPC_PROF_* kinds
    size_t ev_used;
};

static void prof_begin(nmpc_handle* h, int kind)
{
    if (!h->profiling) return;
    if (h->ev_used + 2 > h->ev->size()) {
        cudaEvent_t a, b;
        cudaEventCreate(&a); cudaEventCreate(&b);
        h->ev->push_back(a); h->ev->push_back(b);
    }
    if (h->ev_kind->size() < h->ev_used / 2 + 1) h->ev_kind->push_back(kind);
    (*h->ev_kind)[h->ev_used / 2] = kind;
    cudaEventRecord((*h->ev)[h->ev_used], h->stream);
}

static void prof_end(nmpc_handle* h)
{
    if (!h->profiling) return;
    cudaEventRecord((*h->ev)[h->ev_used + 1], h->stream);
    h->ev_used += 2;
}

// grows a staging buffer; the caller holds the lock that guards it, and the stream that uses it is drained first so
// that no queued kernel or copy still refers to the old allocation
static int ensure_buf(cudaStream_t st, double** buf, size_t* have, size_t bytes)
{
    if (bytes <= *have) return 0;
    if (*buf) { CK(cudaStreamSynchronize(st)); cudaFree(*buf); }
    *buf = nullptr; *have = 0;
    CK(cudaMalloc(buf, bytes));
    *have = bytes;
    return 0;
}
static int ensure_scratch(nmpc_handle* h, size_t bytes) { return ensure_buf(h->stream, &h->scratch, &h->scratch_bytes, bytes); }
static int ensure_uscratch(nmpc_handle* h, size_t bytes) { return ensure_buf(h->ustream, &h->uscratch, &h->uscratch_bytes, bytes); }
#define LOCK(h) std::lock_guard<std::mutex> lock_(*(h)->mtx)
#define ULOCK(h) std::lock_guard<std::mutex> ulock_(*(h)->umtx)

template <typename T>
static cudaError_t dalloc(T** p, size_t n)
{
    cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
    if (e == cudaSuccess) e = cudaMemset(*p, 0, n * sizeof(T));
    return e;
}

extern "C" const char* nmpc_version(void) { return "nmpc_b200 0.2 (sm_100a, fp64, nx=42 nu=30 ng=42)"; }

extern "C" nmpc_handle* nmpc_create(const nmpc_opts* o, int32_t batch, int32_t device, int32_t* err)
{
    int32_t dummy;
    if (!err) err = &dummy;
    *err = NMPC_EINVAL;
    if (!o || batch <= 0 || o->n_nodes < 1 || o->n_nodes > NMPC_MAX_NODES) return nullptr;
    if (o->enable_time_opt && !(o->dt_min > 0.0 && o->dt_min <= o->dt_nodes && o->dt_nodes <= o->dt_max)) return nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device >= ndev) { *err = NMPC_ENODEV; return nullptr; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major < 10) { *err = NMPC_ENODEV; return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { *err = NMPC_ENODEV; return nullptr; }

    nmpc_handle* h = (nmpc_handle*)calloc(1, sizeof(nmpc_handle));
    h->o = *o; h->B = batch; h->N = o->n_nodes; h->device = device;
    h->n_sm = prop.multiProcessorCount; h->wave_plan = 1;
    if (const char* s = getenv("NMPC_WAVE_PLAN")) h->wave_plan = atoi(s) != 0;
    h->ev = new std::vector<cudaEvent_t>(); h->ev_kind = new std::vector<int>();
    h->mtx = new std::mutex(); h->umtx = new std::mutex();
    int cap = 4096;
    if (const char* s = getenv("NMPC_CHUNK")) { int v = atoi(s); if (v > 0) cap = v; }
    h->Bc = batch < cap ? batch : cap;
    const size_t B = batch, N = h->N, Bc = h->Bc;
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->ustream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->cstream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_wave, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_copied, cudaEventDisableTiming);
#define A_(ptr, n) if (e == cudaSuccess) e = dalloc(&h->ptr, (n))
    A_(x, B * (N + 1) * NX); A_(u, B * N * NU); A_(pi, B * N * NX); A_(lam, B * (N + 1) * 2 * NG);
    A_(x0, B * NX); A_(p, B * (N + 1) * NPAR); A_(yref, B * N * NY); A_(yref_e, B * NYE);
    A_(W, B * NY); A_(We, B * NYE);
    A_(qz, B * (N + 1) * NZ); A_(qpi, B * N * NX); A_(qlam, B * (N + 1) * 2 * NG); A_(qt, B * (N + 1) * 2 * NG);
    A_(dtn, B * N); A_(lamt, B * N * 2); A_(qx, B * N * QX_SZ);
    A_(res, B * 4); A_(qp_res, B * 4); A_(cost, B); A_(res_acc, B * 4);
    A_(status, B); A_(done, B); A_(sqp_iter, B); A_(qp_iter, B); A_(qp_status, B); A_(qp_ncond, B); A_(itmp, B);
    A_(vec, Bc * (N + 1) * V_STAGE); A_(cost_stage, Bc * (N + 1));
    A_(rec, ((Bc * (N + 1) + REC_BLK - 1) / REC_BLK) * NMPC_REC_SIZE * REC_BLK);
    A_(SR, Bc * (N + 1) * SR_SZ); A_(FR, Bc * (N + 1) * FR_SZ); A_(sr_flag, 1);
#undef A_
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(nmpc_qp_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, FQ_SMEM_CAP1);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(nmpc_qp_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FQ_SMEM_CAP1);
    if (e == cudaSuccess) {          // the node durations start at their nominal value
        std::vector<double> d(B * N, o->dt_nodes);
        e = cudaMemcpy(h->dtn, d.data(), d.size() * 8, cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        *err = (e == cudaErrorMemoryAllocation) ? NMPC_ENOMEM : (int)e;
        nmpc_destroy(h);
        return nullptr;
    }
    *err = 0;
    return h;
}

extern "C" void nmpc_destroy(nmpc_handle* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->ustream) cudaStreamSynchronize(h->ustream);
    void* ptrs[] = {h->x, h->u, h->pi, h->lam, h->x0, h->p, h->yref, h->yref_e, h->W, h->We, h->qz, h->qpi,
                    h->qlam, h->qt, h->dtn, h->lamt, h->qx, h->tracking, h->sim_qv, h->sim_div, h->res, h->qp_res, h->cost, h->res_acc, h->status, h->done, h->sqp_iter,
                    h->qp_iter, h->qp_status, h->qp_ncond, h->itmp, h->G, h->M, h->C, h->D, h->vec, h->cost_stage,
                    h->rec, h->scratch, h->uscratch, h->SR, h->FR, h->sr_flag};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->ustream) cudaStreamDestroy(h->ustream);
    if (h->cstream) cudaStreamDestroy(h->cstream);
    if (h->ev_wave) cudaEventDestroy(h->ev_wave);
    if (h->ev_copied) cudaEventDestroy(h->ev_copied);
    delete h->mtx; delete h->umtx;
    if (h->ev) { for (cudaEvent_t e : *h->ev) cudaEventDestroy(e); delete h->ev; }
    delete h->ev_kind;
    free(h);
}

extern "C" int nmpc_set_opts(nmpc_handle* h, const nmpc_opts* o)
{
    if (!h || !o || o->n_nodes != h->N) return NMPC_EINVAL;
    LOCK(h);
    if (o->enable_time_opt != h->o.enable_time_opt) return NMPC_EINVAL;      // fixed at nmpc_create (it changes nu)
    h->o = *o;
    return 0;
}

extern "C" int nmpc_get_dims(const nmpc_handle* h, int32_t* batch, int32_t* n_nodes)
{
    if (!h) return NMPC_EINVAL;
    if (batch) *batch = h->B;
    if (n_nodes) *n_nodes = h->N;
    return 0;
}

extern "C" int nmpc_stream(const nmpc_handle* h, void** s)
{
    if (!h || !s) return NMPC_EINVAL;
    *s = (void*)h->stream;
    return 0;
}

extern "C" int64_t nmpc_launch_count(const nmpc_handle* h) { return h ? h->launches + h->ulaunches : 0; }

static int d2h(nmpc_handle* h, void* dst, const void* src, size_t bytes);
static int h2d(nmpc_handle* h, void* dst, const void* src, size_t bytes)
{
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, h->stream));
    return 0;
}

extern "C" int nmpc_set_x0(nmpc_handle* h, const double* qv)
{
    if (!h || !qv) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (int r = ensure_scratch(h, (size_t)h->B * 36 * 8)) return r;
    CK(cudaMemcpyAsync(h->scratch, qv, (size_t)h->B * 36 * 8, cudaMemcpyDefault, h->stream));
    nmpc_set_x0_kernel<<<(h->B + 63) / 64, 64, 0, h->stream>>>(h->scratch, h->B, h->N, h->x0, h->x);
    h->launches++;
    return (int)cudaGetLastError();
}

extern "C" int nmpc_set_x0_full(nmpc_handle* h, const double* x0)
{
    if (!h || !x0) return NMPC_EINVAL;
    LOCK(h);
    if (int r = h2d(h, h->x0, x0, (size_t)h->B * NX * 8)) return r;
    nmpc_copy_x0_kernel<<<(h->B * NX + 127) / 128, 128, 0, h->stream>>>(h->B, h->N, h->x0, h->x);
    h->launches++;
    return (int)cudaGetLastError();
}

extern "C" int nmpc_set_yref(nmpc_handle* h, const double* yref, const double* yref_e)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    if (yref) if (int r = h2d(h, h->yref, yref, (size_t)h->B * h->N * NY * 8)) return r;
    if (yref_e) if (int r = h2d(h, h->yref_e, yref_e, (size_t)h->B * NYE * 8)) return r;
    return 0;
}

extern "C" int nmpc_set_weights(nmpc_handle* h, const double* W, const double* We, int32_t per_problem)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    if (per_problem) {
        if (W) if (int r = h2d(h, h->W, W, (size_t)h->B * NY * 8)) return r;
        if (We) if (int r = h2d(h, h->We, We, (size_t)h->B * NYE * 8)) return r;
        return 0;
    }
    // broadcast: replicate on the host (cost weights change rarely)
    CK(cudaSetDevice(h->device));
    if (W) {
        std::vector<double> w(NY), rep((size_t)h->B * NY);
        CK(cudaMemcpy(w.data(), W, NY * 8, cudaMemcpyDefault));
        for (int b = 0; b < h->B; b++) memcpy(&rep[(size_t)b * NY], w.data(), NY * 8);
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaMemcpy(h->W, rep.data(), rep.size() * 8, cudaMemcpyHostToDevice));
    }
    if (We) {
        std::vector<double> w(NYE), rep((size_t)h->B * NYE);
        CK(cudaMemcpy(w.data(), We, NYE * 8, cudaMemcpyDefault));
        for (int b = 0; b < h->B; b++) memcpy(&rep[(size_t)b * NYE], w.data(), NYE * 8);
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaMemcpy(h->We, rep.data(), rep.size() * 8, cudaMemcpyHostToDevice));
    }
    return 0;
}

extern "C" int nmpc_set_params(nmpc_handle* h, const double* p)
{
    if (!h || !p) return NMPC_EINVAL;
    LOCK(h);
    return h2d(h, h->p, p, (size_t)h->B * (h->N + 1) * NPAR * 8);
}

// QuadrupedAcadosSolver.init on the device: compact planner outputs in, node-major records out (nmpc_init.cuh)
extern "C" int nmpc_init_tick(nmpc_handle* h, const double* base_ref, const double* base_ref_e, const double* joint_ref,
                              int32_t joint_per_problem, const double* step_height, int32_t height_per_problem,
                              const int8_t* cnt_sequence, const int8_t* swing_peak, const double* cnt_locations,
                              const double* p_gain, double cnt_radius, double height_offset, int32_t restrict_cnt)
{
    if (!h || !base_ref || !joint_ref || !step_height || !cnt_sequence || !p_gain) return NMPC_EINVAL;
    if (restrict_cnt && !cnt_locations) return NMPC_EINVAL;           // "Contact plan not provided" (solver.py:377)
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B, N1 = h->N + 1;
    const size_t n_ref = B * 12, n_jr = joint_per_problem ? B * 12 : 12, n_sh = height_per_problem ? B : 1;
    const size_t n_cnt = B * NF * N1, n_loc = cnt_locations ? B * NF * N1 * 3 : 0;
    const size_t cnt_pad = (n_cnt + 7) & ~(size_t)7;
    const size_t bytes = (2 * n_ref + n_jr + n_sh + n_loc) * 8 + 2 * cnt_pad;
    if (int r = ensure_scratch(h, bytes)) return r;
    double* d_ref = h->scratch;
    double* d_ref_e = d_ref + n_ref;
    double* d_jr = d_ref_e + n_ref;
    double* d_sh = d_jr + n_jr;
    double* d_loc = d_sh + n_sh;
    signed char* d_cnt = reinterpret_cast<signed char*>(d_loc + n_loc);
    signed char* d_peak = d_cnt + cnt_pad;
    CK(cudaMemcpyAsync(d_ref, base_ref, n_ref * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_ref_e, base_ref_e ? base_ref_e : base_ref, n_ref * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_jr, joint_ref, n_jr * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_sh, step_height, n_sh * 8, cudaMemcpyDefault, h->stream));
    if (n_loc) CK(cudaMemcpyAsync(d_loc, cnt_locations, n_loc * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_cnt, cnt_sequence, n_cnt, cudaMemcpyDefault, h->stream));
    if (swing_peak) CK(cudaMemcpyAsync(d_peak, swing_peak, n_cnt, cudaMemcpyDefault, h->stream));
    TickArgs a;
    a.base_ref = d_ref; a.base_ref_e = d_ref_e; a.joint_ref = d_jr; a.step_height = d_sh;
    a.cnt = d_cnt; a.peak = swing_peak ? d_peak : nullptr; a.loc = n_loc ? d_loc : nullptr;
    for (int f = 0; f < NF; f++) a.p_gain[f] = p_gain[f];
    a.radius = cnt_radius; a.height_offset = height_offset; a.restrict_cnt = restrict_cnt;
    a.joint_per_problem = joint_per_problem; a.height_per_problem = height_per_problem;
    a.p = h->p; a.yref = h->yref; a.yref_e = h->yref_e; a.B = h->B; a.N = h->N;
    nmpc_init_tick_kernel<<<h->B, 128, 0, h->stream>>>(a);
    h->launches++;
    return (int)cudaGetLastError();
}

// LocomotionMPC.optimize's planning step + QuadrupedAcadosSolver.init on the device for the cyclic contact planner
// (nmpc_init.cuh): contact / swing windows of the gait cycle at i_node, base references from q, the command and the
// tracked pose, then the expansion of nmpc_init_tick.  Every array may be a host or a device pointer.
extern "C" int nmpc_plan_tick(nmpc_handle* h, const int32_t* i_node, const int8_t* gait_seq, const int8_t* peak_seq,
                              int32_t nodes_per_cycle, const double* q, const double* v_des, const double* w_yaw,
                              const double* tracking, int32_t advance_steps, double sim_dt, double nom_height,
                              double height_offset, double t_horizon, const double* joint_ref, double step_height,
                              const double* p_gain, int32_t opt_peak)
{
    if (!h || !i_node || !gait_seq || !peak_seq || nodes_per_cycle < 1 || !q || !v_des || !w_yaw ||
        !joint_ref || !p_gain || advance_steps < 0) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B, N1 = h->N + 1, npc = nodes_per_cycle;
    const size_t n_cnt = B * NF * N1, cnt_pad = (n_cnt + 7) & ~(size_t)7, seq_pad = (NF * npc + 7) & ~(size_t)7;
    // doubles: q 18 | v_des 3 | w 1 | tracking 12 | ref 12 | ref_e 12 per problem, joint_ref 12, step height 1
    const size_t n_d = B * 58 + 13;
    const size_t bytes = n_d * 8 + ((B * 4 + 7) & ~(size_t)7) + 2 * seq_pad + 2 * cnt_pad;
    if (int r = ensure_scratch(h, bytes)) return r;
    double* d_q = h->scratch;
    double* d_vd = d_q + B * 18;
    double* d_w = d_vd + B * 3;
    double* d_tr = d_w + B;
    double* d_ref = d_tr + B * 12;
    double* d_ref_e = d_ref + B * 12;
    double* d_jr = d_ref_e + B * 12;
    double* d_sh = d_jr + 12;
    int32_t* d_node = reinterpret_cast<int32_t*>(d_sh + 1);
    signed char* d_gs = reinterpret_cast<signed char*>(d_node) + ((B * 4 + 7) & ~(size_t)7);
    signed char* d_ps = d_gs + seq_pad;
    signed char* d_cnt = d_ps + seq_pad;
    signed char* d_peak = d_cnt + cnt_pad;
    CK(cudaMemcpyAsync(d_q, q, B * 18 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_vd, v_des, B * 3 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_w, w_yaw, B * 8, cudaMemcpyDefault, h->stream));
    if (!h->tracking) { cudaError_t e = dalloc(&h->tracking, B * 12); if (e != cudaSuccess) return (int)e; }
    if (tracking) CK(cudaMemcpyAsync(h->tracking, tracking, B * 12 * 8, cudaMemcpyDefault, h->stream));
    (void)d_tr;
    CK(cudaMemcpyAsync(d_jr, joint_ref, 12 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_sh, &step_height, 8, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_node, i_node, B * 4, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_gs, gait_seq, NF * npc, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_ps, peak_seq, NF * npc, cudaMemcpyDefault, h->stream));
    PlannerArgs pa;
    pa.i_node = d_node; pa.gait_seq = d_gs; pa.peak_seq = d_ps; pa.q = d_q; pa.v_des = d_vd; pa.w_yaw = d_w;
    pa.tracking = h->tracking; pa.advance_steps = advance_steps; pa.sim_dt = sim_dt; pa.nom_height = nom_height; pa.height_offset = height_offset; pa.t_horizon = t_horizon;
    pa.npc = nodes_per_cycle; pa.N = h->N; pa.B = h->B;
    pa.base_ref = d_ref; pa.base_ref_e = d_ref_e; pa.cnt = d_cnt; pa.peak = d_peak;
    nmpc_plan_refs_kernel<<<h->B, 128, 0, h->stream>>>(pa);
    TickArgs a;
    a.base_ref = d_ref; a.base_ref_e = d_ref_e; a.joint_ref = d_jr; a.step_height = d_sh;
    a.cnt = d_cnt; a.peak = opt_peak ? d_peak : nullptr; a.loc = nullptr;
    for (int f = 0; f < NF; f++) a.p_gain[f] = p_gain[f];
    a.radius = 0.0; a.height_offset = height_offset; a.restrict_cnt = 0;
    a.joint_per_problem = 0; a.height_per_problem = 0;
    a.p = h->p; a.yref = h->yref; a.yref_e = h->yref_e; a.B = h->B; a.N = h->N;
    nmpc_init_tick_kernel<<<h->B, 128, 0, h->stream>>>(a);
    h->launches += 2;
    return (int)cudaGetLastError();
}

// LocomotionMPC.optimize with contact_planner="raibert" on the device (nmpc_init.cuh): centre of mass by the model
// kernel, contact / swing windows, Raibert footholds, the contact-restricted base references, then the expansion of
// nmpc_init_tick with the contact restriction on.  Every array may be a host or a device pointer.
extern "C" int nmpc_plan_tick_raibert(nmpc_handle* h, const int32_t* i_node, const int8_t* gait_seq,
                                      const int8_t* peak_seq, const int8_t* make_seq, int32_t nodes_per_cycle,
                                      const double* qv, const double* v_des, const double* w_yaw, const double* t_cnt,
                                      const nmpc_raibert_params* rp, double nom_height, double height_offset,
                                      const double* joint_ref, double step_height, const double* p_gain,
                                      int32_t opt_peak, double* cnt_locations)
{
    if (!h || !i_node || !gait_seq || !peak_seq || !make_seq || nodes_per_cycle < 1 || !qv || !v_des || !w_yaw ||
        !t_cnt || !rp || !joint_ref || !p_gain) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B, N1 = h->N + 1, npc = nodes_per_cycle;
    const size_t n_cnt = B * NF * N1, cnt_pad = (n_cnt + 7) & ~(size_t)7, seq_pad = (NF * npc + 7) & ~(size_t)7;
    // doubles: qv 36 | com 3 | v_des 3 | w 1 | ref 12 | ref_e 12 | loc 12 (N+1) per problem, t_cnt N+1, joint_ref 12, step height 1
    const size_t n_d = B * (67 + 12 * N1) + N1 + 13;
    const size_t bytes = n_d * 8 + ((B * 4 + 7) & ~(size_t)7) + 3 * seq_pad + 2 * cnt_pad;
    if (int r = ensure_scratch(h, bytes)) return r;
    double* d_qv = h->scratch;
    double* d_com = d_qv + B * 36;
    double* d_vd = d_com + B * 3;
    double* d_w = d_vd + B * 3;
    double* d_ref = d_w + B;
    double* d_ref_e = d_ref + B * 12;
    double* d_loc = d_ref_e + B * 12;
    double* d_tc = d_loc + B * 12 * N1;
    double* d_jr = d_tc + N1;
    double* d_sh = d_jr + 12;
    int32_t* d_node = reinterpret_cast<int32_t*>(d_sh + 1);
    signed char* d_gs = reinterpret_cast<signed char*>(d_node) + ((B * 4 + 7) & ~(size_t)7);
    signed char* d_ps = d_gs + seq_pad;
    signed char* d_ms = d_ps + seq_pad;
    signed char* d_cnt = d_ms + seq_pad;
    signed char* d_peak = d_cnt + cnt_pad;
    CK(cudaMemcpyAsync(d_qv, qv, B * 36 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_vd, v_des, B * 3 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_w, w_yaw, B * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_tc, t_cnt, N1 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_jr, joint_ref, 12 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_sh, &step_height, 8, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_node, i_node, B * 4, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_gs, gait_seq, NF * npc, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_ps, peak_seq, NF * npc, cudaMemcpyDefault, h->stream));
    CK(cudaMemcpyAsync(d_ms, make_seq, NF * npc, cudaMemcpyDefault, h->stream));
    nmpc_kin0_kernel<<<(h->B + 63) / 64, 64, 0, h->stream>>>(d_qv, h->B, nullptr, d_com, nullptr);
    RaibertArgs ra;
    ra.i_node = d_node; ra.gait_seq = d_gs; ra.peak_seq = d_ps; ra.make_seq = d_ms; ra.qv = d_qv; ra.com = d_com;
    ra.v_des = d_vd; ra.w_yaw = d_w; ra.t_cnt = d_tc;
    for (int i = 0; i < NF * 3; i++) ra.rp.offset_hip_b[i] = rp->offset_hip_b[i];
    for (int f = 0; f < NF; f++) ra.rp.stance_ratio[f] = rp->stance_ratio[f];
    ra.rp.nominal_period = rp->nominal_period; ra.rp.foot_size = rp->foot_size;
    ra.rp.height_offset = rp->height_offset; ra.rp.gravity = rp->gravity;
    ra.nom_height = nom_height; ra.height_offset = height_offset;
    ra.npc = nodes_per_cycle; ra.N = h->N; ra.B = h->B;
    ra.base_ref = d_ref; ra.base_ref_e = d_ref_e; ra.cnt = d_cnt; ra.peak = d_peak; ra.loc = d_loc;
    nmpc_plan_raibert_kernel<<<h->B, 128, 0, h->stream>>>(ra);
    TickArgs a;
    a.base_ref = d_ref; a.base_ref_e = d_ref_e; a.joint_ref = d_jr; a.step_height = d_sh;
    a.cnt = d_cnt; a.peak = opt_peak ? d_peak : nullptr; a.loc = d_loc;
    for (int f = 0; f < NF; f++) a.p_gain[f] = p_gain[f];
    a.radius = rp->cnt_radius; a.height_offset = height_offset; a.restrict_cnt = 1;
    a.joint_per_problem = 0; a.height_per_problem = 0;
    a.p = h->p; a.yref = h->yref; a.yref_e = h->yref_e; a.B = h->B; a.N = h->N;
    nmpc_init_tick_kernel<<<h->B, 128, 0, h->stream>>>(a);
    h->launches += 3;
    if (cnt_locations) {
        CK(cudaMemcpyAsync(cnt_locations, d_loc, B * 12 * N1 * 8, cudaMemcpyDefault, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return (int)cudaGetLastError();
}

// the tracked reference pose [B,12] as nmpc_plan_tick holds (and advances) it on the device
extern "C" int nmpc_get_tracking(nmpc_handle* h, double* tracking)
{
    if (!h || !tracking || !h->tracking) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(tracking, h->tracking, (size_t)h->B * 12 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

// cold first guess from x0 and the contact mask already on the device (call after nmpc_set_x0 and nmpc_init_tick / nmpc_set_params)
extern "C" int nmpc_cold_guess(nmpc_handle* h)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    nmpc_cold_guess_kernel<<<h->B, 128, 0, h->stream>>>(h->N, h->x0, h->p, h->x, h->u);
    h->launches++;
    if (h->o.enable_time_opt) {
        nmpc_fill_kernel<<<(h->B * h->N + 255) / 256, 256, 0, h->stream>>>(h->dtn, h->B * h->N, h->o.dt_nodes);
        h->launches++;
    }
    return (int)cudaGetLastError();
}

// problem data as the solver holds it (tests: device-side expansion vs the host packing)
extern "C" int nmpc_get_problem_data(nmpc_handle* h, double* p, double* yref, double* yref_e, double* x0)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B, N = h->N;
    if (p) if (int r = d2h(h, p, h->p, B * (N + 1) * NPAR * 8)) return r;
    if (yref) if (int r = d2h(h, yref, h->yref, B * N * NY * 8)) return r;
    if (yref_e) if (int r = d2h(h, yref_e, h->yref_e, B * NYE * 8)) return r;
    if (x0) if (int r = d2h(h, x0, h->x0, B * NX * 8)) return r;
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_pin_stance_feet(nmpc_handle* h, const int32_t* i_node)
{
    if (!h || !i_node) return NMPC_EINVAL;
    LOCK(h);
    if (int r = h2d(h, h->itmp, i_node, (size_t)h->B * 4)) return r;
    nmpc_pin_feet_kernel<<<(h->B + 63) / 64, 64, 0, h->stream>>>(h->B, h->N, h->x0, h->itmp, h->p);
    h->launches++;
    return (int)cudaGetLastError();
}

extern "C" int nmpc_set_guess(nmpc_handle* h, const double* x, const double* u)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    if (x) if (int r = h2d(h, h->x, x, (size_t)h->B * (h->N + 1) * NX * 8)) return r;
    if (u) if (int r = h2d(h, h->u, u, (size_t)h->B * h->N * NU * 8)) return r;
    if (u && h->o.enable_time_opt) {                // a new input guess restarts the node durations at dt_nodes (solver.py:189-190)
        nmpc_fill_kernel<<<(h->B * h->N + 255) / 256, 256, 0, h->stream>>>(h->dtn, h->B * h->N, h->o.dt_nodes);
        h->launches++;
    }
    return 0;
}

// enable_time_opt: node durations of the iterate, dt [B][N] (inputs["dt"], solver.py:189-190, 325-336).  A null pointer
// resets them to the nominal dt_nodes.
extern "C" int nmpc_set_dt(nmpc_handle* h, const double* dt)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (dt) return h2d(h, h->dtn, dt, (size_t)h->B * h->N * 8);
    nmpc_fill_kernel<<<(h->B * h->N + 255) / 256, 256, 0, h->stream>>>(h->dtn, h->B * h->N, h->o.dt_nodes);
    h->launches++;
    return (int)cudaGetLastError();
}

// dt_node_sol (solver.py:423-427): dt [B][N] (the nominal value everywhere unless enable_time_opt) and, optionally,
// the multipliers of dt_min <= dt <= dt_max, lam_dt [B][N][2] (lower | upper)
extern "C" int nmpc_get_dt(nmpc_handle* h, double* dt, double* lam_dt)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (dt) if (int r = d2h(h, dt, h->dtn, (size_t)h->B * h->N * 8)) return r;
    if (lam_dt) if (int r = d2h(h, lam_dt, h->lamt, (size_t)h->B * h->N * 2 * 8)) return r;
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_shift_warm_start(nmpc_handle* h, const int32_t* start_node, int32_t repeat_last)
{
    if (!h || !start_node) return NMPC_EINVAL;
    LOCK(h);
    if (int r = h2d(h, h->itmp, start_node, (size_t)h->B * 4)) return r;
    const bool topt = h->o.enable_time_opt != 0;
    nmpc_shift_kernel<<<h->B, 128, 0, h->stream>>>(h->N, h->itmp, repeat_last, h->x, h->u, h->pi, h->lam,
                                                   topt ? h->dtn : nullptr, h->lamt, h->o.dt_nodes);
    h->launches++;
    return (int)cudaGetLastError();
}

extern "C" int nmpc_reset_duals(nmpc_handle* h)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B, N = h->N;
    CK(cudaMemsetAsync(h->pi, 0, B * N * NX * 8, h->stream));
    CK(cudaMemsetAsync(h->lam, 0, B * (N + 1) * 2 * NG * 8, h->stream));
    CK(cudaMemsetAsync(h->qz, 0, B * (N + 1) * NZ * 8, h->stream));
    CK(cudaMemsetAsync(h->qpi, 0, B * N * NX * 8, h->stream));
    CK(cudaMemsetAsync(h->qlam, 0, B * (N + 1) * 2 * NG * 8, h->stream));
    CK(cudaMemsetAsync(h->qt, 0, B * (N + 1) * 2 * NG * 8, h->stream));
    CK(cudaMemsetAsync(h->lamt, 0, B * N * 2 * 8, h->stream));
    CK(cudaMemsetAsync(h->qx, 0, B * N * QX_SZ * 8, h->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// solve
// ---------------------------------------------------------------------------------------------
// dense stage blocks (3.6 MB per OCP of the chunk): only the diagnostics below touch them, so they are
// allocated on the first nmpc_export_qp / nmpc_solve_qp call and not by nmpc_create
static int ensure_dense(nmpc_handle* h)
{
    if (h->G) return 0;
    const size_t Bc = h->Bc, N = h->N;
    cudaError_t e = dalloc(&h->G, Bc * N * G_SZ);
    if (e == cudaSuccess) e = dalloc(&h->M, Bc * (N + 1) * M_SZ);
    if (e == cudaSuccess) e = dalloc(&h->C, Bc * (N + 1) * C_SZ);
    if (e == cudaSuccess) e = dalloc(&h->D, Bc * (N + 1) * D_SZ);
    if (e != cudaSuccess) return e == cudaErrorMemoryAllocation ? NMPC_ENOMEM : (int)e;
    return 0;
}

static LinArgs lin_args(nmpc_handle* h, int b0)
{
    LinArgs a;
    a.x = h->x; a.u = h->u; a.p = h->p; a.yref = h->yref; a.yref_e = h->yref_e; a.W = h->W; a.We = h->We;
    a.G = h->G; a.M = h->M; a.C = h->C; a.D = h->D; a.cost_stage = h->cost_stage; a.done = h->done;
    a.rec = h->rec; a.N = h->N; a.b0 = b0; a.dt = h->o.dt_nodes; a.reg = h->o.reg_eps; a.reg_e = h->o.reg_eps_e;
    a.mu_c = h->o.mu_contact;
    a.dtn = h->dtn; a.topt = h->o.enable_time_opt != 0; a.w_dt = h->o.time_opt_weight;
    a.dt_min = h->o.dt_min; a.dt_max = h->o.dt_max;
    return a;
}

// model evaluation (thread per instance x part) followed by the per-stage assembly
static void launch_linearize(nmpc_handle* h, int b0, int Bc)
{
    const int N = h->N, n_inst = Bc * (N + 1);
    ModelArgs ma;
    ma.x = h->x; ma.rec = h->rec; ma.done = h->done; ma.n_inst = n_inst; ma.N = N; ma.b0 = b0;
    nmpc_model_kernel<<<dim3((n_inst + MODEL_THREADS - 1) / MODEL_THREADS, NMPC_MODEL_PARTS), MODEL_THREADS, 0,
                        h->stream>>>(ma);
    LinArgs la = lin_args(h, b0);
    nmpc_linearize_kernel<<<dim3(N + 1, Bc), LIN_THREADS, 0, h->stream>>>(la);
    h->launches += 2;
}

// product path: model evaluation followed by the assembly of the structured stage records
static void launch_linearize_sr(nmpc_handle* h, int b0, int Bc)
{
    const int N = h->N, n_inst = Bc * (N + 1);
    ModelArgs ma;
    ma.x = h->x; ma.rec = h->rec; ma.done = h->done; ma.n_inst = n_inst; ma.N = N; ma.b0 = b0;
    nmpc_model_kernel<<<dim3((n_inst + MODEL_THREADS - 1) / MODEL_THREADS, NMPC_MODEL_PARTS), MODEL_THREADS, 0,
                        h->stream>>>(ma);
    LinArgs la = lin_args(h, b0);
    nmpc_linearize_sr_kernel<<<dim3(N + 1, Bc), LIN_SR_THREADS, 0, h->stream>>>(la, h->SR);
    h->launches += 2;
}

static QpArgs qp_args(nmpc_handle* h, int b0)
{
    QpArgs a;
    a.qz = h->qz; a.qpi = h->qpi; a.qlam = h->qlam; a.qt = h->qt; a.qx = h->qx;
    a.qp_iter = h->qp_iter; a.qp_status = h->qp_status; a.qp_ncond = h->qp_ncond; a.qp_res = h->qp_res;
    a.dx0_ext = nullptr; a.x0 = h->x0; a.x = h->x; a.u = h->u; a.pi = h->pi; a.lam = h->lam;
    a.sqp_done = h->done; a.sqp_iter = h->sqp_iter;
    a.N = h->N; a.b0 = b0; a.max_iter = h->o.max_qp_iter; a.warm = h->o.warm_start_qp;
    a.cond_pred_corr = h->o.cond_pred_corr;
    a.tol = h->o.qp_tol;
    return a;
}

// The product path: ONE fused kernel per QP (nmpc_qp_fused.cuh) on the stage records the linearisation wrote.
static void run_qp_fused(nmpc_handle* h, QpArgs a, int Bc, bool apply, bool final_iterate = false)
{
    const int N = h->N;
    cudaStream_t s = h->stream;
    FqArgs f;
    f.SR = h->SR; f.FR = h->FR; f.vec = h->vec; f.qz = a.qz; f.qpi = a.qpi; f.qlam = a.qlam; f.qt = a.qt; f.qx = a.qx;
    f.qp_iter = a.qp_iter; f.qp_status = a.qp_status; f.qp_ncond = a.qp_ncond; f.qp_res = a.qp_res;
    f.dx0_ext = a.dx0_ext; f.x0 = a.x0;
    f.x = a.x; f.sqp_done = a.sqp_done; f.N = N; f.b0 = a.b0; f.max_iter = a.max_iter; f.warm = a.warm;
    f.cond_pred_corr = a.cond_pred_corr;
    f.apply = apply ? 1 : 0; f.sx = a.x; f.su = a.u; f.spi = a.pi; f.slam = a.lam; f.sqp_iter = a.sqp_iter;
    f.dtn = h->dtn; f.lamt = h->lamt;
    f.tol = a.tol; f.dt = h->o.dt_nodes; f.c2 = 0.5 * h->o.dt_nodes * h->o.dt_nodes;
    const bool topt = h->o.enable_time_opt != 0 && a.dx0_ext == nullptr;      // externally supplied QPs have nu = 30
    // Launch plan.  A CTA's run time hardly depends on how many CTAs share its SM (the kernel is bound by the latency of
    // one OCP's chain: 3.8 / 4.3 / 4.9 / 5.9 ms with 1 / 2 / 3 / 4 CTAs per SM on config 2), and the hardware refills an SM
    // as soon as its CTAs retire: in one launch of 1024 CTAs the last 432 land four to an SM on 108 SMs while 40 SMs
    // idle, and in a long launch the co-resident CTAs drift into different passes and miss the instruction cache.  So
    // the chunk goes out wave by wave: full waves of 4 CTAs per SM, then the rest with the residency capped (through
    // the dynamic shared memory size) at the smallest number of CTAs per SM that holds it, spread over all SMs.
    const int full = 4 * h->n_sm;
    int off = 0;
    while (off < Bc) {
        int n = Bc - off, smem = FQ_SMEM_BYTES;
        if (h->wave_plan) {
            if (n > full) n = full;
            else {
                const int per_sm = (n + h->n_sm - 1) / h->n_sm;
                smem = per_sm >= 4 ? FQ_SMEM_BYTES : per_sm == 3 ? FQ_SMEM_CAP3 : per_sm == 2 ? FQ_SMEM_CAP2 : FQ_SMEM_CAP1;
            }
        }
        f.bl0 = off;
        prof_begin(h, NMPC_PROF_QP_FUSED);
        if (topt) nmpc_qp_fused_kernel<true><<<n, FQ_THREADS, smem, s>>>(f);
        else nmpc_qp_fused_kernel<false><<<n, FQ_THREADS, smem, s>>>(f);
        prof_end(h);
        h->launches++;
        if (apply) {             // the kernel has applied the SQP step of this wave: its rows are final
            if (final_iterate && (h->out_x || h->out_u)) {
                // nmpc_set_output: this wave's x / u leave on the copy stream while the next wave is being solved
                const size_t g0 = (size_t)a.b0 + off, nx = (size_t)(N + 1) * NX * 8, nu = (size_t)N * NU * 8;
                cudaEventRecord(h->ev_wave, s);
                cudaStreamWaitEvent(h->cstream, h->ev_wave, 0);
                if (h->out_x) cudaMemcpyAsync((char*)h->out_x + g0 * nx, (const char*)h->x + g0 * nx, n * nx, cudaMemcpyDefault, h->cstream);
                if (h->out_u) cudaMemcpyAsync((char*)h->out_u + g0 * nu, (const char*)h->u + g0 * nu, n * nu, cudaMemcpyDefault, h->cstream);
                h->out_pending = 1;
            }
        }
        off += n;
    }
}

extern "C" int nmpc_solve(nmpc_handle* h, int32_t* status)
{
    if (!h) return NMPC_EINVAL;
    if (h->o.hpipm_mode != NMPC_HPIPM_SPEED) return NMPC_ENOSYS;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const int B = h->B, N = h->N;
    const int rti = h->o.real_time_it != 0;
    const int max_iter = rti ? 1 : h->o.max_iter;
    nmpc_sqp_begin_kernel<<<(B + 127) / 128, 128, 0, h->stream>>>(B, rti, h->status, h->done, h->sqp_iter,
                                                                 h->qp_iter, h->qp_status, h->qp_ncond, h->res_acc);
    h->launches++;
    for (int b0 = 0; b0 < B; b0 += h->Bc) {
        const int Bc = (B - b0 < h->Bc) ? (B - b0) : h->Bc;
        for (int it = 0; it <= max_iter; it++) {
            if (rti && it == 1) break;
            prof_begin(h, NMPC_PROF_LINEARIZE);
            launch_linearize_sr(h, b0, Bc);
            prof_end(h);
            if (!rti) {
                prof_begin(h, NMPC_PROF_RESIDUAL);
                ResSrArgs ra;
                ra.SR = h->SR; ra.x = h->x; ra.x0 = h->x0; ra.pi = h->pi; ra.lam = h->lam; ra.res_acc = h->res_acc;
                ra.done = h->done; ra.N = N; ra.b0 = b0; ra.dt = h->o.dt_nodes;
                ra.lamt = h->lamt; ra.topt = h->o.enable_time_opt != 0;
                nmpc_residual_sr_kernel<<<dim3(N + 1, Bc), RES_THREADS, 0, h->stream>>>(ra);
                prof_end(h);
                nmpc_sqp_decide_kernel<<<(Bc + 127) / 128, 128, 0, h->stream>>>(
                    b0, Bc, N, it, max_iter, h->o.nlp_tol, h->res_acc, h->cost_stage, h->res, h->cost,
                    h->status, h->done);
                h->launches += 2;
                if (it == max_iter) break;
            } else {
                nmpc_cost_sum_kernel<<<(Bc + 127) / 128, 128, 0, h->stream>>>(b0, Bc, N, h->cost_stage, h->cost);
                h->launches++;
            }
            QpArgs qa = qp_args(h, b0);
            // acados ocp_nlp_sqp / sqp_rti: "(typically) no warm start at first iteration" -- the QP solver's
            // warm_start is forced to 0 for the first QP of a solve unless opts->warm_start_first_qp
            if (it == 0 && !h->o.warm_start_first_qp) qa.warm = 0;
            run_qp_fused(h, qa, Bc, true, /*final_iterate=*/rti || it == max_iter - 1);
            nmpc_sqp_post_qp_kernel<<<(Bc + 127) / 128, 128, 0, h->stream>>>(b0, Bc, h->qp_status, h->status, h->done);
            h->launches++;
        }
    }
    if ((h->out_x || h->out_u) && !h->out_pending) {     // no QP ran (max_iter = 0): the iterate as it is
        if (h->out_x) CK(cudaMemcpyAsync(h->out_x, h->x, (size_t)B * (N + 1) * NX * 8, cudaMemcpyDefault, h->stream));
        if (h->out_u) CK(cudaMemcpyAsync(h->out_u, h->u, (size_t)B * N * NU * 8, cudaMemcpyDefault, h->stream));
    }
    if (h->out_pending) {        // everything that later waits for the solve stream also waits for the streamed outputs
        CK(cudaEventRecord(h->ev_copied, h->cstream));
        CK(cudaStreamWaitEvent(h->stream, h->ev_copied, 0));
        h->out_pending = 0;
    }
    CK(cudaGetLastError());
    if (status) {
        CK(cudaMemcpyAsync(status, h->status, (size_t)B * 4, cudaMemcpyDefault, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

extern "C" int nmpc_set_profiling(nmpc_handle* h, int32_t on)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    h->profiling = on ? 1 : 0;
    h->ev_used = 0;
    return 0;
}

// accumulated device time [ms] and launch counts per kernel kind since the last call; resets
extern "C" int nmpc_get_kernel_times(nmpc_handle* h, double* ms, int64_t* count)
{
    if (!h || !ms || !count) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < NMPC_PROF_KINDS; i++) { ms[i] = 0.0; count[i] = 0; }
    for (size_t i = 0; i + 1 < h->ev_used; i += 2) {
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, (*h->ev)[i], (*h->ev)[i + 1]));
        const int kind = (*h->ev_kind)[i / 2];
        ms[kind] += t; count[kind] += 1;
    }
    h->ev_used = 0;
    return 0;
}

extern "C" int nmpc_synchronize(nmpc_handle* h)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return (int)cudaGetLastError();
}

static int d2h(nmpc_handle* h, void* dst, const void* src, size_t bytes)
{
    CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, h->stream));
    return 0;
}

extern "C" int nmpc_get_solution(nmpc_handle* h, double* x, double* u, double* pi, double* lam)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B, N = h->N;
    if (x) if (int r = d2h(h, x, h->x, B * (N + 1) * NX * 8)) return r;
    if (u) if (int r = d2h(h, u, h->u, B * N * NU * 8)) return r;
    if (pi) if (int r = d2h(h, pi, h->pi, B * N * NX * 8)) return r;
    if (lam) if (int r = d2h(h, lam, h->lam, B * (N + 1) * 2 * NG * 8)) return r;
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

// Output streaming: destinations that every following nmpc_solve fills with the final x / u, wave by wave, on a copy
// stream, while the next wave is still being solved (the D2H copy of parse_sol hidden behind the solve).
extern "C" int nmpc_set_output(nmpc_handle* h, double* x, double* u)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));            // no solve in flight still refers to the old destinations
    CK(cudaStreamSynchronize(h->cstream));
    h->out_x = x; h->out_u = u;
    return 0;
}

// per-problem status of the last solve (host or device pointer); synchronises the solve stream
extern "C" int nmpc_get_status(nmpc_handle* h, int32_t* status)
{
    if (!h || !status) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(status, h->status, (size_t)h->B * 4, cudaMemcpyDefault, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

// packed result rows [B, (N+1)*42 + N*31 + 1] into a DEVICE buffer, enqueued on the solve stream without a
// synchronisation: the send buffer of the gather to the dataset-writer rank
extern "C" int nmpc_pack_solution(nmpc_handle* h, double* out_dev)
{
    if (!h || !out_dev) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, out_dev) != cudaSuccess || at.type != cudaMemoryTypeDevice) { cudaGetLastError(); return NMPC_EINVAL; }
    nmpc_pack_solution_kernel<<<h->B, 128, 0, h->stream>>>(h->N, h->o.dt_nodes, h->o.enable_time_opt ? h->dtn : nullptr,
                                                           h->x, h->u, h->status, out_dev);
    h->launches++;
    return (int)cudaGetLastError();
}

extern "C" int nmpc_get_stats(nmpc_handle* h, nmpc_stats* s)
{
    if (!h || !s) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t B = h->B;
    if (s->sqp_iter) if (int r = d2h(h, s->sqp_iter, h->sqp_iter, B * 4)) return r;
    if (s->qp_iter) if (int r = d2h(h, s->qp_iter, h->qp_iter, B * 4)) return r;
    if (s->qp_status) if (int r = d2h(h, s->qp_status, h->qp_status, B * 4)) return r;
    if (s->res) if (int r = d2h(h, s->res, h->res, B * 4 * 8)) return r;
    if (s->qp_res) if (int r = d2h(h, s->qp_res, h->qp_res, B * 4 * 8)) return r;
    if (s->cost) if (int r = d2h(h, s->cost, h->cost, B * 8)) return r;
    if (s->qp_ncond) if (int r = d2h(h, s->qp_ncond, h->qp_ncond, B * 4)) return r;
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_get_qp_step(nmpc_handle* h, double* dz)
{
    if (!h || !dz) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (int r = d2h(h, dz, h->qz, (size_t)h->B * (h->N + 1) * NZ * 8)) return r;
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_kinematics(nmpc_handle* h, const double* qv, int32_t n, double* feet, double* com, double* hg)
{
    if (!h || !qv || n <= 0) return NMPC_EINVAL;
    ULOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t in_b = (size_t)n * 36 * 8, out_d = (size_t)n * (12 + 3 + 6);
    if (int r = ensure_uscratch(h, in_b + out_d * 8)) return r;
    double* d_in = h->uscratch;
    double* d_feet = h->uscratch + (size_t)n * 36;
    double* d_com = d_feet + (size_t)n * 12;
    double* d_hg = d_com + (size_t)n * 3;
    CK(cudaMemcpyAsync(d_in, qv, in_b, cudaMemcpyDefault, h->ustream));
    nmpc_kin0_kernel<<<(n + 63) / 64, 64, 0, h->ustream>>>(d_in, n, d_feet, d_com, d_hg);
    h->ulaunches++;
    if (feet) CK(cudaMemcpyAsync(feet, d_feet, (size_t)n * 12 * 8, cudaMemcpyDefault, h->ustream));
    if (com) CK(cudaMemcpyAsync(com, d_com, (size_t)n * 3 * 8, cudaMemcpyDefault, h->ustream));
    if (hg) CK(cudaMemcpyAsync(hg, d_hg, (size_t)n * 6 * 8, cudaMemcpyDefault, h->ustream));
    CK(cudaStreamSynchronize(h->ustream));
    return (int)cudaGetLastError();
}

// ---- allocation-free variant of nmpc_kinematics (no reference counterpart)
// The staging buffer nmpc_kinematics takes from cudaMalloc lives here in a file-scope device array, so this entry point
// can be called where cudaMalloc cannot: under stream capture, in serving loops, in harnesses that forbid allocation.
// Layout, in doubles: qv [n][36] | feet [n][12] | com [n][3] | hg [n][6], n <= NMPC_STATIC_MAX_N (57 doubles each).
// Every GPU that loads this library holds the array (1.78 MiB) until the process exits.  In host code the NAME of the
// array is a host-side stand-in, not a device address: the address a kernel can use comes from cudaGetSymbolAddress,
// once per device, with that device current.
static __device__ __align__(256) double g_static_kin_scratch[(size_t)NMPC_STATIC_MAX_N * 57];
#define NMPC_MAX_DEVICES 64
static double* g_static_kin_ptr[NMPC_MAX_DEVICES];      // device address of the array on each device, null until set up
static std::mutex g_static_kin_mtx;                     // guards the table and serialises the calls of this process

extern "C" int nmpc_kinematics_static(const double* qv, int32_t n, double* feet, double* com, double* hg, void* stream)
{
    if (!qv || n <= 0) return NMPC_EINVAL;
    if (n > NMPC_STATIC_MAX_N) {
        fprintf(stderr, "nmpc_kinematics_static: n = %d exceeds the static scratch (NMPC_STATIC_MAX_N = %d instances); "
                        "split the batch or use nmpc_kinematics\n", (int)n, (int)NMPC_STATIC_MAX_N);
        return NMPC_E2BIG;
    }
    std::lock_guard<std::mutex> lock(g_static_kin_mtx);
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NMPC_MAX_DEVICES) return NMPC_ENODEV;
    if (!g_static_kin_ptr[dev]) {                       // set-up, once per device, that device current
        double* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_kin_scratch);
        if (e != cudaSuccess || !p) {
            fprintf(stderr, "nmpc_kinematics_static: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : NMPC_ENODEV;
        }
        g_static_kin_ptr[dev] = p;
    }
    cudaStream_t st = (cudaStream_t)stream;
    double* d_in = g_static_kin_ptr[dev];
    double* d_feet = d_in + (size_t)n * 36;
    double* d_com = d_feet + (size_t)n * 12;
    double* d_hg = d_com + (size_t)n * 3;
    // nothing has to start at zero: the kernel writes every output of instances 0..n-1 and nothing beyond is read back
    CK(cudaMemcpyAsync(d_in, qv, (size_t)n * 36 * 8, cudaMemcpyDefault, st));
    nmpc_kin0_kernel<<<(n + 63) / 64, 64, 0, st>>>(d_in, n, d_feet, d_com, d_hg);
    if (feet) CK(cudaMemcpyAsync(feet, d_feet, (size_t)n * 12 * 8, cudaMemcpyDefault, st));
    if (com) CK(cudaMemcpyAsync(com, d_com, (size_t)n * 3 * 8, cudaMemcpyDefault, st));
    if (hg) CK(cudaMemcpyAsync(hg, d_hg, (size_t)n * 6 * 8, cudaMemcpyDefault, st));
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap == cudaStreamCaptureStatusNone)
        CK(cudaStreamSynchronize(st));                  // outputs complete on return (not under capture: the graph owns the order)
    return (int)cudaGetLastError();
}

extern "C" double nmpc_model_mass(void) { return NMPC_MASS; }

extern "C" int nmpc_inverse_dynamics(nmpc_handle* h, const double* qva, const double* f, int32_t n, double* tau)
{
    if (!h || !qva || !f || !tau || n <= 0) return NMPC_EINVAL;
    ULOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t nd = (size_t)n * (54 + 12 + 12);
    if (int r = ensure_uscratch(h, nd * 8)) return r;
    double* d_qva = h->uscratch;
    double* d_f = d_qva + (size_t)n * 54;
    double* d_tau = d_f + (size_t)n * 12;
    CK(cudaMemcpyAsync(d_qva, qva, (size_t)n * 54 * 8, cudaMemcpyDefault, h->ustream));
    CK(cudaMemcpyAsync(d_f, f, (size_t)n * 12 * 8, cudaMemcpyDefault, h->ustream));
    nmpc_rnea_kernel<<<(n + 63) / 64, 64, 0, h->ustream>>>(d_qva, d_f, n, d_tau);
    h->ulaunches++;
    CK(cudaMemcpyAsync(tau, d_tau, (size_t)n * 12 * 8, cudaMemcpyDefault, h->ustream));
    CK(cudaStreamSynchronize(h->ustream));
    return (int)cudaGetLastError();
}

// plan -> torque for every problem of the batch, from the solution resident on the device (nmpc_plan.cuh)
extern "C" int nmpc_plan_torques(nmpc_handle* h, const double* qv_meas, int32_t n_interp, int32_t step0, int32_t n_steps,
                                 double kp, double kd, double* q_plan, double* v_plan, double* tau_ff, double* tau,
                                 double* action)
{
    if (!h || n_interp < 2 || n_steps <= 0 || step0 < 0 || step0 + n_steps > n_interp || !(kp != 0.0)) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t n = (size_t)h->B * n_steps;
    // scratch: [meas 36 | q 18 | v 18 | tau_ff 12 | tau 12 | action 12] per (problem, step)
    if (int r = ensure_scratch(h, n * 108 * 8)) return r;
    double* d_meas = h->scratch;
    double* d_q = d_meas + n * 36;
    double* d_v = d_q + n * 18;
    double* d_tff = d_v + n * 18;
    double* d_tau = d_tff + n * 12;
    double* d_act = d_tau + n * 12;
    if (qv_meas) CK(cudaMemcpyAsync(d_meas, qv_meas, n * 36 * 8, cudaMemcpyDefault, h->stream));
    PlanArgs a;
    a.x = h->x; a.u = h->u; a.qv_meas = qv_meas ? d_meas : nullptr;
    a.q_plan = q_plan ? d_q : nullptr; a.v_plan = v_plan ? d_v : nullptr; a.tau_ff = tau_ff ? d_tff : nullptr;
    a.tau = tau ? d_tau : nullptr; a.action = action ? d_act : nullptr;
    a.B = h->B; a.N = h->N; a.n_interp = n_interp; a.step0 = step0; a.n_steps = n_steps;
    a.dt = h->o.dt_nodes; a.kp = kp; a.kd = kd;
    a.dtn = h->o.enable_time_opt ? h->dtn : nullptr;
    nmpc_plan_torque_kernel<<<(unsigned)((n + 63) / 64), 64, 0, h->stream>>>(a);
    h->launches++;
    if (q_plan) CK(cudaMemcpyAsync(q_plan, d_q, n * 18 * 8, cudaMemcpyDefault, h->stream));
    if (v_plan) CK(cudaMemcpyAsync(v_plan, d_v, n * 18 * 8, cudaMemcpyDefault, h->stream));
    if (tau_ff) CK(cudaMemcpyAsync(tau_ff, d_tff, n * 12 * 8, cudaMemcpyDefault, h->stream));
    if (tau) CK(cudaMemcpyAsync(tau, d_tau, n * 12 * 8, cudaMemcpyDefault, h->stream));
    if (action) CK(cudaMemcpyAsync(action, d_act, n * 12 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// checkpoint / resume of the solver state: everything a warm-started solve depends on
// ---------------------------------------------------------------------------------------------
struct StatePart { void* ptr; size_t bytes; };
static int state_parts(nmpc_handle* h, StatePart* out)
{
    const size_t B = h->B, N = h->N;
    int n = 0;
    out[n++] = {h->x, B * (N + 1) * NX * 8}; out[n++] = {h->u, B * N * NU * 8}; out[n++] = {h->pi, B * N * NX * 8};
    out[n++] = {h->lam, B * (N + 1) * 2 * NG * 8};
    out[n++] = {h->qz, B * (N + 1) * NZ * 8}; out[n++] = {h->qpi, B * N * NX * 8};
    out[n++] = {h->qlam, B * (N + 1) * 2 * NG * 8}; out[n++] = {h->qt, B * (N + 1) * 2 * NG * 8};
    out[n++] = {h->dtn, B * N * 8}; out[n++] = {h->lamt, B * N * 2 * 8}; out[n++] = {h->qx, B * N * QX_SZ * 8};
    out[n++] = {h->x0, B * NX * 8}; out[n++] = {h->p, B * (N + 1) * NPAR * 8};
    out[n++] = {h->yref, B * N * NY * 8}; out[n++] = {h->yref_e, B * NYE * 8};
    out[n++] = {h->W, B * NY * 8}; out[n++] = {h->We, B * NYE * 8};
    return n;
}

extern "C" int64_t nmpc_state_bytes(nmpc_handle* h)
{
    if (!h) return 0;
    StatePart parts[32];
    const int n = state_parts(h, parts);
    size_t tot = 16;                               // header: B, N, nu build, reserved
    for (int i = 0; i < n; i++) tot += parts[i].bytes;
    return (int64_t)tot;
}

// iterate (primal, dual, QP warm start, node durations) and problem data into / from a host blob of nmpc_state_bytes
extern "C" int nmpc_export_state(nmpc_handle* h, void* blob)
{
    if (!h || !blob) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    StatePart parts[32];
    const int n = state_parts(h, parts);
    int32_t* hdr = (int32_t*)blob;
    hdr[0] = h->B; hdr[1] = h->N; hdr[2] = NU + (h->o.enable_time_opt ? 1 : 0); hdr[3] = 0;
    char* p = (char*)blob + 16;
    for (int i = 0; i < n; i++) { CK(cudaMemcpyAsync(p, parts[i].ptr, parts[i].bytes, cudaMemcpyDeviceToHost, h->stream)); p += parts[i].bytes; }
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_import_state(nmpc_handle* h, const void* blob)
{
    if (!h || !blob) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const int32_t* hdr = (const int32_t*)blob;
    if (hdr[0] != h->B || hdr[1] != h->N || hdr[2] != NU + (h->o.enable_time_opt ? 1 : 0)) return NMPC_EINVAL;
    StatePart parts[32];
    const int n = state_parts(h, parts);
    const char* p = (const char*)blob + 16;
    for (int i = 0; i < n; i++) { CK(cudaMemcpyAsync(parts[i].ptr, p, parts[i].bytes, cudaMemcpyHostToDevice, h->stream)); p += parts[i].bytes; }
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// batched rigid-body rollout (nmpc_rollout.cuh)
// ---------------------------------------------------------------------------------------------
static int ensure_sim(nmpc_handle* h)
{
    if (h->sim_qv) return 0;
    cudaError_t e = dalloc(&h->sim_qv, (size_t)h->B * 36);
    if (e == cudaSuccess) e = dalloc(&h->sim_div, (size_t)h->B);
    return e == cudaSuccess ? 0 : (e == cudaErrorMemoryAllocation ? NMPC_ENOMEM : (int)e);
}

extern "C" int nmpc_sim_set_state(nmpc_handle* h, const double* qv)
{
    if (!h || !qv) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (int r = ensure_sim(h)) return r;
    CK(cudaMemcpyAsync(h->sim_qv, qv, (size_t)h->B * 36 * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemsetAsync(h->sim_div, 0, (size_t)h->B * 4, h->stream));
    return 0;
}

extern "C" int nmpc_sim_get_state(nmpc_handle* h, double* qv, int32_t* diverged)
{
    if (!h || !h->sim_qv) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (qv) CK(cudaMemcpyAsync(qv, h->sim_qv, (size_t)h->B * 36 * 8, cudaMemcpyDefault, h->stream));
    if (diverged) CK(cudaMemcpyAsync(diverged, h->sim_div, (size_t)h->B * 4, cudaMemcpyDefault, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_sim_steps(nmpc_handle* h, const nmpc_sim_params* p, int32_t n_interp, int32_t step0, int32_t n_steps,
                              double* out_qv, double* out_tau, double* out_fc)
{
    if (!h || !p || !h->sim_qv || n_interp < 2 || n_steps <= 0 || step0 < 0 || !(p->sim_dt > 0.0)) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    const size_t n = (size_t)h->B * n_steps;
    double *d_qv = nullptr, *d_tau = nullptr, *d_fc = nullptr;
    if (out_qv || out_tau || out_fc) {
        if (int r = ensure_scratch(h, n * 60 * 8)) return r;
        d_qv = h->scratch; d_tau = d_qv + n * 36; d_fc = d_tau + n * 12;
    }
    SimArgs a;
    a.x = h->x; a.u = h->u; a.dtn = h->o.enable_time_opt ? h->dtn : nullptr;
    a.qv = h->sim_qv; a.out_qv = out_qv ? d_qv : nullptr; a.out_tau = out_tau ? d_tau : nullptr; a.out_fc = out_fc ? d_fc : nullptr;
    a.diverged = h->sim_div;
    a.B = h->B; a.N = h->N; a.n_interp = n_interp; a.step0 = step0; a.n_steps = n_steps; a.dt_nodes = h->o.dt_nodes;
    a.p.sim_dt = p->sim_dt; a.p.kp = p->kp; a.p.kd = p->kd; a.p.k_n = p->k_n; a.p.d_n = p->d_n; a.p.d_t = p->d_t;
    a.p.mu = p->mu; a.p.ground_z = p->ground_z; a.p.tau_max = p->tau_max;
    nmpc_sim_kernel<<<(h->B + 63) / 64, 64, 0, h->stream>>>(a);
    h->launches++;
    if (out_qv) CK(cudaMemcpyAsync(out_qv, d_qv, n * 36 * 8, cudaMemcpyDefault, h->stream));
    if (out_tau) CK(cudaMemcpyAsync(out_tau, d_tau, n * 12 * 8, cudaMemcpyDefault, h->stream));
    if (out_fc) CK(cudaMemcpyAsync(out_fc, d_fc, n * 12 * 8, cudaMemcpyDefault, h->stream));
    if (out_qv || out_tau || out_fc) CK(cudaStreamSynchronize(h->stream));
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// diagnostics: export / import dense QPs in the HPIPM-style layouts of the header
// ---------------------------------------------------------------------------------------------
extern "C" int nmpc_export_qp(nmpc_handle* h, double* BAbt, double* RSQrq, double* DCt, double* lg, double* ug,
                              double* mask)
{
    if (!h) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    if (int r = ensure_dense(h)) return r;
    const int B = h->B, N = h->N;
    std::vector<double> hG((size_t)h->Bc * N * G_SZ), hM((size_t)h->Bc * (N + 1) * M_SZ),
        hC((size_t)h->Bc * (N + 1) * C_SZ), hD((size_t)h->Bc * (N + 1) * D_SZ);
    CK(cudaMemsetAsync(h->done, 0, (size_t)B * 4, h->stream));
    for (int b0 = 0; b0 < B; b0 += h->Bc) {
        const int Bc = (B - b0 < h->Bc) ? (B - b0) : h->Bc;
        launch_linearize(h, b0, Bc);
        CK(cudaMemcpyAsync(hG.data(), h->G, (size_t)Bc * N * G_SZ * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(hM.data(), h->M, (size_t)Bc * (N + 1) * M_SZ * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(hC.data(), h->C, (size_t)Bc * (N + 1) * C_SZ * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(hD.data(), h->D, (size_t)Bc * (N + 1) * D_SZ * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        for (int bl = 0; bl < Bc; bl++) {
            const size_t b = b0 + bl;
            for (int k = 0; k <= N; k++) {
                if (BAbt && k < N) {
                    const double* g = &hG[((size_t)bl * N + k) * G_SZ];
                    double* o = BAbt + (b * N + k) * (NZ + 1) * NX;
                    for (int c = 0; c <= NZ; c++)
                        for (int l = 0; l < NX; l++) o[c * NX + l] = g[l * G_LD + c];
                }
                if (RSQrq) {
                    const double* m = &hM[((size_t)bl * (N + 1) + k) * M_SZ];
                    double* o = RSQrq + (b * (N + 1) + k) * (NZ + 1) * NZ;
                    for (int r = 0; r <= NZ; r++)
                        for (int c = 0; c < NZ; c++)
                            o[r * NZ + c] = (r < NZ && c > r) ? m[c * M_LD + r] : m[r * M_LD + c];
                }
                if (DCt) {
                    const double* cc = &hC[((size_t)bl * (N + 1) + k) * C_SZ];
                    double* o = DCt + (b * (N + 1) + k) * NZ * NG;
                    for (int c = 0; c < NZ; c++)
                        for (int j = 0; j < NG; j++) o[c * NG + j] = cc[j * C_LD + c];
                }
                const double* d = &hD[((size_t)bl * (N + 1) + k) * D_SZ];
                if (lg) memcpy(lg + (b * (N + 1) + k) * NG, d, NG * 8);
                if (ug) memcpy(ug + (b * (N + 1) + k) * NG, d + NG, NG * 8);
                if (mask) memcpy(mask + (b * (N + 1) + k) * 2 * NG, d + 2 * NG, 2 * NG * 8);
            }
        }
    }
    return 0;
}

extern "C" int nmpc_solve_qp(nmpc_handle* h, const double* BAbt, const double* RSQrq, const double* DCt,
                             const double* lg, const double* ug, const double* mask, const double* dx0,
                             double* dz, double* pi, double* lam, double* t, int32_t* qp_iter, int32_t* qp_status)
{
    if (!h || !BAbt || !RSQrq || !DCt || !lg || !ug || !mask || !dx0) return NMPC_EINVAL;
    LOCK(h);
    CK(cudaSetDevice(h->device));
    {   // the stage blocks are repacked on the host below: device pointers would be dereferenced there
        const void* in[] = {BAbt, RSQrq, DCt, lg, ug, mask};
        for (const void* q : in) {
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, q) == cudaSuccess && at.type == cudaMemoryTypeDevice) return NMPC_EINVAL;
            cudaGetLastError();
        }
    }
    if (int r = ensure_dense(h)) return r;
    const int B = h->B, N = h->N;
    if (int r = ensure_scratch(h, (size_t)B * NX * 8)) return r;
    CK(cudaMemcpyAsync(h->scratch, dx0, (size_t)B * NX * 8, cudaMemcpyDefault, h->stream));
    CK(cudaMemsetAsync(h->done, 0, (size_t)B * 4, h->stream));
    std::vector<double> hG((size_t)h->Bc * N * G_SZ, 0.0), hM((size_t)h->Bc * (N + 1) * M_SZ, 0.0),
        hC((size_t)h->Bc * (N + 1) * C_SZ, 0.0), hD((size_t)h->Bc * (N + 1) * D_SZ, 0.0);
    for (int b0 = 0; b0 < B; b0 += h->Bc) {
        const int Bc = (B - b0 < h->Bc) ? (B - b0) : h->Bc;
        for (int bl = 0; bl < Bc; bl++) {
            const size_t b = b0 + bl;
            for (int k = 0; k <= N; k++) {
                if (k < N) {
                    double* g = &hG[((size_t)bl * N + k) * G_SZ];
                    const double* o = BAbt + (b * N + k) * (NZ + 1) * NX;
                    for (int c = 0; c <= NZ; c++)
                        for (int l = 0; l < NX; l++) g[l * G_LD + c] = o[c * NX + l];
                }
                double* m = &hM[((size_t)bl * (N + 1) + k) * M_SZ];
                const double* o = RSQrq + (b * (N + 1) + k) * (NZ + 1) * NZ;
                for (int r = 0; r <= NZ; r++)              // HPIPM convention: lower triangle is the data
                    for (int c = 0; c < NZ; c++)
                        m[r * M_LD + c] = (r < NZ && c > r) ? o[c * NZ + r] : o[r * NZ + c];
                double* cc = &hC[((size_t)bl * (N + 1) + k) * C_SZ];
                const double* oc = DCt + (b * (N + 1) + k) * NZ * NG;
                for (int c = 0; c < NZ; c++)
                    for (int j = 0; j < NG; j++) cc[j * C_LD + c] = oc[c * NG + j];
                double* d = &hD[((size_t)bl * (N + 1) + k) * D_SZ];
                memcpy(d, lg + (b * (N + 1) + k) * NG, NG * 8);
                memcpy(d + NG, ug + (b * (N + 1) + k) * NG, NG * 8);
                memcpy(d + 2 * NG, mask + (b * (N + 1) + k) * 2 * NG, 2 * NG * 8);
            }
        }
        CK(cudaMemcpyAsync(h->G, hG.data(), (size_t)Bc * N * G_SZ * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->M, hM.data(), (size_t)Bc * (N + 1) * M_SZ * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->C, hC.data(), (size_t)Bc * (N + 1) * C_SZ * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->D, hD.data(), (size_t)Bc * (N + 1) * D_SZ * 8, cudaMemcpyHostToDevice, h->stream));
        QpArgs qa = qp_args(h, b0);
        qa.dx0_ext = h->scratch; qa.warm = 0; qa.sqp_done = nullptr; qa.sqp_iter = nullptr;
        // Only QPs with the structure of this OCP family (every QP the solver itself builds) can be solved:
        // the dense blocks are converted into stage records and checked entry by entry against what the
        // records expand to; anything else is refused (there is no generic dense solver in this build).
        int32_t unstructured = 0;
        CK(cudaMemsetAsync(h->sr_flag, 0, 4, h->stream));
        SrArgs sa;
        sa.G = h->G; sa.M = h->M; sa.C = h->C; sa.D = h->D; sa.SR = h->SR; sa.flag = h->sr_flag;
        sa.done = nullptr; sa.N = N; sa.b0 = b0; sa.check = 1; sa.dt = h->o.dt_nodes;
        nmpc_dense2sr_kernel<<<dim3(N + 1, Bc), 128, 0, h->stream>>>(sa);
        h->launches++;
        CK(cudaMemcpyAsync(&unstructured, h->sr_flag, 4, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        h->last_qp_path = unstructured ? -1 : 1;
        if (unstructured) return NMPC_ENOSYS;
        run_qp_fused(h, qa, Bc, false);
        CK(cudaStreamSynchronize(h->stream));
    }
    CK(cudaGetLastError());
    const size_t Bs = B;
    if (dz) CK(cudaMemcpyAsync(dz, h->qz, Bs * (N + 1) * NZ * 8, cudaMemcpyDefault, h->stream));
    if (pi) CK(cudaMemcpyAsync(pi, h->qpi, Bs * N * NX * 8, cudaMemcpyDefault, h->stream));
    if (lam) CK(cudaMemcpyAsync(lam, h->qlam, Bs * (N + 1) * 2 * NG * 8, cudaMemcpyDefault, h->stream));
    if (t) CK(cudaMemcpyAsync(t, h->qt, Bs * (N + 1) * 2 * NG * 8, cudaMemcpyDefault, h->stream));
    if (qp_iter) CK(cudaMemcpyAsync(qp_iter, h->qp_iter, Bs * 4, cudaMemcpyDefault, h->stream));
    if (qp_status) CK(cudaMemcpyAsync(qp_status, h->qp_status, Bs * 4, cudaMemcpyDefault, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int nmpc_last_qp_path(const nmpc_handle* h) { return h ? h->last_qp_path : -1; }

// debug builds (-DNMPC_TIMING): per-phase cycle counts of CTA 0 of nmpc_qp_fused_kernel (24 counters)
extern "C" int nmpc_debug_phase_cycles(unsigned long long* out24, int reset)
{
#ifdef NMPC_TIMING
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpyFromSymbol(out24, g_fq_cycles, sizeof(unsigned long long) * 32));
    if (reset) { unsigned long long z[32] = {0}; CK(cudaMemcpyToSymbol(g_fq_cycles, z, sizeof(z))); }
    return 0;
#else
    (void)out24; (void)reset;
    return NMPC_ENOSYS;
#endif
}
