// Single-filter engine + its C ABI (include/aruco_slam_b200.h, "single filter" section).
// Host side: id -> index association (dict semantics of EKF:75-78 / EKF:253-254), staging
// of one frame's observations, and the per-frame launch sequence documented in
// single_kernels.cuh.  All arithmetic runs in the CUDA kernels; there is no CPU fallback.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <new>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "front.cuh"
#include "pipe.cuh"
#include "single_kernels.cuh"

using namespace asb;

#ifndef ASB_SMALL_MAP_DEFAULT
#define ASB_SMALL_MAP_DEFAULT 192
#endif

namespace {
constexpr uint32_t FILTER_MAGIC = 0x41534246;  // "ASBF"
constexpr int N_SLOTS = 8;                     // staging ring depth

struct FrameRec {  // one frame of a device-resident stream
    int64_t off;
    int k, n0, n1;
    std::vector<int32_t> new_ids;
};
}  // namespace

struct asb_filter {
    uint32_t magic = FILTER_MAGIC;
    int kind = 0, quat_mode = 0, device = 0;
    int D = 3, R = 3, W = 13;
    Noise nz{};
    int cap_lm = 0, ncap = 0, ld = 0, max_obs = 0, kc_max = 0;
    int nlm = 0;
    std::unordered_map<int32_t, int> map;
    std::vector<int32_t> id_order;

    cudaStream_t own_stream = nullptr, stream = nullptr;
    double *d_state = nullptr, *d_P = nullptr, *d_Vr = nullptr, *d_Vc = nullptr, *d_blocks = nullptr, *d_nu = nullptr;
    double *d_dx[2] = {nullptr, nullptr}, *dx_last = nullptr, *d_unc = nullptr;
    int* d_status = nullptr;
    unsigned long long* d_prof = nullptr;
    double* d_gate = nullptr; int* d_gate_i = nullptr; int gate_cap = 0;   // asb_correct_poses staging
    // opt-in moving-average motion model (README.md:55-56; absent from the reference code)
    int motion_window = 0; long long motion_frames = 0; double* d_hist = nullptr;
    // staging ring (host pinned + device)
    // per-frame staging, N_SLOTS slots of stage_stride bytes: [poses k x 6 f64 | idx k x i32 | isnew k x u8], each part
    // 16-byte aligned, so that one frame's observations go to the device in ONE copy
    unsigned char* h_stage = nullptr; unsigned char* d_stage = nullptr; size_t stage_stride = 0;
    // one-launch frames: the kernel reads the staged observations straight from the page-locked slot and writes the
    // posterior state + status into this page-locked mirror (ncap doubles, then the status int)
    double* h_mirror = nullptr;
    cudaEvent_t slot_ev[N_SLOTS]{};
    float* h_corner = nullptr; float* d_corner = nullptr;   // asb_observe_corners: (slot, max_obs, 4, 2) float32
    cudaEvent_t ev_pose = nullptr;                            // the IPPE poses of that call are back on the host
    int slot = 0;
    int last_k = 0;
    const int* last_idx = nullptr;
    // resident stream
    int* s_idx = nullptr; double* s_pose = nullptr; unsigned char* s_new = nullptr;
    std::vector<FrameRec> frames;
    int next_frame = 0;
    // The resident stream bakes landmark indices against the map as it was at upload: every map mutation outside
    // asb_stream_run bumps map_gen, and a run on a stale upload is refused (ASB_ERR_NO_STREAM).
    unsigned long long map_gen = 0, stream_gen = 0;
    int sweep_grid = 0, smem_optin = 0;
    // pipelined engine (pipe.cuh): covariance ping-pong pair, side stream for k_f1
    double *d_P2 = nullptr, *d_At = nullptr, *d_y = nullptr, *d_Vt = nullptr;
    int* d_T = nullptr;
    unsigned int* d_counter = nullptr;
    cudaStream_t side = nullptr;
    cudaEvent_t ev_ready = nullptr, ev_f1 = nullptr;
    bool state_on_main = false;    // ... and nothing is queued behind it on the main stream (one-launch frame)
    bool state_at_ready = false;   // the state vector is final at ev_ready (only a sweep is queued behind it):
                                   // asb_get_state need not wait for that sweep
    bool lazy = false;         // d_P2 holds the input of the sweep launched last, d_Vr its panel
    bool side_waited = false;
    int prof_slot = 0, spare_ctas = 2;
    unsigned long long* d_ctaprof = nullptr;   // 3 x 1024: per-CTA start / end / smid of the last sweep         // timeline of the last two updates (asb_timeline)
    int prev_m4 = 0, prev_add_q = 0;
    // stream mode (asb_stream_run): the sweep of update t is launched AFTER k_f1 of update t+1
    // (programmatic dependent launches, see launch_frame); until then it is kept here
    struct PendingSweep {
        bool valid = false;
        int n1 = 0, m4 = 0;
        SweepExtra ex{};
        const double* pin = nullptr;
        double* pout = nullptr;
    } pending;
    unsigned int* d_tail = nullptr;   // [0,1] tile counters, [2,3] range counters, used alternately by successive sweeps (sweep.cuh)
    int reserve_sm = -1, reserve_ctas = 2;   // per-frame mode: the SM kept free for k_f1 (the highest %smid seen at creation)
    unsigned int sweep_seq = 0;
    int tail_tiles = 2;        // tiles per CTA handed out dynamically at the end of a sweep
    int use_vt = 0;            // 1: k_f2 also writes the panel by state column for k_f1 (bulk copies instead of a gather); measured slower, kept off
    int lower_only = 0;        // pipelined engine: the sweeps keep only the lower triangle of P up to date
    bool defer = false;        // set by asb_stream_run around its frame loop
    int pdl = 1;               // stream mode launches its kernels as programmatic dependents of each other

    // observe() on a map of at most small_n_max state dims runs as ONE launch of the resident batched kernel
    // (common.cuh SmallFrame); larger maps take the pipelined k_f1 / k_f2 / k_sweep engine.  0 = never.
    int small_n_max = ASB_SMALL_MAP_DEFAULT;

    int n_state() const { return CAM + D * nlm; }
};

static int check(asb_filter* f) {
    if (!f || f->magic != FILTER_MAGIC) return fail(ASB_ERR_BAD_HANDLE, "bad asb_filter handle");
    return ASB_OK;
}

#define ASB_CHECK_HANDLE(f)             \
    do {                                \
        int _rc = check(f);             \
        if (_rc != ASB_OK) return _rc;  \
        cudaSetDevice((f)->device);     \
    } while (0)

template <typename T>
static cudaError_t dmalloc(T** p, size_t count) {
    return cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T));
}

// ---------------------------------------------------------------------------------------
// launch sequence for one frame whose observations are at (idx, pose, isnew) on the device
// ---------------------------------------------------------------------------------------
// persistent sweep: 2 CTAs / SM (64 x 64 tiles) while both fit in shared memory, else
// 1 CTA / SM (128 x 64 tiles)
// P[i][j] = P[j][i] for i < j (32 x 32 tiles through shared memory): the covariance getter of an engine
// whose sweeps keep only the lower triangle.
__global__ void __launch_bounds__(256) k_symmetrize(double* P, int ld, int n) {
    __shared__ double tile[32][33];
    const int bi = blockIdx.y, bj = blockIdx.x;   // destination tile (rows bi, columns bj), bj >= bi
    if (bj < bi) return;
    for (int r = threadIdx.y; r < 32; r += 8) {   // source tile (rows bj, columns bi)
        const int i = 32 * bj + r, j = 32 * bi + threadIdx.x;
        tile[r][threadIdx.x] = (i < n && j < n) ? P[(long long)i * ld + j] : 0.0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += 8) {
        const int i = 32 * bi + r, j = 32 * bj + threadIdx.x;
        if (i < n && j < n && i < j) P[(long long)i * ld + j] = tile[threadIdx.x][r];
    }
}

// <<<>>> with (optionally) the programmatic-stream-serialization attribute: the kernel may be
// scheduled as soon as every CTA of the previous kernel of the stream has executed
// griddepcontrol.launch_dependents (or exited); it orders itself with griddepcontrol.wait.
template <typename... KA, typename... A>
static void launch_k(void (*kern)(KA...), int grid, int block, size_t smem, cudaStream_t st, bool pdl, A&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, static_cast<KA>(args)...);
}

// Moving-average motion model (README.md:55-56: X_k|k-1 = X_k-1 + (X_k-1 - X_k-n) / n on the camera position), opt-in.
// hist is a ring of the last `window` posterior camera positions; frame = number of positions pushed so far.
__global__ void k_motion(double* __restrict__ state, double* __restrict__ hist, long long frame, int window) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const double cur[3] = {state[0], state[1], state[2]};
    double* slot = hist + 3 * (frame % window);
    const long long have = frame + 1 < window ? frame + 1 : window;     // positions available including this one
    if (have >= 2) {
        // X_{k-n} with n = have: the oldest of the `have` positions held once this one is pushed
        const double* old = frame + 1 >= window ? hist + 3 * ((frame + 1) % window) : hist;
        const double o[3] = {old[0], old[1], old[2]};
        for (int c = 0; c < 3; ++c) state[c] = cur[c] + (cur[c] - o[c]) / (double)have;
    }
    slot[0] = cur[0]; slot[1] = cur[1]; slot[2] = cur[2];
}

// highest %smid of the device (SM ids need not be 0 .. count-1): the SM the per-frame sweeps keep free
__global__ void k_smid_probe(unsigned int* out) {
    unsigned int sm;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
    if (threadIdx.x == 0) atomicMax(out, sm);
}

static void launch_sweep(asb_filter* f, int n1, int m4, int dbg, const SweepExtra& ex_in, int config = -1,
                         const double* pin = nullptr, double* pout = nullptr, int spare_ctas = 0, bool pdl = false,
                         bool dynamic_tail = false, bool reserve = false) {
    SweepExtra ex = ex_in;
    ex.reserve_sm = -1;
    // hybrid tile schedule (sweep.cuh): all but the last tail_tiles tiles of a CTA's share are a
    // contiguous range, the rest are handed out one at a time
    auto schedule = [&](int ntiles, int grid) {
        if (!dynamic_tail || f->tail_tiles <= 0 || !f->d_tail) return;
        ex.tail = f->d_tail + (f->sweep_seq & 1);
        ex.tail_reset = f->d_tail + ((f->sweep_seq + 1) & 1);
        int ranges = grid;
        // only when the sweep fills the GPU: a small grid leaves SMs free anyway, and its few CTAs must not retire
        if (reserve && f->reserve_sm >= 0 && grid >= 2 * f->sweep_grid) {
            // per-frame mode: the CTAs that land on SM reserve_sm retire (k_f1 of the next frame gets that SM);
            // the other CTAs claim the static ranges from a counter (sweep.cuh)
            ex.reserve_sm = f->reserve_sm;
            ex.range_ctr = f->d_tail + 2 + (f->sweep_seq & 1);
            ex.range_reset = f->d_tail + 2 + ((f->sweep_seq + 1) & 1);
            ranges = std::max(1, grid - f->reserve_ctas);
            ex.n_ranges = ranges;
        }
        ex.n_static = std::max(0, ntiles / ranges - f->tail_tiles);
        ++f->sweep_seq;
    };
    if (!pin) pin = f->d_P;
    if (!pout) pout = f->d_P;
    const bool dual = config < 0 ? (2 * (SweepDual::smem_bytes(m4) + 1024) <= (size_t)f->smem_optin)
                                 : config == 1;
    if (dual) {
        const int ntiles = SweepDual::num_tiles(n1);
        const int grid = std::min(ntiles, 2 * f->sweep_grid - ((reserve && f->reserve_sm >= 0) ? 0 : spare_ctas));
        schedule(ntiles, grid);
        if (ex.reserve_sm >= 0)
            launch_k(k_sweep<SweepDual, true>, grid, SweepDual::THREADS, SweepDual::smem_bytes(m4), f->stream, pdl,
                     pin, pout, f->ld, f->d_Vr, f->d_Vc, m4, ntiles, dbg, ex);
        else
            launch_k(k_sweep<SweepDual>, grid, SweepDual::THREADS, SweepDual::smem_bytes(m4), f->stream, pdl,
                     pin, pout, f->ld, f->d_Vr, f->d_Vc, m4, ntiles, dbg, ex);
    } else {
        const int ntiles = SweepBig::num_tiles(n1);
        const int grid = std::min(ntiles, f->sweep_grid - (spare_ctas ? 1 : 0));
        schedule(ntiles, grid);
        launch_k(k_sweep<SweepBig>, grid, SweepBig::THREADS, SweepBig::smem_bytes(m4), f->stream, pdl,
                 pin, pout, f->ld, f->d_Vr, f->d_Vc, m4, ntiles, dbg, ex);
    }
}

// Launch the sweep kept back by the stream-mode engine (after_f1: as a programmatic dependent of
// the k_f1 just launched, so that k_f1 already owns its SM when the sweep's CTAs are placed).
static void flush_pending(asb_filter* f, bool after_f1) {
    if (!f->pending.valid) return;
    f->pending.valid = false;
    launch_sweep(f, f->pending.n1, f->pending.m4, 2 << 4, f->pending.ex, -1, f->pending.pin, f->pending.pout,
                 f->spare_ctas, after_f1, true);
}

// Side stream waits until the state, dx and panel of the previous update are final (k_f2 done);
// it does NOT wait for the sweep launched after that k_f2.
static void pipe_wait(asb_filter* f) {
    if (f->side_waited) return;
    if (!f->lazy) cudaEventRecord(f->ev_ready, f->stream);   // everything enqueued so far
    cudaStreamWaitEvent(f->side, f->ev_ready, 0);
    f->side_waited = true;
}

template <int KIND>
static int launch_frame(asb_filter* f, const int* idx, const double* pose, const unsigned char* isnew,
                        int k, int n0, int n1, bool do_predict, bool do_update) {
    constexpr int R = Dims<KIND>::R, W = CAM + Dims<KIND>::D;
    cudaStream_t st = f->stream;
    const int ld = f->ld;
    if (do_update && k > 0) {
        // pipelined update (pipe.cuh): per chunk k_f1 on the side stream, k_f2 + k_sweep on the main one
        int cur = 0;
        for (int c0 = 0; c0 < k; c0 += f->kc_max) {
            const int kc = std::min(f->kc_max, k - c0);
            const int m4 = round_up(R * kc, 4);
            const bool first = c0 == 0, last = c0 + kc >= k;
            // Stream mode: everything on the main stream in the order
            //   k_f2(t) -> k_f1(t+1) -> k_sweep(t) -> k_f2(t+1) -> k_f1(t+2) -> k_sweep(t+1) ...
            // each a programmatic dependent of its predecessor.  k_f1 asks for (nearly) all the shared
            // memory of an SM and starts BEFORE the sweep it overlaps, so it has that SM to itself and
            // the sweep's CTAs (two per SM on the others) never share the FP64 pipe with it.
            // Per-frame calls (observe()): the next frame is not known when the sweep is launched, so
            // k_f1 goes to the side stream and squeezes in next to one sweep CTA.
            const bool ordered = f->defer && f->pdl;
            if (ordered) {
                if (!f->lazy) flush_pending(f, false);
            } else {
                flush_pending(f, false);
                pipe_wait(f);
                f->side_waited = false;
            }
            F1Args a;
            a.state = f->d_state; a.P = f->lazy ? f->d_P2 : f->d_P; a.PT_dense = nullptr; a.Vc_tiled = nullptr; a.ld = ld;
            a.n0 = first ? n0 : n1; a.n1 = n1; a.nz = f->nz;
            a.idx = idx + c0; a.poses = pose + 6 * (size_t)c0; a.isnew = isnew ? isnew + c0 : nullptr;
            a.kc = kc; a.m4 = m4;
            a.first_chunk = first ? 1 : 0;
            a.add_q = (first && do_predict) ? 1 : 0;
            a.use_delta = first ? 0 : 1;
            a.dx_in = f->d_dx[cur];
            a.lazy = f->lazy ? 1 : 0; a.m4p = f->prev_m4; a.lazy_q = f->prev_add_q;
            // the previous panel's columns T: gathered from the tiled V_J panel (k_f2 then writes no third copy)
            a.Vp = f->use_vt ? f->d_Vt : nullptr;
            a.Vc_tiled = f->use_vt ? nullptr : f->d_Vc;
            a.At = f->d_At; a.y = f->d_y; a.T = f->d_T;
            a.blocks_out = f->d_blocks + (size_t)c0 * R * W; a.nu_out = f->d_nu + (size_t)R * c0;
            a.status = f->d_status;
            a.dbg = 0;
            unsigned long long* prof = f->d_prof + 16 * f->prof_slot;
            f->prof_slot ^= 1;
            a.prof = prof;
            if (ordered) {
                launch_k(k_f1<KIND, true>, 1, FRONT_THREADS, (size_t)f->smem_optin - 2048, st, true, a);
                flush_pending(f, true);
            } else {
                // the sweep of the previous frame (if one is running) keeps SM reserve_sm free: the exclusive
                // variant asks for a whole SM's shared memory, so that is where it lands
                k_f1<KIND, true><<<1, FRONT_THREADS, (size_t)f->smem_optin - 2048, f->side>>>(a);
                cudaEventRecord(f->ev_f1, f->side);
                cudaStreamWaitEvent(st, f->ev_f1, 0);
            }
            F2Args b;
            b.state = f->d_state; b.P = f->d_P; b.ld = ld; b.n0 = a.n0; b.n1 = n1; b.nz = f->nz;
            b.kc = kc; b.m4 = m4; b.first_chunk = a.first_chunk; b.add_q = a.add_q; b.use_delta = a.use_delta;
            b.At = f->d_At; b.y = f->d_y; b.T = f->d_T;
            b.dx_in = f->d_dx[cur]; b.dx_out = f->d_dx[cur ^ 1];
            b.Vr = f->d_Vr; b.Vc = f->d_Vc; b.Vt = f->use_vt ? f->d_Vt : nullptr;
            b.lower_only = f->lower_only;
            b.apply_kind = last ? KIND : -1;
            b.as_written = (KIND == 0 && f->quat_mode == ASB_QUAT_AS_WRITTEN) ? 1 : 0;
            b.nlm = (n1 - CAM) / Dims<KIND>::D;
            b.counter = f->d_counter;
            b.prof = prof + 8;
            const int nblk = 4 * ((n1 + 127) / 128);  // whole 128-row sweep tiles are covered, F2_COLS = 32 columns per CTA
            launch_k(k_f2<KIND>, nblk, F2_THREADS, F2Smem<KIND>::bytes(kc, m4), st, ordered, b);
            if (!ordered) cudaEventRecord(f->ev_ready, st);
            cur ^= 1;
            SweepExtra ex{};
            ex.n = n1; ex.add_q = a.add_q; ex.nz = f->nz;
            ex.apply_kind = -1;
            ex.lower_only = f->lower_only;
            ex.prof = prof + 12;
            ex.cta_prof = f->d_ctaprof;
            // the sweep leaves room for k_f1 of the next update: two CTAs fewer than fill the GPU
            if (ordered) {
                f->pending.valid = true;
                f->pending.n1 = n1; f->pending.m4 = m4; f->pending.ex = ex;
                f->pending.pin = f->d_P; f->pending.pout = f->d_P2;
            } else {
                // contiguous tile ranges here: with the dynamic tail the lone sweep CTA next to k_f1 would
                // take MORE tiles and slow k_f1 (the critical kernel of this mode) further
                launch_sweep(f, n1, m4, 2 << 4, ex, -1, f->d_P, f->d_P2, f->spare_ctas, false, true, true);
            }
            std::swap(f->d_P, f->d_P2);
            f->lazy = true; f->prev_m4 = m4; f->prev_add_q = a.add_q;
        }
        f->dx_last = f->d_dx[cur];
        f->state_on_main = false;
        f->state_at_ready = !(f->defer && f->pdl);   // per-frame mode: ev_ready was recorded right after the last k_f2
    } else if (!do_update || k == 0) {
        flush_pending(f, false);
        f->lazy = false;
        f->state_at_ready = false;
        f->state_on_main = false;
        // predict() / add_marker() on their own: diagonal add, covariance growth, new states
        long long work = (long long)n0 + (long long)(n1 - n0) * (n1 + n0);
        int grid = (int)std::min<long long>(1 + work / 1024, 592);
        k_prep<KIND><<<grid, 256, 0, st>>>(f->d_state, f->d_P, ld, n0, n1, f->nz, idx, pose, isnew, k,
                                           do_predict ? 1 : 0);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(ASB_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return ASB_OK;
}

static int launch_frame_dyn(asb_filter* f, const int* idx, const double* pose, const unsigned char* isnew,
                            int k, int n0, int n1, bool do_predict, bool do_update) {
    return f->kind == ASB_KIND_EKF ? launch_frame<0>(f, idx, pose, isnew, k, n0, n1, do_predict, do_update)
                                   : launch_frame<1>(f, idx, pose, isnew, k, n0, n1, do_predict, do_update);
}

// Stage k observations into the next ring slot and copy them to the device (async).
// Corner-fed frames (asb_observe_corners): `poses` is null and the slot's device poses come from the
// IPPE kernel (ippe.cu) run on the same stream; they are also copied back into the slot's pinned
// host poses (ev_pose) for the caller.
struct CornerFeed {
    const float* corners;
    const double* camera_matrix;
    const double* dist_coeffs;
    double marker_size;
};

static int stage(asb_filter* f, const std::vector<int>& idx, const double* poses,
                 const std::vector<unsigned char>& isnew, int k, int** d_idx, double** d_pose,
                 unsigned char** d_new, cudaStream_t st, const CornerFeed* feed = nullptr,
                 double** h_pose_slot = nullptr) {
    const int s = f->slot;
    f->slot = (f->slot + 1) % N_SLOTS;
    ASB_CUDA(cudaEventSynchronize(f->slot_ev[s]));
    // slot layout, compact in k: poses at 0, idx at off_idx, isnew at off_new (one contiguous copy)
    const size_t off_idx = round_up16((size_t)k * 48), off_new = off_idx + round_up16((size_t)k * 4);
    const size_t bytes = off_new + (size_t)k;
    unsigned char* hs = f->h_stage + (size_t)s * f->stage_stride;
    unsigned char* ds = f->d_stage + (size_t)s * f->stage_stride;
    std::memcpy(hs + off_idx, idx.data(), k * sizeof(int));
    std::memcpy(hs + off_new, isnew.data(), k);
    if (feed) {
        const size_t o = (size_t)s * f->max_obs;
        std::memcpy(f->h_corner + 8 * o, feed->corners, (size_t)k * 8 * sizeof(float));
        ASB_CUDA(cudaMemcpyAsync(f->d_corner + 8 * o, f->h_corner + 8 * o, (size_t)k * 8 * sizeof(float),
                                 cudaMemcpyHostToDevice, st));
        ASB_CUDA(cudaMemcpyAsync(ds + off_idx, hs + off_idx, bytes - off_idx, cudaMemcpyHostToDevice, st));
        int rc = asb_ippe_square_device(f->d_corner + 8 * o, k, feed->camera_matrix, feed->dist_coeffs,
                                        feed->marker_size, reinterpret_cast<double*>(ds), nullptr, nullptr, st);
        if (rc != ASB_OK) return rc;
        ASB_CUDA(cudaMemcpyAsync(hs, ds, (size_t)k * 6 * sizeof(double), cudaMemcpyDeviceToHost, st));
        ASB_CUDA(cudaEventRecord(f->ev_pose, st));
        if (h_pose_slot) *h_pose_slot = reinterpret_cast<double*>(hs);
    } else {
        std::memcpy(hs, poses, (size_t)k * 6 * sizeof(double));
        ASB_CUDA(cudaMemcpyAsync(ds, hs, bytes, cudaMemcpyHostToDevice, st));
    }
    ASB_CUDA(cudaEventRecord(f->slot_ev[s], st));
    *d_idx = reinterpret_cast<int*>(ds + off_idx);
    *d_pose = reinterpret_cast<double*>(ds);
    *d_new = ds + off_new;
    return ASB_OK;
}

static int poll_status(asb_filter* f) {
    int st = 0;
    ASB_CUDA(cudaMemcpyAsync(&st, f->d_status, sizeof(int), cudaMemcpyDeviceToHost, f->stream));
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    if (st != 0) {
        ASB_CUDA(cudaMemsetAsync(f->d_status, 0, sizeof(int), f->stream));
        return fail(st, "innovation covariance S was not positive definite / finite");
    }
    return ASB_OK;
}

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
// Orientation-ambiguity gate (BaseFilter.correct_poses, base_filter.py:274-325), one thread per observation.
// rot_ml = R(q_cam) * from_euler('xyz', pose[3:6]); angle = |rot_ml * inv(R(q_landmark))| in degrees; the
// observation is kept (choice 1) when angle < threshold.  poses2 != null: the second IPPE solution
// (cv2.solvePnPGeneric order) is tried when the first fails (choice 2).  Unknown ids (index < 0) are dropped
// (choice 0), as the reference's `continue` does.
__global__ void k_gate(const double* __restrict__ state, const int* __restrict__ lm_index,
                       const double* __restrict__ poses1, const double* __restrict__ poses2, int k, int lm_dims,
                       double threshold_deg, int* __restrict__ choice) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= k) return;
    int out = 0;
    const int li = lm_index[o];
    if (li >= 0) {
        double qc[4] = {state[3], state[4], state[5], state[6]};
        quat_normalize(qc);
        const double* lm = state + CAM + lm_dims * li;
        double ql[4] = {lm[3], lm[4], lm[5], lm[6]};
        quat_normalize(ql);
        const double qli[4] = {ql[0], -ql[1], -ql[2], -ql[3]};
        for (int c = 0; c < 2 && out == 0; ++c) {
            const double* pose = c == 0 ? poses1 + 6 * o : (poses2 ? poses2 + 6 * o : nullptr);
            if (!pose) break;
            double qcl[4], qml[4], qd[4];
            euler_xyz_quat(pose + 3, qcl);
            hamilton(qc, qcl, qml);
            hamilton(qml, qli, qd);
            const double v = sqrt(qd[1] * qd[1] + qd[2] * qd[2] + qd[3] * qd[3]);
            double angle = 2.0 * atan2(v, fabs(qd[0])) * (180.0 / 3.14159265358979323846);
            angle = fmin(angle, 360.0 - angle);
            if (angle < threshold_deg) out = c + 1;
        }
    }
    choice[o] = out;
}


extern "C" {

// Everything of asb_filter_create that can fail after the handle exists; the caller destroys the handle on error,
// so a failed cudaMalloc (e.g. the second 289 MB covariance buffer at C3 size) strands nothing.
static int filter_build(asb_filter* f, int kind, int quat_mode, const double pose0[10], const asb_noise* cfg,
                        int max_landmarks, int device) {
    f->kind = kind;
    f->quat_mode = quat_mode;
    f->device = device;
    f->D = kind == ASB_KIND_EKF ? 3 : 10;
    f->R = kind == ASB_KIND_EKF ? 3 : 7;
    f->W = CAM + f->D;
    f->nz = to_noise(kind, cfg);
    f->cap_lm = max_landmarks;
    f->ncap = CAM + f->D * max_landmarks;
    f->ld = round_up(f->ncap, PAD);
    f->max_obs = std::max(64, std::min(max_landmarks * 2, 4096));
    f->kc_max = kind == ASB_KIND_EKF ? FrontLimits<0>::KC_MAX : FrontLimits<1>::KC_MAX;  // front.cuh
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    f->sweep_grid = prop.multiProcessorCount;
    f->smem_optin = (int)prop.sharedMemPerBlockOptin;
    ASB_CUDA(cudaStreamCreateWithFlags(&f->own_stream, cudaStreamNonBlocking));
    f->stream = f->own_stream;
    const size_t ld = f->ld;
    ASB_CUDA(dmalloc(&f->d_state, ld));
    ASB_CUDA(dmalloc(&f->d_P, ld * ld));
    {
        ASB_CUDA(dmalloc(&f->d_P2, ld * ld));
        ASB_CUDA(cudaMemsetAsync(f->d_P2, 0, ld * ld * sizeof(double), f->stream));
        ASB_CUDA(cudaStreamCreateWithFlags(&f->side, cudaStreamNonBlocking));
        ASB_CUDA(cudaEventCreateWithFlags(&f->ev_ready, cudaEventDisableTiming));
        ASB_CUDA(cudaEventCreateWithFlags(&f->ev_f1, cudaEventDisableTiming));
        const size_t nt_max = CAM + (size_t)f->D * f->kc_max;
        ASB_CUDA(dmalloc(&f->d_At, nt_max * M_MAX));
        ASB_CUDA(dmalloc(&f->d_y, (size_t)M_MAX));
        ASB_CUDA(dmalloc(&f->d_Vt, ld * F1_LDT));
        ASB_CUDA(cudaMemsetAsync(f->d_Vt, 0, ld * F1_LDT * sizeof(double), f->stream));
        ASB_CUDA(dmalloc(&f->d_T, nt_max));
        ASB_CUDA(dmalloc(&f->d_counter, 1));
        f->lower_only = 1;
        if (std::getenv("ASB_CTA_PROF")) {
            ASB_CUDA(dmalloc(&f->d_ctaprof, 3 * 1024));
            ASB_CUDA(cudaMemsetAsync(f->d_ctaprof, 0, 3 * 1024 * sizeof(unsigned long long), f->stream));
        }
        ASB_CUDA(cudaMemsetAsync(f->d_counter, 0, sizeof(unsigned int), f->stream));
    }
    ASB_CUDA(dmalloc(&f->d_tail, 8));
    ASB_CUDA(cudaMemsetAsync(f->d_tail, 0, 8 * sizeof(unsigned int), f->stream));
    k_smid_probe<<<4 * f->sweep_grid, 32, 0, f->stream>>>(f->d_tail + 4);   // d_tail[4] = highest %smid
    ASB_CUDA(cudaGetLastError());
    {
        unsigned int top = 0;
        ASB_CUDA(cudaMemcpyAsync(&top, f->d_tail + 4, sizeof top, cudaMemcpyDeviceToHost, f->stream));
        ASB_CUDA(cudaStreamSynchronize(f->stream));
        f->reserve_sm = (int)top;
    }
    ASB_CUDA(dmalloc(&f->d_Vr, (size_t)M_MAX * V_LDB * (ld / V_BLK)));
    ASB_CUDA(dmalloc(&f->d_Vc, (size_t)M_MAX * V_LDB * (ld / V_BLK)));
    ASB_CUDA(cudaMemsetAsync(f->d_Vr, 0, (size_t)M_MAX * V_LDB * (ld / V_BLK) * sizeof(double), f->stream));
    ASB_CUDA(cudaMemsetAsync(f->d_Vc, 0, (size_t)M_MAX * V_LDB * (ld / V_BLK) * sizeof(double), f->stream));
    ASB_CUDA(dmalloc(&f->d_blocks, (size_t)f->max_obs * f->R * f->W));
    ASB_CUDA(dmalloc(&f->d_nu, (size_t)f->max_obs * f->R));
    ASB_CUDA(dmalloc(&f->d_dx[0], ld));
    ASB_CUDA(dmalloc(&f->d_dx[1], ld));
    f->dx_last = f->d_dx[0];
    ASB_CUDA(dmalloc(&f->d_unc, 16));
    ASB_CUDA(dmalloc(&f->d_status, 1));
    ASB_CUDA(dmalloc(&f->d_prof, 32));
    ASB_CUDA(cudaMemsetAsync(f->d_prof, 0, 32 * sizeof(unsigned long long), f->stream));
    f->stage_stride = round_up16((size_t)f->max_obs * 48) + round_up16((size_t)f->max_obs * 4) + round_up16((size_t)f->max_obs);
    ASB_CUDA(dmalloc(&f->d_stage, (size_t)N_SLOTS * f->stage_stride));
    ASB_CUDA(cudaMallocHost(&f->h_stage, (size_t)N_SLOTS * f->stage_stride));
    ASB_CUDA(cudaMallocHost(&f->h_mirror, ((size_t)f->ncap + 2) * sizeof(double)));
    ASB_CUDA(cudaMallocHost(&f->h_corner, (size_t)N_SLOTS * f->max_obs * 8 * sizeof(float)));
    ASB_CUDA(dmalloc(&f->d_corner, (size_t)N_SLOTS * f->max_obs * 8));
    ASB_CUDA(cudaEventCreateWithFlags(&f->ev_pose, cudaEventDisableTiming));
    for (auto& e : f->slot_ev) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ASB_CUDA(cudaMemsetAsync(f->d_P, 0, ld * ld * sizeof(double), f->stream));
    ASB_CUDA(cudaMemsetAsync(f->d_state, 0, ld * sizeof(double), f->stream));
    ASB_CUDA(cudaMemsetAsync(f->d_dx[0], 0, ld * sizeof(double), f->stream));
    ASB_CUDA(cudaMemsetAsync(f->d_dx[1], 0, ld * sizeof(double), f->stream));
    ASB_CUDA(cudaMemsetAsync(f->d_status, 0, sizeof(int), f->stream));
    // state = pose0; P = init_cam * I10  (EKF:46-48)
    std::vector<double> p0(CAM * f->ld, 0.0);
    for (int i = 0; i < CAM; ++i) p0[(size_t)i * f->ld + i] = f->nz.init_cam;
    ASB_CUDA(cudaMemcpyAsync(f->d_P, p0.data(), p0.size() * sizeof(double), cudaMemcpyHostToDevice, f->stream));
    double pose_default[CAM] = {0, 0, 0, 1, 0, 0, 0, 0, 0, 0};
    ASB_CUDA(cudaMemcpyAsync(f->d_state, pose0 ? pose0 : pose_default, CAM * sizeof(double),
                             cudaMemcpyHostToDevice, f->stream));
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    // opt in to large dynamic shared memory
    const int big = (int)prop.sharedMemPerBlockOptin - 2048;  // static smem (mbarriers) counts too
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepBig>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepDual>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepDual, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f2<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f2<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    // One shared-memory carve-out (the maximum) for every kernel of the update: an SM only changes
    // its L1 / shared split when idle, so kernels that prefer different splits neither overlap
    // (k_f1 next to the sweep) nor follow each other without a drain.
    {
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepBig>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepDual>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepDual, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<0, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<0, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<1, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<1, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f2<0>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f2<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_prep<0>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_prep<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    }
    return ASB_OK;
}

int asb_filter_create(int kind, int quat_mode, const double pose0[10], const asb_noise* cfg,
                      int max_landmarks, int device, asb_filter** out) {
    if (!out) return fail(ASB_ERR_BAD_ARG, "out is NULL");
    *out = nullptr;
    if (kind != ASB_KIND_EKF && kind != ASB_KIND_ROT) return fail(ASB_ERR_BAD_ARG, "kind %d", kind);
    if (quat_mode != ASB_QUAT_AS_WRITTEN && quat_mode != ASB_QUAT_SCALAR_FIRST)
        return fail(ASB_ERR_BAD_ARG, "quat_mode %d", quat_mode);
    if (max_landmarks < 1) return fail(ASB_ERR_BAD_ARG, "max_landmarks %d", max_landmarks);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(ASB_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(ASB_ERR_BAD_ARG, "device %d of %d", device, ndev);
    ASB_CUDA(cudaSetDevice(device));
    asb_filter* f = new (std::nothrow) asb_filter();
    if (!f) return fail(ASB_ERR_BAD_ARG, "out of host memory");
    const int rc = filter_build(f, kind, quat_mode, pose0, cfg, max_landmarks, device);
    if (rc != ASB_OK) {
        asb_filter_destroy(f);      // frees whatever was allocated (null members are skipped by the runtime)
        cudaGetLastError();         // drop the error state the clean-up of null handles may have left
        return rc;
    }
    *out = f;
    return ASB_OK;
}

int asb_filter_destroy(asb_filter* f) {
    ASB_CHECK_HANDLE(f);
    cudaStreamSynchronize(f->stream);
    if (f->side) { cudaStreamSynchronize(f->side); cudaStreamDestroy(f->side); }
    if (f->ev_ready) cudaEventDestroy(f->ev_ready);
    if (f->ev_f1) cudaEventDestroy(f->ev_f1);
    cudaFree(f->d_tail); cudaFree(f->d_P2); cudaFree(f->d_At); cudaFree(f->d_y); cudaFree(f->d_T); cudaFree(f->d_counter); cudaFree(f->d_ctaprof); cudaFree(f->d_Vt);
    cudaFree(f->d_state); cudaFree(f->d_P); cudaFree(f->d_Vr); cudaFree(f->d_Vc); cudaFree(f->d_blocks); cudaFree(f->d_nu);
    cudaFree(f->d_dx[0]); cudaFree(f->d_dx[1]); cudaFree(f->d_unc);
    cudaFree(f->d_status); cudaFree(f->d_prof); cudaFree(f->d_stage);
    cudaFree(f->s_idx); cudaFree(f->s_pose); cudaFree(f->s_new); cudaFree(f->d_gate); cudaFree(f->d_gate_i); cudaFree(f->d_hist);
    cudaFreeHost(f->h_stage); cudaFreeHost(f->h_mirror);
    cudaFreeHost(f->h_corner); cudaFree(f->d_corner);
    if (f->ev_pose) cudaEventDestroy(f->ev_pose);
    for (auto& e : f->slot_ev) cudaEventDestroy(e);
    cudaStreamDestroy(f->own_stream);
    f->magic = 0;
    delete f;
    return ASB_OK;
}

int asb_filter_set_stream(asb_filter* f, void* cuda_stream) {
    ASB_CHECK_HANDLE(f);
    f->lazy = false;   // the next k_f1 reads the current covariance buffer, not the previous sweep's input
    f->state_at_ready = false;
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    f->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : f->own_stream;
    return ASB_OK;
}

int asb_filter_synchronize(asb_filter* f) {
    ASB_CHECK_HANDLE(f);
    return poll_status(f);
}

static int observe_impl(asb_filter* f, const int32_t* ids, const double* poses, int k, bool add_new,
                        bool do_predict, const CornerFeed* feed = nullptr, double* poses_out = nullptr) {
    if (k < 0 || (k > 0 && (!ids || (!poses && !feed)))) return fail(ASB_ERR_BAD_ARG, "bad ids/poses/k");
    if (k > f->max_obs) return fail(ASB_ERR_CAPACITY, "%d observations > per-frame capacity %d", k, f->max_obs);
    std::vector<int> idx(k);
    std::vector<unsigned char> isnew(k, 0);
    const int n0 = f->n_state();
    if (add_new) {
        int fresh = 0;
        {
            std::unordered_map<int32_t, int> seen;
            for (int o = 0; o < k; ++o)
                if (!f->map.count(ids[o]) && !seen.count(ids[o])) { seen[ids[o]] = 1; ++fresh; }
        }
        if (f->nlm + fresh > f->cap_lm)
            return fail(ASB_ERR_CAPACITY, "%d landmarks > max_landmarks %d", f->nlm + fresh, f->cap_lm);
        for (int o = 0; o < k; ++o) {
            auto it = f->map.find(ids[o]);
            if (it == f->map.end()) {
                idx[o] = f->nlm;
                isnew[o] = 1;
                f->map[ids[o]] = f->nlm++;
                f->id_order.push_back(ids[o]);
                ++f->map_gen;
            } else {
                idx[o] = it->second;
            }
        }
    } else {
        for (int o = 0; o < k; ++o) {
            auto it = f->map.find(ids[o]);
            if (it == f->map.end()) return fail(ASB_ERR_UNKNOWN_ID, "unknown landmark id %d", ids[o]);
            idx[o] = it->second;
        }
    }
    const int n1 = f->n_state();
    int* d_idx = nullptr; double* d_pose = nullptr; unsigned char* d_new = nullptr;
    if (k > 0 && add_new && do_predict && f->side && n1 <= f->small_n_max &&
        small_frame_smem(f->kind, f->nlm, k) <= (size_t)f->smem_optin) {
        // small map: the whole frame in one launch on the main stream (no side stream, no ping-pong)
        flush_pending(f, false);
        if (f->motion_window > 0)
            k_motion<<<1, 32, 0, f->stream>>>(f->d_state, f->d_hist, f->motion_frames++, f->motion_window);
        double* h_slot = nullptr;
        int rc = ASB_OK;
        int mapped_slot = -1;
        if (feed) {
            rc = stage(f, idx, poses, isnew, k, &d_idx, &d_pose, &d_new, f->stream, feed, &h_slot);
            if (rc != ASB_OK) return rc;
        } else {
            // no copy operation: the kernel reads the k observations from the page-locked slot itself (pinned host
            // memory is device-addressable under unified addressing); the slot is free again when the kernel is done
            const int s = f->slot;
            f->slot = (f->slot + 1) % N_SLOTS;
            ASB_CUDA(cudaEventSynchronize(f->slot_ev[s]));
            const size_t off_idx = round_up16((size_t)k * 48), off_new = off_idx + round_up16((size_t)k * 4);
            unsigned char* hs = f->h_stage + (size_t)s * f->stage_stride;
            std::memcpy(hs, poses, (size_t)k * 6 * sizeof(double));
            std::memcpy(hs + off_idx, idx.data(), k * sizeof(int));
            std::memcpy(hs + off_new, isnew.data(), k);
            d_pose = reinterpret_cast<double*>(hs);
            d_idx = reinterpret_cast<int*>(hs + off_idx);
            d_new = hs + off_new;
            mapped_slot = s;
        }
        SmallFrame a{};
        a.kind = f->kind; a.as_written = (f->kind == ASB_KIND_EKF && f->quat_mode == ASB_QUAT_AS_WRITTEN) ? 1 : 0;
        a.state = f->d_state; a.P = f->d_P; a.ld = f->ld;
        a.idx = d_idx; a.isnew = d_new; a.poses = d_pose; a.k = k;
        a.nlm0 = (n0 - CAM) / f->D; a.nlm1 = f->nlm;
        a.nz = f->nz;
        a.status = f->d_status; a.nu_out = f->d_nu; a.blocks_out = f->d_blocks; a.dx_out = f->d_dx[0];
        a.host_state = f->h_mirror; a.host_status = reinterpret_cast<int*>(f->h_mirror + f->ncap);
        rc = small_frame_launch(a, f->stream);
        if (rc != ASB_OK) return rc;
        if (mapped_slot >= 0) ASB_CUDA(cudaEventRecord(f->slot_ev[mapped_slot], f->stream));
        ASB_CUDA(cudaEventRecord(f->ev_ready, f->stream));
        f->state_at_ready = true;    // asb_get_state waits for this kernel only
        f->state_on_main = true;
        f->lazy = false;             // the next pipelined update reads d_P as it is
        f->side_waited = false;
        f->dx_last = f->d_dx[0];
        f->last_k = k;
        f->last_idx = d_idx;
        if (feed && poses_out) {
            ASB_CUDA(cudaEventSynchronize(f->ev_pose));
            std::memcpy(poses_out, h_slot, (size_t)k * 6 * sizeof(double));
        }
        return ASB_OK;
    }
    if (k > 0) {
        // a pipelined update reads its observations in k_f1 on the side stream
        const bool piped = true;
        if (piped) pipe_wait(f);
        if (do_predict && f->motion_window > 0) {   // on the side stream, ahead of this frame's k_f1
            k_motion<<<1, 32, 0, f->side>>>(f->d_state, f->d_hist, f->motion_frames++, f->motion_window);
        }
        double* h_slot = nullptr;
        int rc = stage(f, idx, poses, isnew, k, &d_idx, &d_pose, &d_new, piped ? f->side : f->stream, feed, &h_slot);
        if (rc != ASB_OK) return rc;
        f->last_k = k;
        f->last_idx = d_idx;
        rc = launch_frame_dyn(f, d_idx, d_pose, d_new, k, n0, n1, do_predict, true);
        if (rc == ASB_OK && feed && poses_out) {
            // only the IPPE kernel and its 48 k bytes are waited for; the update stays queued
            ASB_CUDA(cudaEventSynchronize(f->ev_pose));
            std::memcpy(poses_out, h_slot, (size_t)k * 6 * sizeof(double));
        }
        return rc;
    }
    f->last_k = k;
    f->last_idx = d_idx;
    return launch_frame_dyn(f, d_idx, d_pose, d_new, k, n0, n1, do_predict, true);
}

int asb_observe(asb_filter* f, const int32_t* ids, const double* poses, int k) {
    ASB_CHECK_HANDLE(f);
    return observe_impl(f, ids, poses, k, true, true);
}

// process_frame's hot path in one call (BASE:196-204): estimate_pose_of_markers on the k detected
// corner sets and observe(ids, those poses), without the poses leaving the device in between.
int asb_observe_corners(asb_filter* f, const int32_t* ids, const float* corners, int k,
                        const double camera_matrix[9], const double dist_coeffs[5], double marker_size,
                        double* poses_out) {
    ASB_CHECK_HANDLE(f);
    if (k <= 0 || !ids || !corners || !camera_matrix || !dist_coeffs) return fail(ASB_ERR_BAD_ARG, "bad ids/corners/k/intrinsics");
    CornerFeed feed{corners, camera_matrix, dist_coeffs, marker_size};
    return observe_impl(f, ids, nullptr, k, true, true, &feed, poses_out);
}

int asb_set_small_map_limit(asb_filter* f, int n_state_max) {
    ASB_CHECK_HANDLE(f);
    if (n_state_max < 0) return fail(ASB_ERR_BAD_ARG, "n_state_max must be >= 0");
    f->small_n_max = n_state_max;
    return ASB_OK;
}

int asb_predict(asb_filter* f) {
    ASB_CHECK_HANDLE(f);
    const int n = f->n_state();
    return launch_frame_dyn(f, nullptr, nullptr, nullptr, 0, n, n, true, false);
}

int asb_update(asb_filter* f, const int32_t* ids, const double* poses, int k) {
    ASB_CHECK_HANDLE(f);
    if (k <= 0) return fail(ASB_ERR_BAD_ARG, "update with no observations");
    return observe_impl(f, ids, poses, k, false, false);
}

int asb_add_marker(asb_filter* f, int32_t id, const double pose[6], const double* uncertainity) {
    ASB_CHECK_HANDLE(f);
    if (!pose) return fail(ASB_ERR_BAD_ARG, "pose is NULL");
    if (f->nlm + 1 > f->cap_lm) return fail(ASB_ERR_CAPACITY, "max_landmarks %d reached", f->cap_lm);
    const int n0 = f->n_state();
    std::vector<int> idx(1, f->nlm);
    std::vector<unsigned char> isnew(1, 1);
    f->map[id] = f->nlm++;  // dict assignment: a known id is re-bound (EKF:253)
    ++f->map_gen;
    f->id_order.push_back(id);
    const int n1 = f->n_state();
    int* d_idx; double* d_pose; unsigned char* d_new;
    int rc = stage(f, idx, pose, isnew, 1, &d_idx, &d_pose, &d_new, f->stream);
    if (rc != ASB_OK) return rc;
    rc = launch_frame_dyn(f, d_idx, d_pose, d_new, 1, n0, n1, false, false);
    if (rc != ASB_OK) return rc;
    if (uncertainity) {
        ASB_CUDA(cudaStreamSynchronize(f->stream));  // d_unc reuse
        ASB_CUDA(cudaMemcpyAsync(f->d_unc, uncertainity, f->D * sizeof(double), cudaMemcpyHostToDevice, f->stream));
        k_set_diag<<<1, 32, 0, f->stream>>>(f->d_P, f->ld, n0, f->D, f->d_unc);
        ASB_CUDA(cudaStreamSynchronize(f->stream));
    }
    return ASB_OK;
}

int asb_stream_upload(asb_filter* f, const int32_t* ids, const double* poses, const int64_t* offsets,
                      int n_frames) {
    ASB_CHECK_HANDLE(f);
    if (!ids || !poses || !offsets || n_frames < 1) return fail(ASB_ERR_BAD_ARG, "bad stream");
    const int64_t total = offsets[n_frames];
    std::vector<int> idx(total);
    std::vector<unsigned char> isnew(total, 0);
    std::unordered_map<int32_t, int> sim = f->map;
    int nlm = f->nlm;
    f->frames.clear();
    f->frames.reserve(n_frames);
    for (int t = 0; t < n_frames; ++t) {
        FrameRec fr;
        fr.off = offsets[t];
        fr.k = (int)(offsets[t + 1] - offsets[t]);
        if (fr.k > f->max_obs) { f->frames.clear(); return fail(ASB_ERR_CAPACITY, "frame %d: %d observations > %d", t, fr.k, f->max_obs); }
        fr.n0 = CAM + f->D * nlm;
        for (int64_t o = fr.off; o < fr.off + fr.k; ++o) {
            auto it = sim.find(ids[o]);
            if (it == sim.end()) {
                if (nlm + 1 > f->cap_lm) { f->frames.clear(); return fail(ASB_ERR_CAPACITY, "stream needs > %d landmarks", f->cap_lm); }
                idx[o] = nlm;
                isnew[o] = 1;
                sim[ids[o]] = nlm++;
                fr.new_ids.push_back(ids[o]);
            } else {
                idx[o] = it->second;
            }
        }
        fr.n1 = CAM + f->D * nlm;
        f->frames.push_back(std::move(fr));
    }
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    cudaFree(f->s_idx); cudaFree(f->s_pose); cudaFree(f->s_new);
    f->s_idx = nullptr; f->s_pose = nullptr; f->s_new = nullptr;
    ASB_CUDA(dmalloc(&f->s_idx, (size_t)std::max<int64_t>(total, 1)));
    ASB_CUDA(dmalloc(&f->s_pose, (size_t)std::max<int64_t>(total, 1) * 6));
    ASB_CUDA(dmalloc(&f->s_new, (size_t)std::max<int64_t>(total, 1)));
    ASB_CUDA(cudaMemcpy(f->s_idx, idx.data(), total * sizeof(int), cudaMemcpyHostToDevice));
    ASB_CUDA(cudaMemcpy(f->s_pose, poses, (size_t)total * 6 * sizeof(double), cudaMemcpyHostToDevice));
    ASB_CUDA(cudaMemcpy(f->s_new, isnew.data(), total, cudaMemcpyHostToDevice));
    f->next_frame = 0;
    f->stream_gen = f->map_gen;
    return ASB_OK;
}

int asb_stream_run(asb_filter* f, int t0, int t1) {
    ASB_CHECK_HANDLE(f);
    if (f->motion_window > 0) return fail(ASB_ERR_BAD_ARG, "the moving-average motion model runs with per-frame observe() only");
    if (f->frames.empty()) return fail(ASB_ERR_NO_STREAM, "stream_run before stream_upload");
    if (f->stream_gen != f->map_gen)
        return fail(ASB_ERR_NO_STREAM, "the landmark map changed since stream_upload (observe / add_marker / set_landmarks): upload again");
    if (t0 != f->next_frame || t1 < t0 || t1 > (int)f->frames.size())
        return fail(ASB_ERR_BAD_ARG, "frames must run in order: next is %d, got [%d,%d)", f->next_frame, t0, t1);
    f->defer = true;
    for (int t = t0; t < t1; ++t) {
        const FrameRec& fr = f->frames[t];
        for (int32_t id : fr.new_ids) {
            f->map[id] = f->nlm++;
            f->id_order.push_back(id);
        }
        f->last_k = fr.k;
        f->last_idx = f->s_idx + fr.off;
        if (fr.k == 0) continue;  // BASE:197-204: frames without detections do not predict
        int rc = launch_frame_dyn(f, f->s_idx + fr.off, f->s_pose + 6 * fr.off, f->s_new + fr.off, fr.k, fr.n0,
                                  fr.n1, true, true);
        if (rc != ASB_OK) { f->defer = false; flush_pending(f, false); return rc; }
        f->next_frame = t + 1;
    }
    f->defer = false;
    if (f->pending.valid) {   // last k_f2: what a per-frame k_f1 on the side stream waits for
        cudaEventRecord(f->ev_ready, f->stream);
        f->state_at_ready = true;
    }
    flush_pending(f, false);   // nothing stays pending across API calls
    f->next_frame = t1;
    return ASB_OK;
}

int asb_get_dims(asb_filter* f, int* n_landmarks, int* n_state) {
    ASB_CHECK_HANDLE(f);
    if (n_landmarks) *n_landmarks = f->nlm;
    if (n_state) *n_state = f->n_state();
    return ASB_OK;
}

int asb_get_state(asb_filter* f, double* out, int n_state) {
    ASB_CHECK_HANDLE(f);
    if (!out || n_state != f->n_state()) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, f->n_state());
    if (f->side && f->state_at_ready) {
        // The state (and the status flag) are final once the last k_f2 is done; the covariance sweep
        // queued behind it keeps running while the host reads the pose and prepares the next frame.
        // (After a one-launch small-map frame nothing is queued behind: the copies stay on the main stream.)
        int st = 0;
        if (f->state_on_main) {
            // the kernel left the posterior state and the status in the page-locked mirror
            ASB_CUDA(cudaEventSynchronize(f->ev_ready));
            st = *reinterpret_cast<const int*>(f->h_mirror + f->ncap);
            std::memcpy(out, f->h_mirror, (size_t)n_state * sizeof(double));
        } else {
            ASB_CUDA(cudaStreamWaitEvent(f->side, f->ev_ready, 0));
            ASB_CUDA(cudaMemcpyAsync(&st, f->d_status, sizeof(int), cudaMemcpyDeviceToHost, f->side));
            ASB_CUDA(cudaMemcpyAsync(out, f->d_state, n_state * sizeof(double), cudaMemcpyDeviceToHost, f->side));
            ASB_CUDA(cudaStreamSynchronize(f->side));
        }
        if (st != 0) {
            ASB_CUDA(cudaMemsetAsync(f->d_status, 0, sizeof(int), f->stream));
            return fail(st, "innovation covariance S was not positive definite / finite");
        }
        return ASB_OK;
    }
    int rc = poll_status(f);
    ASB_CUDA(cudaMemcpy(out, f->d_state, n_state * sizeof(double), cudaMemcpyDeviceToHost));
    return rc;
}

int asb_set_state(asb_filter* f, const double* in, int n_state) {
    ASB_CHECK_HANDLE(f);
    f->lazy = false;   // the next k_f1 reads the current covariance buffer, not the previous sweep's input
    f->state_at_ready = false;
    if (!in || n_state != f->n_state()) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, f->n_state());
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    ASB_CUDA(cudaMemcpy(f->d_state, in, n_state * sizeof(double), cudaMemcpyHostToDevice));
    return ASB_OK;
}

int asb_get_cov(asb_filter* f, double* out, int n_state) {
    ASB_CHECK_HANDLE(f);
    if (!out || n_state != f->n_state()) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, f->n_state());
    if (f->lower_only) {   // the upper triangle is stale: rebuild it from the lower one (in place)
        const int nt = (n_state + 31) / 32;
        k_symmetrize<<<dim3(nt, nt), dim3(32, 8), 0, f->stream>>>(f->d_P, f->ld, n_state);
        ASB_CUDA(cudaGetLastError());
    }
    int rc = poll_status(f);
    ASB_CUDA(cudaMemcpy2D(out, (size_t)n_state * sizeof(double), f->d_P, (size_t)f->ld * sizeof(double),
                          (size_t)n_state * sizeof(double), n_state, cudaMemcpyDeviceToHost));
    return rc;
}

int asb_set_cov(asb_filter* f, const double* in, int n_state) {
    ASB_CHECK_HANDLE(f);
    f->lazy = false;   // the next k_f1 reads the current covariance buffer, not the previous sweep's input
    f->state_at_ready = false;
    if (!in || n_state != f->n_state()) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, f->n_state());
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    ASB_CUDA(cudaMemcpy2D(f->d_P, (size_t)f->ld * sizeof(double), in, (size_t)n_state * sizeof(double),
                          (size_t)n_state * sizeof(double), n_state, cudaMemcpyHostToDevice));
    return ASB_OK;
}

int asb_get_cov_diag(asb_filter* f, double* out, int n_state) {
    ASB_CHECK_HANDLE(f);
    if (!out || n_state != f->n_state()) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, f->n_state());
    int rc = poll_status(f);
    ASB_CUDA(cudaMemcpy2D(out, sizeof(double), f->d_P, (size_t)(f->ld + 1) * sizeof(double), sizeof(double),
                          n_state, cudaMemcpyDeviceToHost));
    return rc;
}

int asb_get_landmark_ids(asb_filter* f, int32_t* ids, int capacity) {
    ASB_CHECK_HANDLE(f);
    if (!ids || capacity < f->nlm) return fail(ASB_ERR_BAD_ARG, "capacity %d < %d", capacity, f->nlm);
    std::copy(f->id_order.begin(), f->id_order.end(), ids);
    return ASB_OK;
}

int asb_set_landmarks(asb_filter* f, const int32_t* ids, int n) {
    ASB_CHECK_HANDLE(f);
    f->lazy = false;   // the next k_f1 reads the current covariance buffer, not the previous sweep's input
    f->state_at_ready = false;
    if (n < 0 || (n > 0 && !ids)) return fail(ASB_ERR_BAD_ARG, "bad ids");
    if (n > f->cap_lm) return fail(ASB_ERR_CAPACITY, "%d landmarks > max_landmarks %d", n, f->cap_lm);
    f->map.clear();
    f->id_order.assign(ids, ids + n);
    for (int i = 0; i < n; ++i) f->map[ids[i]] = i;
    ++f->map_gen;
    f->nlm = n;
    return ASB_OK;
}

int asb_device_views(asb_filter* f, void** d_state, void** d_cov, int* ld) {
    ASB_CHECK_HANDLE(f);
    f->lazy = false;   // the next k_f1 reads the current covariance buffer, not the previous sweep's input
    f->state_at_ready = false;
    if (d_state) *d_state = f->d_state;
    if (d_cov) *d_cov = f->d_P;
    if (ld) *ld = f->ld;
    return ASB_OK;
}

int asb_bench_sweep(asb_filter* f, int m4, int config, int dbg, int reps, double* us_per_launch) {
    ASB_CHECK_HANDLE(f);
    f->lazy = false;   // the next k_f1 reads the current covariance buffer, not the previous sweep's input
    f->state_at_ready = false;
    if (!us_per_launch || m4 < 4 || m4 > M_MAX || (m4 & 3) || reps < 1) return fail(ASB_ERR_BAD_ARG, "bad args");
    const int n1 = f->n_state();
    cudaEvent_t e0, e1;
    ASB_CUDA(cudaEventCreate(&e0));
    ASB_CUDA(cudaEventCreate(&e1));
    SweepExtra ex{};
    ex.n = n1; ex.apply_kind = -1;
    ex.pol = (dbg >> 12) & 0x1ff;          // experiments: cache policies, serpentine walk
    for (int rep = -2; rep < reps; ++rep) {
        if (rep == 0) ASB_CUDA(cudaEventRecord(e0, f->stream));
        ex.reverse = (dbg & (1 << 25)) ? 1 : ((dbg & (1 << 24)) ? (rep & 1) : 0);
        launch_sweep(f, n1, m4, dbg & 0xfff, ex, config);
    }
    ASB_CUDA(cudaEventRecord(e1, f->stream));
    ASB_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    ASB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *us_per_launch = 1e3 * ms / reps;
    return ASB_OK;
}

int asb_front_profile(asb_filter* f, double us[12]) {
    ASB_CHECK_HANDLE(f);
    if (!us) return fail(ASB_ERR_BAD_ARG, "us is NULL");
    unsigned long long t[16];
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    ASB_CUDA(cudaMemcpy(t, f->d_prof + 16 * (f->prof_slot ^ 1), sizeof t, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 12; ++i) us[i] = 0.0;
    {
        // k_f1: [index set, issue copies, model + wait, lazy operand + Q + S, Cholesky, outputs]; k_f2: [loads, product, stores]
        for (int i = 0; i < 5; ++i) us[i] = 1e-3 * (double)(long long)(t[i + 1] - t[i]);
        for (int i = 0; i < 3; ++i) us[8 + i] = 1e-3 * (double)(long long)(t[9 + i] - t[8 + i]);
        us[5] = 1e-3 * (double)(long long)(t[5] - t[0]);    // k_f1 start -> end
        us[6] = 1e-3 * (double)(long long)(t[8] - t[0]);    // k_f1 start -> k_f2 start (same update)
        us[7] = 1e-3 * (double)(long long)(t[11] - t[0]);   // k_f1 start -> k_f2 end
    }
    return ASB_OK;
}

int asb_timeline(asb_filter* f, double us[32]) {
    ASB_CHECK_HANDLE(f);
    if (!us) return fail(ASB_ERR_BAD_ARG, "us is NULL");
    unsigned long long t[32];
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    ASB_CUDA(cudaMemcpy(t, f->d_prof, sizeof t, cudaMemcpyDeviceToHost));
    unsigned long long t0 = ~0ull;
    for (int i = 0; i < 32; ++i) if (t[i] && t[i] < t0) t0 = t[i];
    // previous update first: the slot written last is prof_slot ^ 1
    for (int u = 0; u < 2; ++u)
        for (int i = 0; i < 16; ++i) {
            const unsigned long long v = t[16 * (u == 0 ? f->prof_slot : f->prof_slot ^ 1) + i];
            us[16 * u + i] = v ? 1e-3 * (double)(long long)(v - t0) : -1.0;
        }
    return ASB_OK;
}

int asb_cta_profile(asb_filter* f, double* us, int n) {
    ASB_CHECK_HANDLE(f);
    if (!us || n < 1 || n > 1024) return fail(ASB_ERR_BAD_ARG, "bad args");
    if (!f->d_ctaprof) return fail(ASB_ERR_BAD_ARG, "set ASB_CTA_PROF=1 before creating the filter");
    std::vector<unsigned long long> t(3 * 1024);
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    ASB_CUDA(cudaMemcpy(t.data(), f->d_ctaprof, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    unsigned long long t0 = ~0ull;
    for (int b = 0; b < n; ++b) if (t[3 * b] && t[3 * b] < t0) t0 = t[3 * b];
    for (int b = 0; b < n; ++b) {
        us[3 * b] = t[3 * b] ? 1e-3 * (double)(long long)(t[3 * b] - t0) : -1.0;
        us[3 * b + 1] = t[3 * b + 1] ? 1e-3 * (double)(long long)(t[3 * b + 1] - t0) : -1.0;
        us[3 * b + 2] = (double)t[3 * b + 2];
    }
    return ASB_OK;
}

int asb_get_last_update(asb_filter* f, double* nu, double* blocks, double* dx, int* m) {
    ASB_CHECK_HANDLE(f);
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    const int k = f->last_k;
    if (m) *m = f->R * k;
    if (nu && k) ASB_CUDA(cudaMemcpy(nu, f->d_nu, (size_t)f->R * k * sizeof(double), cudaMemcpyDeviceToHost));
    if (blocks && k)
        ASB_CUDA(cudaMemcpy(blocks, f->d_blocks, (size_t)k * f->R * f->W * sizeof(double), cudaMemcpyDeviceToHost));
    if (dx) ASB_CUDA(cudaMemcpy(dx, f->dx_last, (size_t)f->n_state() * sizeof(double), cudaMemcpyDeviceToHost));
    return ASB_OK;
}

int asb_correct_poses(asb_filter* f, const int32_t* ids, const double* poses, const double* poses2_or_null, int k,
                      double threshold_deg, int32_t* choice_out) {
    ASB_CHECK_HANDLE(f);
    if (f->kind != ASB_KIND_ROT) return fail(ASB_ERR_BAD_ARG, "correct_poses needs landmark orientations (EKF_Rotations)");
    if (k < 0 || (k > 0 && (!ids || !poses || !choice_out))) return fail(ASB_ERR_BAD_ARG, "bad correct_poses arguments");
    if (k == 0) return ASB_OK;
    if (k > f->gate_cap) {
        ASB_CUDA(cudaStreamSynchronize(f->stream));
        cudaFree(f->d_gate); cudaFree(f->d_gate_i);
        f->d_gate = nullptr; f->d_gate_i = nullptr; f->gate_cap = 0;
        const int cap = std::max(k, 64);
        ASB_CUDA(cudaMalloc(&f->d_gate, (size_t)cap * 12 * sizeof(double)));
        ASB_CUDA(cudaMalloc(&f->d_gate_i, (size_t)cap * 2 * sizeof(int)));
        f->gate_cap = cap;
    }
    std::vector<int> idx(k);
    for (int o = 0; o < k; ++o) {
        auto it = f->map.find(ids[o]);
        idx[o] = it == f->map.end() ? -1 : it->second;
    }
    double* d_p1 = f->d_gate;
    double* d_p2 = poses2_or_null ? f->d_gate + (size_t)f->gate_cap * 6 : nullptr;
    int* d_idx = f->d_gate_i;
    int* d_choice = f->d_gate_i + f->gate_cap;
    ASB_CUDA(cudaMemcpyAsync(d_idx, idx.data(), (size_t)k * sizeof(int), cudaMemcpyHostToDevice, f->stream));
    ASB_CUDA(cudaMemcpyAsync(d_p1, poses, (size_t)k * 6 * sizeof(double), cudaMemcpyHostToDevice, f->stream));
    if (d_p2) ASB_CUDA(cudaMemcpyAsync(d_p2, poses2_or_null, (size_t)k * 6 * sizeof(double), cudaMemcpyHostToDevice, f->stream));
    k_gate<<<(k + 63) / 64, 64, 0, f->stream>>>(f->d_state, d_idx, d_p1, d_p2, k, f->D, threshold_deg, d_choice);
    ASB_CUDA(cudaGetLastError());
    ASB_CUDA(cudaMemcpyAsync(choice_out, d_choice, (size_t)k * sizeof(int), cudaMemcpyDeviceToHost, f->stream));
    ASB_CUDA(cudaStreamSynchronize(f->stream));   // idx (host vector) and choice_out are done with
    return ASB_OK;
}

int asb_set_motion_model(asb_filter* f, int window) {
    ASB_CHECK_HANDLE(f);
    if (window < 0 || window > 4096) return fail(ASB_ERR_BAD_ARG, "window %d", window);
    ASB_CUDA(cudaStreamSynchronize(f->stream));
    if (f->side) ASB_CUDA(cudaStreamSynchronize(f->side));
    cudaFree(f->d_hist);
    f->d_hist = nullptr;
    f->motion_window = window;
    f->motion_frames = 0;
    if (window > 0) {
        ASB_CUDA(cudaMalloc(&f->d_hist, (size_t)window * 3 * sizeof(double)));
        ASB_CUDA(cudaMemset(f->d_hist, 0, (size_t)window * 3 * sizeof(double)));
    }
    return ASB_OK;
}

}  // extern "C"
