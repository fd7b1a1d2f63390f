// Batched engine host side + C ABI ("batched independent filters" section of
// include/aruco_slam_b200.h).  See batch_kernels.cuh for the kernel.
#include <algorithm>
#include <cstdlib>
#include <new>
#include <vector>

#include "batch_kernels.cuh"
#include "common.cuh"
#include "wide_kernels.cuh"

using namespace asb;

namespace {
constexpr uint32_t BATCH_MAGIC = 0x41534242;  // "ASBB"
}

struct asb_batch {
    uint32_t magic = BATCH_MAGIC;
    int kind = 0, quat_mode = 0, device = 0, D = 3, R = 3;
    Noise nz{};
    int n_seq = 0, max_lm = 0, max_obs = 0, ncap = 0;
    long long state_stride = 0, cov_stride = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    double *d_states = nullptr, *d_covs = nullptr, *d_poses = nullptr, *d_dense = nullptr, *d_pose0 = nullptr;
    int *d_nlm = nullptr, *d_tab = nullptr, *d_status = nullptr, *d_ids = nullptr;
    long long* d_off = nullptr;
    unsigned long long* d_prof = nullptr;
    bool profiling = false;   // asb_batch_profile was called: later runs use the instrumented instantiation of k_batch
    int n_frames = 0;
    int obs_seqs = 0;          // distinct observation streams on the device (n_seq, or fewer after an upload_tiled)
    long long n_obs = 0;
    int sm_count = 0, smem_optin = 0;
#ifndef ASB_BATCH_CK
#define ASB_BATCH_CK 4
#endif
    int ck = ASB_BATCH_CK;       // markers per chunk
    int threads = 128;
    // Pipelined host path: asb_batch_upload copies the observations in N_GROUPS sequence groups on a
    // copy stream; the asb_batch_run that follows launches one kernel per group as soon as that
    // group's bytes have landed (two run streams, so the tail of one group's kernel is filled by the
    // next), i.e. all but the first group's H2D time hides behind compute.
    static constexpr int N_GROUPS = 8;
    cudaStream_t copy_stream = nullptr, run_streams[2] = {nullptr, nullptr};
    cudaEvent_t ev_group[N_GROUPS]{}, ev_main = nullptr, ev_done[2] = {nullptr, nullptr};
    bool fresh_upload = false;
    int groups = 1, up_groups = 1;   // sequence groups of the pipelined host path / of the last upload
    // Wide engine (wide_kernels.cuh): frame-synchronous kernels, covariance tiled in HBM.  Chosen at creation
    // when the resident engine's shared-memory footprint exceeds an SM (EKF_Rotations at 32 / 64 markers),
    // or on request (asb_batch_set_engine before the first upload).
    bool wide = false;
    int engine_req = 0;            // 0 auto, 1 resident, 2 wide
    int *w_meta = nullptr, *w_idx = nullptr;
    double *w_delta = nullptr, *w_H = nullptr, *w_Hp = nullptr, *w_L = nullptr, *w_Dinv = nullptr, *w_y = nullptr, *w_V = nullptr;
    long long w_vstride = 0;
    int w_nb64 = 0;
    static constexpr int W_MAX_GROUPS = 4;
    int w_groups = 1;
    cudaStream_t w_streams[W_MAX_GROUPS]{};
    cudaEvent_t w_done[W_MAX_GROUPS]{};
    std::vector<int> kmax;         // per frame: max observations over the sequences (host, from the offsets)
};

static int check(asb_batch* b) {
    if (!b || b->magic != BATCH_MAGIC) return fail(ASB_ERR_BAD_HANDLE, "bad asb_batch handle");
    return ASB_OK;
}
#define ASB_CHECK_BATCH(b)              \
    do {                                \
        int _rc = check(b);             \
        if (_rc != ASB_OK) return _rc;  \
        cudaSetDevice((b)->device);     \
    } while (0)

template <typename T>
static cudaError_t dmalloc(T** p, size_t count) {
    return cudaMalloc(reinterpret_cast<void**>(p), std::max<size_t>(count, 1) * sizeof(T));
}

// shared-memory bytes of k_batch excluding the packed covariance (same carve-up as the kernel)
static size_t batch_smem_base(int D, int R, int CK, int max_lm, int max_obs) {
    const int W = CAM + D, MC = R * CK, LDS_S = (MC + 1) | 1;
    const int ncap = CAM + D * max_lm, npad = batch_npad(ncap);
    size_t dbl = (size_t)npad * 2 + (size_t)MC * npad + (size_t)2 * CK * ((R * W + 1) & ~1) + 2 * ((MC + 3) & ~3) +
                 (size_t)(MC + 1) * LDS_S + ((MC + 3) & ~3) + (size_t)max_obs * 6;
    size_t ints = (size_t)max_obs * 3 + max_lm + 4;
    return dbl * 8 + ints * 4 + 16;
}

template <int KIND, int CK, int NT, bool PROF = false>
static int launch_batch(asb_batch* b, const BatchParams& p, size_t smem, int grid, cudaStream_t st) {
    auto kern = k_batch<KIND, CK, NT, false, PROF>;
    ASB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ASB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    // persistent grid: exactly the CTAs that are co-resident (registers can bind before shared memory)
    int occ = 0;
    ASB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, smem));
    if (occ < 1) return fail(ASB_ERR_CAPACITY, "k_batch does not fit on an SM (%zu B shared memory)", smem);
    grid = std::min(grid, b->sm_count * occ);
    kern<<<grid, NT, smem, st>>>(p);
    ASB_CUDA(cudaGetLastError());
    return ASB_OK;
}

template <int KIND, int CK>
static int launch_batch_nt(asb_batch* b, const BatchParams& p, size_t smem, int grid, cudaStream_t st) {
    if (b->profiling) {   // phase-profile instantiation (asb_batch_profile has been called on this handle)
        if (b->threads >= 256) return launch_batch<KIND, CK, 256, true>(b, p, smem, grid, st);
        return launch_batch<KIND, CK, 128, true>(b, p, smem, grid, st);
    }
    if (b->threads >= 256) return launch_batch<KIND, CK, 256>(b, p, smem, grid, st);
    return launch_batch<KIND, CK, 128>(b, p, smem, grid, st);
}

static int launch_batch_dyn(asb_batch* b, const BatchParams& p, size_t smem, int grid, cudaStream_t st) {
    if (b->kind == ASB_KIND_EKF) return launch_batch_nt<0, ASB_BATCH_CK>(b, p, smem, grid, st);
    return launch_batch_nt<1, 4>(b, p, smem, grid, st);
}

// ---- wide engine, host side ------------------------------------------------------------------------
template <int KIND> static size_t wide_model_smem(int max_lm, int max_obs) {
    using WD = WideDims<KIND>;
    const int cap_obs = std::max(max_obs, (int)WCK);
    return ((size_t)CAM * WD::TW + (size_t)cap_obs * 6) * 8 + ((size_t)cap_obs * 3 + max_lm + 4) * 4;
}
template <int KIND> static size_t wide_chol_smem() {
    using WD = WideDims<KIND>;
    return ((size_t)tri(WD::NB) * 64 + WD::M8 + 64) * 8;
}
template <int KIND> static size_t wide_panel_smem() {
    using WD = WideDims<KIND>;
    return ((size_t)WD::M8 * W_VLD + 2 * (size_t)WD::LROW) * 8 + WCK * 4;
}
template <int KIND> static size_t wide_sweep_smem() {
    return ((size_t)WideDims<KIND>::M8 * W_VLD + (size_t)W_NST * W_SLAB * W_VLD) * 8;
}

static int resident_fits(const asb_batch* b) {
    const size_t base = batch_smem_base(b->D, b->R, b->ck, b->max_lm, b->max_obs);
    return base + (size_t)tiled_doubles(b->ncap) * sizeof(double) <= (size_t)b->smem_optin;
}

static void wide_free(asb_batch* b) {
    cudaFree(b->w_meta); cudaFree(b->w_idx); cudaFree(b->w_delta); cudaFree(b->w_H); cudaFree(b->w_Hp);
    b->w_Hp = nullptr;
    cudaFree(b->w_L); cudaFree(b->w_Dinv); cudaFree(b->w_y); cudaFree(b->w_V);
    b->w_meta = b->w_idx = nullptr;
    b->w_delta = b->w_H = b->w_L = b->w_Dinv = b->w_y = b->w_V = nullptr;
}

template <int KIND> static int wide_alloc_kind(asb_batch* b) {
    using WD = WideDims<KIND>;
    const size_t ns = (size_t)b->n_seq;
    b->w_nb64 = (b->ncap + 63) / 64;
    b->w_vstride = (long long)b->w_nb64 * WD::M8 * W_VLD;
    ASB_CUDA(dmalloc(&b->w_meta, ns * 8));
    ASB_CUDA(dmalloc(&b->w_idx, ns * b->max_obs));
    ASB_CUDA(dmalloc(&b->w_delta, ns * b->state_stride));
    ASB_CUDA(dmalloc(&b->w_H, ns * WCK * WD::R * WD::W));
    ASB_CUDA(dmalloc(&b->w_Hp, ns * WCK * 8 * WD::HLS));
    ASB_CUDA(dmalloc(&b->w_L, ns * tri(WD::NB) * 64));
    ASB_CUDA(dmalloc(&b->w_Dinv, ns * WD::NB * 64));
    ASB_CUDA(dmalloc(&b->w_y, ns * WD::M8));
    ASB_CUDA(dmalloc(&b->w_V, ns * (size_t)b->w_vstride));
    for (auto& st : b->w_streams) if (!st) ASB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& e : b->w_done) if (!e) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    b->w_groups = b->n_seq >= 1024 ? 2 : 1;
    ASB_CUDA(cudaMemset(b->w_meta, 0, ns * 8 * sizeof(int)));
    ASB_CUDA(cudaMemset(b->w_V, 0, ns * (size_t)b->w_vstride * sizeof(double)));
    return ASB_OK;
}

// (re)allocate the covariance for the chosen engine: packed lower triangle (resident) or 8 x 8 tiles (wide)
static int alloc_covs(asb_batch* b) {
    cudaFree(b->d_covs);
    b->d_covs = nullptr;
    wide_free(b);
    b->cov_stride = b->wide ? (long long)tiled_doubles(b->ncap)
                            : (((long long)b->ncap * (b->ncap + 1) / 2 + 3) & ~3LL);
    ASB_CUDA(dmalloc(&b->d_covs, (size_t)b->n_seq * b->cov_stride));
    if (b->wide) return b->kind == ASB_KIND_EKF ? wide_alloc_kind<0>(b) : wide_alloc_kind<1>(b);
    return ASB_OK;
}

template <int KIND> static int wide_run(asb_batch* b, int t0, int t1) {
    WideParams p;
    p.states = b->d_states; p.covs = b->d_covs; p.nlm = b->d_nlm; p.lm_ids = b->d_tab; p.status = b->d_status;
    p.ids = b->d_ids; p.poses = b->d_poses; p.offsets = b->d_off;
    p.state_stride = b->state_stride; p.cov_stride = b->cov_stride;
    p.n_seq = b->n_seq; p.n_frames = b->n_frames; p.max_lm = b->max_lm; p.max_obs = b->max_obs;
    p.obs_seqs = b->obs_seqs;
    p.as_written = (b->kind == ASB_KIND_EKF && b->quat_mode == ASB_QUAT_AS_WRITTEN) ? 1 : 0;
    p.nz = b->nz;
    p.meta = b->w_meta; p.idx = b->w_idx; p.delta = b->w_delta; p.H = b->w_H; p.Hp = b->w_Hp; p.L = b->w_L; p.Dinv = b->w_Dinv;
    p.y = b->w_y; p.V = b->w_V; p.v_stride = b->w_vstride; p.nb64 = b->w_nb64;
    const size_t sm_model = wide_model_smem<KIND>(b->max_lm, b->max_obs), sm_chol = wide_chol_smem<KIND>(),
                 sm_panel = wide_panel_smem<KIND>(), sm_sweep = wide_sweep_smem<KIND>();
    if (sm_model > (size_t)b->smem_optin) return fail(ASB_ERR_CAPACITY, "k_wide_model needs %zu B shared memory", sm_model);
    ASB_CUDA(cudaFuncSetAttribute(k_wide_model<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_model));
    ASB_CUDA(cudaFuncSetAttribute(k_wide_chol<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_chol));
    ASB_CUDA(cudaFuncSetAttribute(k_wide_panel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_panel));
    ASB_CUDA(cudaFuncSetAttribute(k_wide_sweep<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_sweep));
    // Sequence groups on separate streams: the latency-bound kernels of one group (model, chol) overlap the
    // tensor-pipe-bound ones (panel, sweep) of the others.  Frames stay ordered inside a group (stream order).
    const int ngroups = b->w_groups;
    ASB_CUDA(cudaEventRecord(b->ev_main, b->stream));
    auto group_stream = [&](int gidx) { return ngroups == 1 ? b->stream : b->w_streams[gidx]; };
    auto group_first = [&](int gidx) { return (int)((long long)b->n_seq * gidx / ngroups); };
    if (ngroups > 1)
        for (int gidx = 0; gidx < ngroups; ++gidx) ASB_CUDA(cudaStreamWaitEvent(group_stream(gidx), b->ev_main, 0));
    // frames outermost, groups inside: a long run queues thousands of launches, and issued group by group the
    // first group's backlog would fill the launch queue before the second stream gets anything to overlap with
    for (int t = t0; t < t1; ++t) {
        const int rounds = (b->kmax[t] + WCK - 1) / WCK;
        for (int r = 0; r < rounds; ++r)
            for (int gidx = 0; gidx < ngroups; ++gidx) {
                const int s0 = group_first(gidx), s1 = group_first(gidx + 1);
                if (s1 == s0) continue;
                cudaStream_t st = group_stream(gidx);
                p.seq0 = s0; p.t = t; p.round = r;
                const dim3 g_strips(b->w_nb64, s1 - s0);
                k_wide_model<KIND><<<s1 - s0, 128, sm_model, st>>>(p);
                k_wide_chol<KIND><<<s1 - s0, 128, sm_chol, st>>>(p);
                k_wide_panel<KIND><<<g_strips, 256, sm_panel, st>>>(p);
                k_wide_sweep<KIND><<<g_strips, W_SWEEP_THREADS, sm_sweep, st>>>(p);
            }
    }
    for (int gidx = 0; gidx < ngroups; ++gidx) {
        const int s0 = group_first(gidx), s1 = group_first(gidx + 1);
        if (s1 == s0) continue;
        cudaStream_t st = group_stream(gidx);
        p.seq0 = s0;
        k_wide_apply<KIND><<<s1 - s0, 64, 0, st>>>(p);
        if (ngroups > 1) {
            ASB_CUDA(cudaEventRecord(b->w_done[gidx], st));
            ASB_CUDA(cudaStreamWaitEvent(b->stream, b->w_done[gidx], 0));
        }
    }
    ASB_CUDA(cudaGetLastError());
    return ASB_OK;
}

// ---- one frame of a single small-map filter on the resident kernel (declared in common.cuh) ----------------
namespace asb {

size_t small_frame_smem(int kind, int nlm1, int k) {
    const int D = kind == ASB_KIND_EKF ? 3 : 10, R = kind == ASB_KIND_EKF ? 3 : 7;
    return batch_smem_base(D, R, 4, nlm1, k) + (size_t)tiled_doubles(CAM + D * nlm1) * sizeof(double);
}

template <int KIND, int NT>
static int small_launch(const BatchParams& p, size_t smem, cudaStream_t st) {
    auto kern = k_batch<KIND, 4, NT, true>;
    static size_t granted[64] = {};   // per device: the largest dynamic shared-memory size asked for so far
    int dev = 0;
    ASB_CUDA(cudaGetDevice(&dev));
    if (smem > granted[dev & 63]) {
        ASB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        granted[dev & 63] = smem;
    }
    kern<<<1, NT, smem, st>>>(p);
    ASB_CUDA(cudaGetLastError());
    return ASB_OK;
}

int small_frame_launch(const SmallFrame& a, cudaStream_t st) {
    BatchParams p{};
    p.states = a.state; p.covs = a.P; p.ld = a.ld;
    p.status = a.status;
    p.poses = a.poses;
    p.n_seq = 1; p.n_frames = 1; p.t0 = 0; p.t1 = 1; p.seq0 = 0; p.seq1 = 1; p.obs_seqs = 1;
    p.max_lm = a.nlm1; p.max_obs = a.k;
    p.as_written = a.as_written;
    p.nz = a.nz;
    p.one_idx = a.idx; p.one_new = a.isnew; p.one_k = a.k; p.one_nlm0 = a.nlm0; p.one_nlm1 = a.nlm1;
    p.nu_out = a.nu_out; p.blocks_out = a.blocks_out; p.dx_out = a.dx_out;
    p.one_state_host = a.host_state; p.one_status_host = a.host_status;
    const size_t smem = small_frame_smem(a.kind, a.nlm1, a.k);
    // 256 threads at every size: the CTA has its SM to itself, and measured against 128 threads it is faster from
    // N = 40 (41.4 vs 42.6 us per frame) to N = 150 (99.6 vs 125.8 us, EKF_Rotations) - tools/small_map_exp.py
    return a.kind == ASB_KIND_EKF ? small_launch<0, 256>(p, smem, st) : small_launch<1, 256>(p, smem, st);
}

}  // namespace asb

extern "C" {

int asb_batch_reset(asb_batch* b) {
    ASB_CHECK_BATCH(b);
    if (b->wide) {   // the wide engine relies on every covariance entry outside the live block being zero
        ASB_CUDA(cudaMemsetAsync(b->d_covs, 0, (size_t)b->n_seq * b->cov_stride * sizeof(double), b->stream));
        ASB_CUDA(cudaMemsetAsync(b->w_meta, 0, (size_t)b->n_seq * 8 * sizeof(int), b->stream));
    }
    k_batch_init<<<b->n_seq, 64, 0, b->stream>>>(b->d_states, b->d_covs, b->d_nlm, b->d_status, b->state_stride,
                                                 b->cov_stride, b->n_seq, b->d_pose0, b->nz.init_cam, b->wide ? 1 : 0);
    ASB_CUDA(cudaGetLastError());
    return ASB_OK;
}

// The part of asb_batch_create that can fail once the handle exists (the caller destroys the handle on error).
static int batch_build(asb_batch* b, int kind, int quat_mode, const double pose0[10], const asb_noise* cfg, int n_seq,
                       int max_landmarks, int max_obs_per_frame, int device) {
    b->kind = kind; b->quat_mode = quat_mode; b->device = device;
    b->D = kind == ASB_KIND_EKF ? 3 : 10;
    b->R = kind == ASB_KIND_EKF ? 3 : 7;
    b->ck = kind == ASB_KIND_EKF ? ASB_BATCH_CK : 4;   // rows per chunk R * ck <= 31 (single-warp Cholesky)
    b->nz = to_noise(kind, cfg);
    b->n_seq = n_seq; b->max_lm = max_landmarks; b->max_obs = max_obs_per_frame;
    b->ncap = CAM + b->D * max_landmarks;
    b->state_stride = round_up(b->ncap, 4);
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    b->sm_count = prop.multiProcessorCount;
    b->smem_optin = (int)prop.sharedMemPerBlockOptin;
    b->threads = b->ncap > 150 ? 256 : 128;
    ASB_CUDA(cudaStreamCreateWithFlags(&b->own_stream, cudaStreamNonBlocking));
    b->stream = b->own_stream;
    ASB_CUDA(cudaStreamCreateWithFlags(&b->copy_stream, cudaStreamNonBlocking));
    for (auto& st : b->run_streams) ASB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& e : b->ev_group) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : b->ev_done) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ASB_CUDA(cudaEventCreateWithFlags(&b->ev_main, cudaEventDisableTiming));
    // a group must still fill the GPU a few times over (persistent grid: up to 3 CTAs per SM)
    b->groups = std::max(1, std::min((int)asb_batch::N_GROUPS, n_seq / (3 * b->sm_count)));
    ASB_CUDA(dmalloc(&b->d_states, (size_t)n_seq * b->state_stride));
    b->wide = !resident_fits(b);
    {
        int rc = alloc_covs(b);
        if (rc != ASB_OK) return rc;
    }
    ASB_CUDA(dmalloc(&b->d_nlm, (size_t)n_seq));
    ASB_CUDA(dmalloc(&b->d_prof, 16));
    ASB_CUDA(cudaMemset(b->d_prof, 0, 16 * sizeof(unsigned long long)));
    ASB_CUDA(dmalloc(&b->d_tab, (size_t)n_seq * max_landmarks));
    ASB_CUDA(dmalloc(&b->d_status, (size_t)n_seq));
    ASB_CUDA(dmalloc(&b->d_dense, (size_t)b->ncap * b->ncap));
    ASB_CUDA(cudaMemset(b->d_nlm, 0, (size_t)n_seq * sizeof(int)));
    ASB_CUDA(cudaMemset(b->d_status, 0, (size_t)n_seq * sizeof(int)));
    ASB_CUDA(cudaMemset(b->d_tab, 0, (size_t)n_seq * max_landmarks * sizeof(int)));
    const double pose_default[CAM] = {0, 0, 0, 1, 0, 0, 0, 0, 0, 0};
    ASB_CUDA(dmalloc(&b->d_pose0, (size_t)CAM));
    ASB_CUDA(cudaMemcpy(b->d_pose0, pose0 ? pose0 : pose_default, CAM * sizeof(double), cudaMemcpyHostToDevice));
    {
        int rc = asb_batch_reset(b);
        if (rc != ASB_OK) return rc;
    }
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    return ASB_OK;
}

int asb_batch_create(int kind, int quat_mode, const double pose0[10], const asb_noise* cfg, int n_seq,
                     int max_landmarks, int max_obs_per_frame, int device, asb_batch** out) {
    if (!out) return fail(ASB_ERR_BAD_ARG, "out is NULL");
    *out = nullptr;
    if (kind != ASB_KIND_EKF && kind != ASB_KIND_ROT) return fail(ASB_ERR_BAD_ARG, "kind %d", kind);
    if (quat_mode != ASB_QUAT_AS_WRITTEN && quat_mode != ASB_QUAT_SCALAR_FIRST)
        return fail(ASB_ERR_BAD_ARG, "quat_mode %d", quat_mode);
    if (n_seq < 1 || max_landmarks < 1 || max_obs_per_frame < 1) return fail(ASB_ERR_BAD_ARG, "bad sizes");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(ASB_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(ASB_ERR_BAD_ARG, "device %d of %d", device, ndev);
    ASB_CUDA(cudaSetDevice(device));
    asb_batch* b = new (std::nothrow) asb_batch();
    if (!b) return fail(ASB_ERR_BAD_ARG, "out of host memory");
    const int rc = batch_build(b, kind, quat_mode, pose0, cfg, n_seq, max_landmarks, max_obs_per_frame, device);
    if (rc != ASB_OK) {
        asb_batch_destroy(b);
        cudaGetLastError();
        return rc;
    }
    *out = b;
    return ASB_OK;
}

int asb_batch_destroy(asb_batch* b) {
    ASB_CHECK_BATCH(b);
    cudaStreamSynchronize(b->stream);
    cudaFree(b->d_states); cudaFree(b->d_covs); cudaFree(b->d_nlm); cudaFree(b->d_tab);
    cudaFree(b->d_status); cudaFree(b->d_ids); cudaFree(b->d_poses); cudaFree(b->d_off);
    cudaFree(b->d_dense); cudaFree(b->d_pose0);
    wide_free(b);
    for (auto& st : b->w_streams) if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
    for (auto& e : b->w_done) if (e) cudaEventDestroy(e);
    cudaStreamDestroy(b->own_stream);
    cudaStreamSynchronize(b->copy_stream);
    cudaStreamDestroy(b->copy_stream);
    for (auto& st : b->run_streams) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
    for (auto& e : b->ev_group) cudaEventDestroy(e);
    for (auto& e : b->ev_done) cudaEventDestroy(e);
    cudaEventDestroy(b->ev_main);
    b->magic = 0;
    delete b;
    return ASB_OK;
}

int asb_batch_set_engine(asb_batch* b, int engine) {
    ASB_CHECK_BATCH(b);
    if (engine < 0 || engine > 2) return fail(ASB_ERR_BAD_ARG, "engine %d", engine);
    const bool want_wide = engine == 2 || (engine == 0 && !resident_fits(b));
    if (engine == 1 && !resident_fits(b))
        return fail(ASB_ERR_CAPACITY, "the resident engine needs the covariance in shared memory (N = %d does not fit)", b->ncap);
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    b->engine_req = engine;
    if (want_wide != b->wide) {
        b->wide = want_wide;
        int rc = alloc_covs(b);
        if (rc != ASB_OK) return rc;
    }
    return asb_batch_reset(b);   // switching engines starts the filters over (the covariance layout differs)
}

int asb_batch_get_engine(asb_batch* b, int* engine) {
    ASB_CHECK_BATCH(b);
    if (!engine) return fail(ASB_ERR_BAD_ARG, "engine is NULL");
    *engine = b->wide ? 2 : 1;
    return ASB_OK;
}

int asb_batch_set_stream(asb_batch* b, void* cuda_stream) {
    ASB_CHECK_BATCH(b);
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    b->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : b->own_stream;
    return ASB_OK;
}

int asb_batch_synchronize(asb_batch* b) {
    ASB_CHECK_BATCH(b);
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    return ASB_OK;
}

}  // extern "C"

static int batch_upload(asb_batch* b, const int32_t* ids, const double* poses, const int64_t* offsets,
                        int n_frames, int obs_seqs) {
    ASB_CHECK_BATCH(b);
    if (obs_seqs < 1 || obs_seqs > b->n_seq || b->n_seq % obs_seqs) return fail(ASB_ERR_BAD_ARG, "%d streams do not tile %d sequences", obs_seqs, b->n_seq);
    if (!ids || !poses || !offsets || n_frames < 1) return fail(ASB_ERR_BAD_ARG, "bad stream");
    const long long nrec = (long long)obs_seqs * n_frames;
    const long long total = offsets[nrec];
    if (total > b->n_obs || !b->d_ids) {
        ASB_CUDA(cudaStreamSynchronize(b->stream));
        cudaFree(b->d_ids); cudaFree(b->d_poses);
        b->d_ids = nullptr; b->d_poses = nullptr;
        ASB_CUDA(dmalloc(&b->d_ids, (size_t)total));
        ASB_CUDA(dmalloc(&b->d_poses, (size_t)total * 6));
        b->n_obs = total;
    }
    if (n_frames != b->n_frames || obs_seqs != b->obs_seqs || !b->d_off) {
        ASB_CUDA(cudaStreamSynchronize(b->stream));
        cudaFree(b->d_off);
        b->d_off = nullptr;
        ASB_CUDA(dmalloc(&b->d_off, (size_t)nrec + 1));
        b->n_frames = n_frames;
        b->obs_seqs = obs_seqs;
    }
    static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
    // group g = sequences [n_seq g / G, n_seq (g+1) / G): its ids / poses / offsets are contiguous
    // (the stream is packed sequence-major).  The copies wait for whatever the main stream has
    // queued so far (a run that still reads the old observations).
    ASB_CUDA(cudaEventRecord(b->ev_main, b->stream));
    ASB_CUDA(cudaStreamWaitEvent(b->copy_stream, b->ev_main, 0));
    const int ngr = obs_seqs == b->n_seq ? b->groups : 1;   // tiled streams: one copy (every group reads all of it)
    b->up_groups = ngr;
    for (int g = 0; g < ngr; ++g) {
        const long long s0 = (long long)obs_seqs * g / ngr, s1 = (long long)obs_seqs * (g + 1) / ngr;
        const long long r0 = s0 * n_frames, r1 = s1 * n_frames;
        const long long o0 = offsets[r0], o1 = offsets[r1];
        const long long nrecs = r1 - r0 + (g == ngr - 1 ? 1 : 0);
        ASB_CUDA(cudaMemcpyAsync(b->d_off + r0, offsets + r0, (size_t)nrecs * sizeof(int64_t), cudaMemcpyHostToDevice, b->copy_stream));
        if (g < ngr - 1)   // a kernel reads offsets[r1] as the end of its last record
            ASB_CUDA(cudaMemcpyAsync(b->d_off + r1, offsets + r1, sizeof(int64_t), cudaMemcpyHostToDevice, b->copy_stream));
        ASB_CUDA(cudaMemcpyAsync(b->d_ids + o0, ids + o0, (size_t)(o1 - o0) * sizeof(int), cudaMemcpyHostToDevice, b->copy_stream));
        ASB_CUDA(cudaMemcpyAsync(b->d_poses + 6 * o0, poses + 6 * o0, (size_t)(o1 - o0) * 6 * sizeof(double), cudaMemcpyHostToDevice, b->copy_stream));
        ASB_CUDA(cudaEventRecord(b->ev_group[g], b->copy_stream));
    }
    b->fresh_upload = true;
    {   // rounds per frame = ceil(max_s k(s, t) / WCK): the wide engine's frame-synchronous launches need it on the
        // host (kept for the resident engine too: asb_batch_set_engine may switch after the upload)
        b->kmax.assign(n_frames, 0);
        for (long long q = 0; q < obs_seqs; ++q) {
            const int64_t* o = offsets + q * n_frames;
            for (int t = 0; t < n_frames; ++t) b->kmax[t] = std::max(b->kmax[t], (int)(o[t + 1] - o[t]));
        }
        for (int t = 0; t < n_frames; ++t) b->kmax[t] = std::min(b->kmax[t], b->max_obs);
    }
    return ASB_OK;
}

int asb_batch_upload(asb_batch* b, const int32_t* ids, const double* poses, const int64_t* offsets,
                     int n_frames) {
    int rc = check(b);
    if (rc != ASB_OK) return rc;
    return batch_upload(b, ids, poses, offsets, n_frames, b->n_seq);
}

int asb_batch_upload_tiled(asb_batch* b, const int32_t* ids, const double* poses, const int64_t* offsets,
                           int n_frames, int n_streams) {
    return batch_upload(b, ids, poses, offsets, n_frames, n_streams);
}

extern "C" {

int asb_batch_run(asb_batch* b, int t0, int t1) {
    ASB_CHECK_BATCH(b);
    if (!b->d_off) return fail(ASB_ERR_NO_STREAM, "batch_run before batch_upload");
    if (t0 < 0 || t1 < t0 || t1 > b->n_frames) return fail(ASB_ERR_BAD_ARG, "bad frame range [%d,%d)", t0, t1);
    if (b->wide) {
        if ((int)b->kmax.size() != b->n_frames) return fail(ASB_ERR_NO_STREAM, "batch_run before batch_upload");
        if (b->fresh_upload) {
            for (int g = 0; g < b->up_groups; ++g) ASB_CUDA(cudaStreamWaitEvent(b->stream, b->ev_group[g], 0));
            b->fresh_upload = false;
        }
        return b->kind == ASB_KIND_EKF ? wide_run<0>(b, t0, t1) : wide_run<1>(b, t0, t1);
    }
    BatchParams p;
    p.states = b->d_states; p.covs = b->d_covs; p.nlm = b->d_nlm; p.lm_ids = b->d_tab; p.status = b->d_status;
    p.ids = b->d_ids; p.poses = b->d_poses; p.offsets = b->d_off;
    p.state_stride = b->state_stride; p.cov_stride = b->cov_stride;
    p.n_seq = b->n_seq; p.n_frames = b->n_frames; p.t0 = t0; p.t1 = t1;
    p.max_lm = b->max_lm; p.max_obs = b->max_obs; p.obs_seqs = b->obs_seqs;
    p.as_written = (b->kind == ASB_KIND_EKF && b->quat_mode == ASB_QUAT_AS_WRITTEN) ? 1 : 0;
    p.nz = b->nz;
    p.prof = b->d_prof;
    const size_t base = batch_smem_base(b->D, b->R, b->ck, b->max_lm, b->max_obs);
    const size_t pbytes = (size_t)tiled_doubles(b->ncap) * sizeof(double);   // tile-blocked copy (batch_kernels.cuh)
    const size_t smem = base + pbytes;
    if (smem > (size_t)b->smem_optin)
        return fail(ASB_ERR_CAPACITY, "batched filter needs %zu B shared memory > %d", smem, b->smem_optin);
    const int per_sm = std::max(1, std::min(8, (int)((size_t)(b->smem_optin + 1024) / (smem + 1024))));
    if (!b->fresh_upload || b->up_groups == 1) {
        // observations already resident: one persistent launch over all sequences
        if (b->fresh_upload) ASB_CUDA(cudaStreamWaitEvent(b->stream, b->ev_group[0], 0));
        b->fresh_upload = false;
        p.seq0 = 0; p.seq1 = b->n_seq;
        return launch_batch_dyn(b, p, smem, std::min(b->n_seq, b->sm_count * per_sm), b->stream);
    }
    // straight after an upload: one launch per sequence group, each gated on that group's copy
    b->fresh_upload = false;
    ASB_CUDA(cudaEventRecord(b->ev_main, b->stream));
    for (auto& st : b->run_streams) ASB_CUDA(cudaStreamWaitEvent(st, b->ev_main, 0));
    for (int g = 0; g < b->up_groups; ++g) {
        cudaStream_t st = b->run_streams[g & 1];
        ASB_CUDA(cudaStreamWaitEvent(st, b->ev_group[g], 0));
        p.seq0 = (int)((long long)b->n_seq * g / b->up_groups);
        p.seq1 = (int)((long long)b->n_seq * (g + 1) / b->up_groups);
        int rc = launch_batch_dyn(b, p, smem, std::min(p.seq1 - p.seq0, b->sm_count * per_sm), st);
        if (rc != ASB_OK) return rc;
    }
    for (int i = 0; i < 2; ++i) {
        ASB_CUDA(cudaEventRecord(b->ev_done[i], b->run_streams[i]));
        ASB_CUDA(cudaStreamWaitEvent(b->stream, b->ev_done[i], 0));
    }
    return ASB_OK;
}

int asb_batch_profile(asb_batch* b, double clocks[10]) {
    ASB_CHECK_BATCH(b);
    if (!clocks) return fail(ASB_ERR_BAD_ARG, "clocks is NULL");
    unsigned long long t[16];
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    b->profiling = true;
    ASB_CUDA(cudaMemcpy(t, b->d_prof, sizeof t, cudaMemcpyDeviceToHost));
    ASB_CUDA(cudaMemset(b->d_prof, 0, sizeof t));
    for (int i = 0; i < 10; ++i) clocks[i] = (double)t[i];
    return ASB_OK;
}

int asb_batch_get_dims(asb_batch* b, int seq, int* n_landmarks, int* n_state) {
    ASB_CHECK_BATCH(b);
    if (seq < 0 || seq >= b->n_seq) return fail(ASB_ERR_BAD_ARG, "seq %d", seq);
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    int nlm = 0;
    ASB_CUDA(cudaMemcpy(&nlm, b->d_nlm + seq, sizeof(int), cudaMemcpyDeviceToHost));
    if (n_landmarks) *n_landmarks = nlm;
    if (n_state) *n_state = CAM + b->D * nlm;
    return ASB_OK;
}

static int seq_status(asb_batch* b, int seq) {
    int st = 0;
    ASB_CUDA(cudaMemcpy(&st, b->d_status + seq, sizeof(int), cudaMemcpyDeviceToHost));
    if (st != 0) return fail(st, "sequence %d: %s", seq, st == ASB_ERR_NOT_SPD ? "S not positive definite" : "capacity exceeded");
    return ASB_OK;
}

int asb_batch_get_state(asb_batch* b, int seq, double* out, int n_state) {
    ASB_CHECK_BATCH(b);
    int nlm, n;
    int rc = asb_batch_get_dims(b, seq, &nlm, &n);
    if (rc != ASB_OK) return rc;
    if (!out || n_state != n) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, n);
    ASB_CUDA(cudaMemcpy(out, b->d_states + (size_t)seq * b->state_stride, n * sizeof(double), cudaMemcpyDeviceToHost));
    return seq_status(b, seq);
}

int asb_batch_get_cov(asb_batch* b, int seq, double* out, int n_state) {
    ASB_CHECK_BATCH(b);
    int nlm, n;
    int rc = asb_batch_get_dims(b, seq, &nlm, &n);
    if (rc != ASB_OK) return rc;
    if (!out || n_state != n) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, n);
    const long long cnt = (long long)n * n;
    if (b->wide)
        k_unpack_tiled<<<(unsigned)((cnt + 255) / 256), 256, 0, b->stream>>>(b->d_covs + (size_t)seq * b->cov_stride, n, b->d_dense);
    else
        k_unpack<<<(unsigned)((cnt + 255) / 256), 256, 0, b->stream>>>(b->d_covs + (size_t)seq * b->cov_stride, n, b->d_dense);
    ASB_CUDA(cudaGetLastError());
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    ASB_CUDA(cudaMemcpy(out, b->d_dense, cnt * sizeof(double), cudaMemcpyDeviceToHost));
    return seq_status(b, seq);
}

// ---- checkpoint / re-synchronisation setters (SURVEY 8c ii: the as-written EKF is compared one step at a time
//      with the prior state copied from the oracle; the headline workload runs that mode) --------------------
int asb_batch_set_landmarks(asb_batch* b, int seq, const int32_t* ids_in_index_order, int n) {
    ASB_CHECK_BATCH(b);
    if (seq < 0 || seq >= b->n_seq) return fail(ASB_ERR_BAD_ARG, "seq %d", seq);
    if (n < 0 || n > b->max_lm || (n > 0 && !ids_in_index_order)) return fail(ASB_ERR_CAPACITY, "%d landmarks > max_landmarks %d", n, b->max_lm);
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    if (n) ASB_CUDA(cudaMemcpy(b->d_tab + (size_t)seq * b->max_lm, ids_in_index_order, (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
    ASB_CUDA(cudaMemcpy(b->d_nlm + seq, &n, sizeof(int), cudaMemcpyHostToDevice));
    const int zero = 0;
    ASB_CUDA(cudaMemcpy(b->d_status + seq, &zero, sizeof(int), cudaMemcpyHostToDevice));
    return ASB_OK;
}

int asb_batch_set_state(asb_batch* b, int seq, const double* in, int n_state) {
    ASB_CHECK_BATCH(b);
    int nlm, n;
    int rc = asb_batch_get_dims(b, seq, &nlm, &n);
    if (rc != ASB_OK) return rc;
    if (!in || n_state != n) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, n);
    ASB_CUDA(cudaMemcpy(b->d_states + (size_t)seq * b->state_stride, in, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
    return ASB_OK;
}

int asb_batch_set_cov(asb_batch* b, int seq, const double* in, int n_state) {
    ASB_CHECK_BATCH(b);
    int nlm, n;
    int rc = asb_batch_get_dims(b, seq, &nlm, &n);
    if (rc != ASB_OK) return rc;
    if (!in || n_state != n) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, n);
    // dense row-major n x n (the lower triangle is taken) -> the engine's layout: packed lower triangle
    // (resident) or 8 x 8 tiles with full diagonal tiles (wide); entries outside the live block stay zero
    std::vector<double> buf((size_t)b->cov_stride, 0.0);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) {
            const double v = in[(size_t)i * n + j];
            if (b->wide) {
                buf[((size_t)(tri(i >> 3) + (j >> 3)) << 6) + ((i & 7) << 3) + (j & 7)] = v;
                if ((i >> 3) == (j >> 3)) buf[((size_t)(tri(i >> 3) + (j >> 3)) << 6) + ((j & 7) << 3) + (i & 7)] = v;
            } else {
                buf[(size_t)tri(i) + j] = v;
            }
        }
    ASB_CUDA(cudaMemcpy(b->d_covs + (size_t)seq * b->cov_stride, buf.data(), buf.size() * sizeof(double), cudaMemcpyHostToDevice));
    return ASB_OK;
}

int asb_batch_get_landmark_ids(asb_batch* b, int seq, int32_t* ids, int capacity) {
    ASB_CHECK_BATCH(b);
    int nlm, n;
    int rc = asb_batch_get_dims(b, seq, &nlm, &n);
    if (rc != ASB_OK) return rc;
    if (!ids || capacity < nlm) return fail(ASB_ERR_BAD_ARG, "capacity %d < %d", capacity, nlm);
    ASB_CUDA(cudaMemcpy(ids, b->d_tab + (size_t)seq * b->max_lm, nlm * sizeof(int), cudaMemcpyDeviceToHost));
    return ASB_OK;
}

int asb_batch_get_cameras(asb_batch* b, double* out, int32_t* status) {
    ASB_CHECK_BATCH(b);
    if (!out) return fail(ASB_ERR_BAD_ARG, "out is NULL");
    ASB_CUDA(cudaStreamSynchronize(b->stream));
    ASB_CUDA(cudaMemcpy2D(out, CAM * sizeof(double), b->d_states, (size_t)b->state_stride * sizeof(double),
                          CAM * sizeof(double), b->n_seq, cudaMemcpyDeviceToHost));
    if (status) ASB_CUDA(cudaMemcpy(status, b->d_status, (size_t)b->n_seq * sizeof(int), cudaMemcpyDeviceToHost));
    return ASB_OK;
}

int asb_batch_device_views(asb_batch* b, void** d_states, void** d_covs, int* state_stride, int* cov_ld) {
    ASB_CHECK_BATCH(b);
    if (d_states) *d_states = b->d_states;
    if (d_covs) *d_covs = b->d_covs;
    if (state_stride) *state_stride = (int)b->state_stride;
    if (cov_ld) *cov_ld = (int)b->cov_stride;  // packed lower triangle: doubles per sequence
    return ASB_OK;
}

}  // extern "C"
