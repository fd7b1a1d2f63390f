// Row-sharded single filter: ONE very large map (BASELINE config C3) spread over the GPUs of a
// box, one process per GPU (north_star: "the covariance rows are sharded, with the
// small PH^T / S panels exchanged by NCCL over NVLink").  C ABI: "row-sharded filter" section of
// include/aruco_slam_b200.h.
//
// Rank g owns the full rows [row0, row0 + rows) of P (rows = a whole number of 64-column panel
// blocks, the same on every rank), stored as a rows x ldp block; the state vector, the id map and
// the observations are replicated.  Because P is symmetric, everything a rank needs of the rows
// T (the touched state indices) for ITS columns is in its own rows: P[T, c] = P[c, T] for c in
// its row range.  Per update chunk (<= kc_max markers), on every rank:
//
//   k_sh_ptt     own rows of P[T,T] (others zero)                      ---- ncclAllReduce (nT x nT, <= 64 KB)
//   k_f1         S = H P[T,T] H^T + R, Cholesky, y = L^-1 nu, A = L^-1 H_T   (pipe.cuh, redundantly per rank)
//   k_sh_f2      V[:, own] = A P[own, T]^T, dx[own] = V^T y, own panel blocks
//                                                                      ---- ncclAllGather (V_J panel 8 m4 N B, dx)
//   k_sh_apply   state update + quaternion reset with the gathered dx (replicated)
//   k_sweep      rect mode: P[own, :] -= V[:, own]^T V (+ diag Q)        (sweep.cuh)
//
// All collectives are on the filter's stream between the kernels (no host synchronisation).  When
// the next chunk is known (stream mode, or a frame of more than kc_max markers) its front is taken
// off the critical path: see shard_chunks() below.  The rectangular sweep cannot use the symmetry
// (each rank computes full rows: 2 N^2 m / G flops per rank), so at G = 2 a rank does as much FP64
// work as the single-GPU triangle sweep; DESIGN.md §6 has the measurements.
#include <dlfcn.h>
#include <nccl.h>   // types and prototypes only: libnccl.so.2 is opened at run time (the one torch already loaded)

#include <algorithm>
#include <cstdlib>
#include <new>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "front.cuh"
#include "pipe.cuh"

using namespace asb;

namespace {
constexpr uint32_t SHARD_MAGIC = 0x41534253;  // "ASBS"
constexpr int SH_SLOTS = 8;

struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    bool ok = false;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.ok ? &api : nullptr;
    tried = true;
    // The copy that is already in the process wins (from Python that is torch's bundled NCCL: two
    // different libnccl.so.2 in one process break whichever is resolved second); ASB_NCCL_LIB overrides.
    if (const char* path = std::getenv("ASB_NCCL_LIB")) api.handle = dlopen(path, RTLD_NOW);
    if (!api.handle) api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!api.handle) api.handle = dlopen("libnccl.so.2", RTLD_NOW);
    if (!api.handle) return nullptr;
#define ASB_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(api.handle, "nccl" #name))
    ASB_SYM(GetUniqueId); ASB_SYM(CommInitRank); ASB_SYM(CommDestroy); ASB_SYM(AllReduce); ASB_SYM(AllGather);
    ASB_SYM(GroupStart); ASB_SYM(GroupEnd); ASB_SYM(GetErrorString);
#undef ASB_SYM
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather &&
             api.GroupStart && api.GroupEnd && api.GetErrorString;
    return api.ok ? &api : nullptr;
}

#define ASB_NCCL(expr)                                                                            \
    do {                                                                                          \
        ncclResult_t _r = (expr);                                                                 \
        if (_r != ncclSuccess)                                                                    \
            return ::asb::fail(ASB_ERR_CUDA, "%s:%d %s -> NCCL: %s", __FILE__, __LINE__, #expr,   \
                               nccl_api()->GetErrorString(_r));                                   \
    } while (0)

struct ShFrame {
    int64_t off;
    int k, n0, n1;
    std::vector<int32_t> new_ids;
};

// The partition: panel blocks of 64 state columns, the same number on every rank.
struct Partition { int blocks_per_rank, rows, row0, ldp; };
Partition make_partition(int kind, int max_landmarks, int world, int rank) {
    const int d = kind == ASB_KIND_EKF ? 3 : 10;
    const int ncap = CAM + d * max_landmarks;
    const int nblk = (ncap + V_BLK - 1) / V_BLK;
    Partition p;
    p.blocks_per_rank = (nblk + world - 1) / world;
    p.rows = p.blocks_per_rank * V_BLK;
    p.row0 = rank * p.rows;
    p.ldp = p.rows * world;
    return p;
}
}  // namespace

struct asb_shard {
    uint32_t magic = SHARD_MAGIC;
    int kind = 0, quat_mode = 0, device = 0, rank = 0, world = 1;
    int D = 3, R = 3, W = 13;
    Noise nz{};
    int cap_lm = 0, ncap = 0, max_obs = 0, kc_max = 0;
    Partition part{};
    int nlm = 0;
    std::unordered_map<int32_t, int> map;
    std::vector<int32_t> id_order;
    ncclComm_t comm = nullptr;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    double *d_state = nullptr, *d_P = nullptr, *d_PTpart = nullptr, *d_PT = nullptr, *d_At = nullptr, *d_y = nullptr;
    double *d_Vr = nullptr, *d_Vc = nullptr, *d_dx[2] = {nullptr, nullptr}, *d_blocks = nullptr, *d_nu = nullptr;
    int *d_T = nullptr, *d_status = nullptr;
    int* h_idx = nullptr; double* h_pose = nullptr; unsigned char* h_new = nullptr;
    int* d_idx = nullptr; double* d_pose = nullptr; unsigned char* d_new = nullptr;
    cudaEvent_t slot_ev[SH_SLOTS]{};
    int slot = 0;
    int* s_idx = nullptr; double* s_pose = nullptr; unsigned char* s_new = nullptr;
    std::vector<ShFrame> frames;
    int next_frame = 0;
    int sm_count = 0, smem_optin = 0;
    unsigned int* d_tail = nullptr;   // two tile counters used alternately by successive sweeps
    unsigned int sweep_seq = 0;
    int tail_tiles = 2;
    int cur = 0;        // which of d_dx[2] holds the latest all-gathered dx
    int overlap = 1;    // look-ahead pipeline: the front of chunk i + 1 overlaps the sweep of chunk i
    int n_state() const { return CAM + D * nlm; }
};

namespace asb {

// ---- own rows of P[T,T] ------------------------------------------------------------------
struct ShPttArgs {
    const double* P;   // rows x ldp
    int ldp, row0, rows, n0;
    const int* idx;
    int kc;
    double* out;       // nT x nT: P[T[a], T[b]] where this rank owns row T[a] and both are < n0, else 0
};

template <int KIND>
__global__ void __launch_bounds__(128) k_sh_ptt(ShPttArgs p) {
    constexpr int D = Dims<KIND>::D;
    const int nT = CAM + D * p.kc;
    auto touched = [&](int a) { return a < CAM ? a : CAM + D * p.idx[(a - CAM) / D] + (a - CAM) % D; };
    const int a = blockIdx.x;
    const int i = touched(a);
    const bool mine = i >= p.row0 && i < p.row0 + p.rows && i < p.n0;
    for (int b = threadIdx.x; b < nT; b += blockDim.x) {
        const int j = touched(b);
        p.out[(size_t)a * nT + b] = (mine && j < p.n0) ? p.P[(size_t)(i - p.row0) * p.ldp + j] : 0.0;
    }
}

// ---- V[:, own columns] = A P[own, T]^T ------------------------------------------------------
struct ShF2Args {
    double* P;                   // rows x ldp (growth stores on the first chunk)
    int ldp, row0, rows;
    int n0, n1;
    Noise nz;
    int kc, m4;
    int first_chunk, add_q, use_delta;
    const double* At;            // (nT, m4) from k_f1
    const double* y;             // (m4)
    const int* T;                // (nT)
    const double* dx_in;         // full length (ldp)
    double* dx_out;              // full length; this rank stores [row0, row0 + rows)
    double* Vr;                  // LOCAL panel blocks (V_I order): block b = local rows [64 b, 64 b + 64)
    double* Vc;                  // GLOBAL panel (V_J order): this rank stores blocks row0 / 64 ...
};

template <int KIND>
struct ShF2Smem {
    static constexpr int D = Dims<KIND>::D;
    static size_t bytes(int kc) {
        const size_t nT = CAM + D * kc, kp = (nT + 3) & ~(size_t)3;
        size_t dbl = kp * F2_LDB + kp * F2_LDA + (size_t)64 * F2_LDV + 64 + 8 * F2_COLS;
        return dbl * sizeof(double) + (nT + 4) * sizeof(int);
    }
};

template <int KIND>
__global__ void __launch_bounds__(F2_THREADS, 2)
k_sh_f2(ShF2Args p) {
    constexpr int D = Dims<KIND>::D;
    constexpr int NT = F2_THREADS;
    extern __shared__ __align__(16) unsigned char fsm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kc = p.kc, m4 = p.m4, nT = CAM + D * kc, kp = (nT + 3) & ~3;
    const int n0 = p.n0, n1 = p.n1;
    const int l0 = blockIdx.x * F2_COLS;          // first local row (= column of V) of this CTA
    const int j0 = p.row0 + l0;                   // ... as a global state index
    const int lblock = blockIdx.x >> 1, half = blockIdx.x & 1;

    double* s_stage = reinterpret_cast<double*>(fsm);        // kp x F2_LDB: P_eff[T, own columns]
    double* s_At = s_stage + (size_t)kp * F2_LDB;            // kp x F2_LDA
    double* s_V = s_At + (size_t)kp * F2_LDA;                // 64 x F2_LDV
    double* s_y = s_V + (size_t)64 * F2_LDV;                 // 64
    double* s_red = s_y + 64;                                // 8 x F2_COLS
    int* s_T = reinterpret_cast<int*>(s_red + 8 * F2_COLS);  // nT

    pdl_wait();   // launched as a programmatic dependent of the previous chunk's sweep (which triggers at its end)
    for (int a = tid; a < nT; a += NT) s_T[a] = p.T[a];
    for (int a = tid; a < 64; a += NT) s_y[a] = a < m4 ? p.y[a] : 0.0;
    {
        const int pairs = m4 >> 1;
        for (int e = tid; e < nT * pairs; e += NT) {
            const int tb = e / pairs, a2 = e - tb * pairs;
            cp_async16(s_At + (size_t)tb * F2_LDA + 2 * a2, p.At + (size_t)tb * m4 + 2 * a2);
        }
        const int padc = 64 - m4;
        for (int e = tid; e < nT * padc; e += NT) {
            const int tb = e / padc, a = m4 + (e - tb * padc);
            s_At[(size_t)tb * F2_LDA + a] = 0.0;
        }
        for (int e = tid; e < (kp - nT) * F2_LDA; e += NT) s_At[(size_t)nT * F2_LDA + e] = 0.0;
        for (int e = tid; e < (kp - nT) * F2_LDB; e += NT) s_stage[(size_t)nT * F2_LDB + e] = 0.0;
    }
    __syncthreads();
    // stage[a][cc] = P_eff[T[a]][j0 + cc] = P[own row j0 + cc][T[a]] (symmetry); >= n0: this frame's constants
    for (int e = tid; e < nT * F2_COLS; e += NT) {
        const int a = e / F2_COLS, cc = e - a * F2_COLS;
        const int i = s_T[a], c = j0 + cc;
        double v;
        if (c < n0 && i < n0) v = p.P[(size_t)(l0 + cc) * p.ldp + i];
        else v = (c == i && c < n1) ? p.nz.init_lm : 0.0;
        if (p.add_q && c == i) v += process_noise(p.nz, c);
        s_stage[(size_t)a * F2_LDB + cc] = v;
    }
    cp_async_wait_all();
    __syncthreads();

    {   // same DMMA product as k_f2 (pipe.cuh)
        const int g = lane >> 2, t4 = lane & 3;
        const int a0 = 8 * warp;
        if (a0 < m4) {
            double acc[4][2];
#pragma unroll
            for (int ct = 0; ct < 4; ++ct) acc[ct][0] = acc[ct][1] = 0.0;
            const double* ap = s_At + (size_t)t4 * F2_LDA + a0 + g;
            const double* bp = s_stage + (size_t)t4 * F2_LDB + g;
#pragma unroll 2
            for (int kk = 0; kk < kp; kk += 4) {
                const double av = ap[(size_t)kk * F2_LDA];
                double bv[4];
#pragma unroll
                for (int ct = 0; ct < 4; ++ct) bv[ct] = bp[(size_t)kk * F2_LDB + 8 * ct];
#pragma unroll
                for (int ct = 0; ct < 4; ++ct) dmma884(acc[ct][0], acc[ct][1], av, bv[ct]);
            }
#pragma unroll
            for (int ct = 0; ct < 4; ++ct)
                *reinterpret_cast<double2*>(s_V + (size_t)(a0 + g) * F2_LDV + 8 * ct + 2 * t4) = make_double2(acc[ct][0], acc[ct][1]);
        }
    }
    __syncthreads();

    {   // dx[own] (+)= V^T y; entries >= n1 are defined (zero): the whole range is all-gathered
        double part = 0.0;
        for (int a = warp; a < m4; a += NT / 32) part += s_V[(size_t)a * F2_LDV + lane] * s_y[a];
        s_red[warp * F2_COLS + lane] = part;
    }
    __syncthreads();
    if (tid < F2_COLS) {
        double dsum = 0.0;
#pragma unroll
        for (int w = 0; w < NT / 32; ++w) dsum += s_red[w * F2_COLS + tid];
        const int j = j0 + tid;
        p.dx_out[j] = j < n1 ? (p.use_delta ? p.dx_in[j] : 0.0) + dsum : 0.0;
    }
    {
        double* vr = p.Vr + (size_t)lblock * m4 * V_LDB;
        double* vc = p.Vc + (size_t)(p.row0 / V_BLK + lblock) * m4 * V_LDB;
        for (int e = tid; e < m4 * F2_COLS; e += NT) {
            const int a = e / F2_COLS, pos = F2_COLS * half + (e - a * F2_COLS);
            const int cr = (pos & ~31) + 4 * (pos & 7) + ((pos & 31) >> 3);
            const int cc = (pos & ~15) + 4 * ((pos & 7) >> 1) + 2 * ((pos & 15) >> 3) + (pos & 1);
            vr[a * V_LDB + pos] = s_V[(size_t)a * F2_LDV + (cr - F2_COLS * half)];
            vc[a * V_LDB + pos] = s_V[(size_t)a * F2_LDV + (cc - F2_COLS * half)];
        }
    }
    // covariance growth (EKF:279-290) for this CTA's rows: new rows get init_lm on the diagonal, old
    // rows zeros in the new columns
    if (p.first_chunk && n1 > n0) {
        for (int r = 0; r < F2_COLS; ++r) {
            const int gi = j0 + r;
            if (gi >= n1) break;
            double* prow = p.P + (size_t)(l0 + r) * p.ldp;
            if (gi >= n0) {
                for (int j = tid; j < n1; j += NT) prow[j] = (j == gi) ? p.nz.init_lm : 0.0;
            } else {
                for (int j = n0 + tid; j < n1; j += NT) prow[j] = 0.0;
            }
        }
    }
}

}  // namespace asb

static int check(asb_shard* s) {
    if (!s || s->magic != SHARD_MAGIC) return fail(ASB_ERR_BAD_HANDLE, "bad asb_shard handle");
    return ASB_OK;
}
#define ASB_CHECK_SHARD(s)              \
    do {                                \
        int _rc = check(s);             \
        if (_rc != ASB_OK) return _rc;  \
        cudaSetDevice((s)->device);     \
    } while (0)

template <typename T>
static cudaError_t dmalloc(T** p, size_t count) {
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(p), std::max<size_t>(count, 1) * sizeof(T));
    if (e == cudaSuccess) e = cudaMemset(*p, 0, std::max<size_t>(count, 1) * sizeof(T));
    return e;
}

// One update chunk (<= kc_max markers of one frame) as the launch sequence sees it.
struct ShChunk {
    const int* idx;
    const double* pose;
    const unsigned char* isnew;
    int kc, n0, n1;      // n0: state dims before this chunk's augmentation (= n1 after the first chunk of a frame)
    bool first, last;    // first / last chunk of its frame
};

static void append_chunks(const asb_shard* s, std::vector<ShChunk>& out, const int* idx, const double* pose,
                          const unsigned char* isnew, int k, int n0, int n1) {
    for (int c0 = 0; c0 < k; c0 += s->kc_max) {
        ShChunk c;
        c.kc = std::min(s->kc_max, k - c0);
        c.idx = idx + c0; c.pose = pose + 6 * (size_t)c0; c.isnew = isnew ? isnew + c0 : nullptr;
        c.first = c0 == 0; c.last = c0 + c.kc >= k;
        c.n0 = c.first ? n0 : n1; c.n1 = n1;
        out.push_back(c);
    }
}

namespace asb {
// state update + quaternion reset (EKF:134-152 / ROT:143-177) with the all-gathered dx, replicated on every rank
__global__ void __launch_bounds__(256) k_sh_apply(double* state, const double* dx, int kind, int as_written, int nlm) {
    apply_state(state, dx, kind, as_written, nlm, threadIdx.x, blockDim.x);
}
}  // namespace asb

// The launch sequence for a run of chunks.  With a NEXT chunk known (stream mode, or a frame of
// more than kc_max markers) the front of chunk i+1 is taken off the critical path exactly as in the
// single-GPU engine (pipe.cuh):
//
//   k_sh_f2(i) -> k_sh_ptt(i+1) -> NCCL group { all-gather V(i), all-gather dx, all-reduce P_in[T(i+1),T(i+1)] }
//     -> k_sh_apply -> k_f1(i+1) -> k_sweep(i) -> k_sh_f2(i+1) ...
//
// k_sh_ptt(i+1) reads the covariance BEFORE sweep i (its all-reduce rides in the same NCCL group
// as the panel all-gather, so the overlap costs no extra collective launch); k_f1(i+1) brings that
// block up to date with the panel of chunk i (lazy operand: P[T,T] + Q - V^T V, the panel columns
// taken from the all-gathered tiled panel) and runs WHILE k_sweep(i) streams the row block: it owns
// an SM (nearly all of its shared memory) and the sweep is its programmatic dependent launch,
// waiting for it only at its end.  The first chunk of a run has no look-ahead: k_sh_ptt, a separate
// all-reduce and k_f1 run serially in front of it.
template <int KIND>
static int shard_chunks(asb_shard* s, const std::vector<ShChunk>& chunks) {
    constexpr int R = Dims<KIND>::R, W = CAM + Dims<KIND>::D, D = Dims<KIND>::D;
    NcclApi* nc = nccl_api();
    cudaStream_t st = s->stream;
    const Partition& pt = s->part;
    const int as_written = (KIND == 0 && s->quat_mode == ASB_QUAT_AS_WRITTEN) ? 1 : 0;
    const size_t f1_big = (size_t)s->smem_optin - 2048;
    bool have_front = false;   // k_f1 of the current chunk has already been launched (overlapped with the previous sweep)
    long long obs_row = 0;     // rows of d_blocks / d_nu used so far in this frame

    auto launch_ptt = [&](const ShChunk& c) {
        ShPttArgs g;
        g.P = s->d_P; g.ldp = pt.ldp; g.row0 = pt.row0; g.rows = pt.rows; g.n0 = c.n0;
        g.idx = c.idx; g.kc = c.kc; g.out = s->d_PTpart;
        k_sh_ptt<KIND><<<CAM + D * c.kc, 128, 0, st>>>(g);
    };
    auto launch_f1 = [&](const ShChunk& c, const ShChunk* prev, long long row) {
        const int m4 = round_up(R * c.kc, 4);
        F1Args a{};
        a.state = s->d_state; a.P = nullptr; a.PT_dense = s->d_PT; a.ld = pt.ldp;
        a.n0 = c.n0; a.n1 = c.n1; a.nz = s->nz;
        a.idx = c.idx; a.poses = c.pose; a.isnew = c.isnew;
        a.kc = c.kc; a.m4 = m4;
        a.first_chunk = c.first ? 1 : 0;
        a.add_q = c.first ? 1 : 0;
        a.use_delta = c.first ? 0 : 1;
        a.dx_in = s->d_dx[s->cur];
        a.lazy = prev ? 1 : 0; a.Vp = nullptr;
        a.Vc_tiled = prev ? s->d_Vc : nullptr;
        a.m4p = prev ? round_up(R * prev->kc, 4) : 0;
        a.lazy_q = (prev && prev->first) ? 1 : 0;
        a.At = s->d_At; a.y = s->d_y; a.T = s->d_T;
        a.blocks_out = s->d_blocks + (size_t)row * R * W; a.nu_out = s->d_nu + (size_t)R * row;
        a.status = s->d_status;
        a.dbg = 0; a.prof = nullptr;
        // overlapped with a sweep: ask for the whole SM so that no sweep CTA shares its FP64 pipe
        k_f1<KIND, true><<<1, FRONT_THREADS, prev ? f1_big : F1Smem<KIND>::bytes(c.kc, m4), st>>>(a);
    };

    for (size_t i = 0; i < chunks.size(); ++i) {
        const ShChunk& c = chunks[i];
        const ShChunk* next = (i + 1 < chunks.size() && s->overlap) ? &chunks[i + 1] : nullptr;
        const int m4 = round_up(R * c.kc, 4);
        if (c.first) obs_row = 0;
        if (!have_front) {
            const int nT = CAM + D * c.kc;
            launch_ptt(c);
            ASB_NCCL(nc->AllReduce(s->d_PTpart, s->d_PT, (size_t)nT * nT, ncclDouble, ncclSum, s->comm, st));
            launch_f1(c, nullptr, obs_row);
        }
        obs_row += c.kc;
        ShF2Args b;
        b.P = s->d_P; b.ldp = pt.ldp; b.row0 = pt.row0; b.rows = pt.rows;
        b.n0 = c.n0; b.n1 = c.n1; b.nz = s->nz; b.kc = c.kc; b.m4 = m4;
        b.first_chunk = c.first ? 1 : 0; b.add_q = c.first ? 1 : 0; b.use_delta = c.first ? 0 : 1;
        b.At = s->d_At; b.y = s->d_y; b.T = s->d_T;
        b.dx_in = s->d_dx[s->cur]; b.dx_out = s->d_dx[s->cur ^ 1];
        b.Vr = s->d_Vr; b.Vc = s->d_Vc;
        {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)(pt.rows / F2_COLS));
            cfg.blockDim = dim3(F2_THREADS);
            cfg.dynamicSmemBytes = ShF2Smem<KIND>::bytes(c.kc);
            cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at;
            cfg.numAttrs = have_front ? 1 : 0;   // the kernel before it in the stream is the previous chunk's sweep
            cudaLaunchKernelEx(&cfg, k_sh_f2<KIND>, b);
        }
        s->cur ^= 1;
        if (next) launch_ptt(*next);   // the covariance as it is BEFORE this chunk's sweep (after its growth stores)
        // the exchange step: every rank's V_J panel blocks, its slice of dx, and the next chunk's P[T,T]
        const size_t vcount = (size_t)pt.blocks_per_rank * m4 * V_LDB;
        ASB_NCCL(nc->GroupStart());
        ASB_NCCL(nc->AllGather(s->d_Vc + (size_t)s->rank * vcount, s->d_Vc, vcount, ncclDouble, s->comm, st));
        ASB_NCCL(nc->AllGather(s->d_dx[s->cur] + pt.row0, s->d_dx[s->cur], (size_t)pt.rows, ncclDouble, s->comm, st));
        if (next) {
            const int nTn = CAM + D * next->kc;
            ASB_NCCL(nc->AllReduce(s->d_PTpart, s->d_PT, (size_t)nTn * nTn, ncclDouble, ncclSum, s->comm, st));
        }
        ASB_NCCL(nc->GroupEnd());
        if (c.last) k_sh_apply<<<1, 256, 0, st>>>(s->d_state, s->d_dx[s->cur], KIND, as_written, (c.n1 - CAM) / D);
        if (next) launch_f1(*next, &c, next->first ? 0 : obs_row);
        // rectangular sweep of the own rows that are live (< n1)
        const int nI = std::max(0, std::min(pt.blocks_per_rank, (c.n1 - pt.row0 + V_BLK - 1) / V_BLK));
        const int nJ = (c.n1 + V_BLK - 1) / V_BLK;
        const int ntiles = nI * nJ;
        SweepExtra ex{};
        ex.n = c.n1; ex.add_q = c.first ? 1 : 0; ex.nz = s->nz;
        ex.apply_kind = -1;
        ex.rect = 1; ex.row0 = pt.row0; ex.nJ = std::max(nJ, 1);
        ex.reserve_sm = -1;   // no SM reservation here (sweep.cuh)
        const int grid = std::max(1, std::min(ntiles, 2 * (s->sm_count - (next ? 1 : 0))));
        if (s->tail_tiles > 0 && ntiles > grid) {   // contiguous range + dynamic tail (sweep.cuh)
            ex.tail = s->d_tail + (s->sweep_seq & 1);
            ex.tail_reset = s->d_tail + ((s->sweep_seq + 1) & 1);
            ex.n_static = std::max(0, ntiles / grid - s->tail_tiles);
            ++s->sweep_seq;
        }
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3(SweepDual::THREADS);
        cfg.dynamicSmemBytes = SweepDual::smem_bytes(m4);
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = next ? 1 : 0;   // programmatic dependent of the k_f1 just launched
        cudaLaunchKernelEx(&cfg, k_sweep<SweepDual>, (const double*)s->d_P, s->d_P, pt.ldp, (const double*)s->d_Vr,
                           (const double*)s->d_Vc, m4, ntiles, 2 << 4, ex);
        have_front = next != nullptr;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(ASB_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return ASB_OK;
}

static int shard_chunks_dyn(asb_shard* s, const std::vector<ShChunk>& chunks) {
    if (chunks.empty()) return ASB_OK;
    return s->kind == ASB_KIND_EKF ? shard_chunks<0>(s, chunks) : shard_chunks<1>(s, chunks);
}

static int shard_frame_dyn(asb_shard* s, const int* idx, const double* pose, const unsigned char* isnew, int k,
                           int n0, int n1) {
    std::vector<ShChunk> chunks;
    append_chunks(s, chunks, idx, pose, isnew, k, n0, n1);
    return shard_chunks_dyn(s, chunks);
}

extern "C" {

int asb_shard_partition(int kind, int max_landmarks, int world, int rank, int* row0, int* row1, int* ld) {
    if ((kind != ASB_KIND_EKF && kind != ASB_KIND_ROT) || max_landmarks < 1 || world < 1 || rank < 0 || rank >= world)
        return fail(ASB_ERR_BAD_ARG, "bad kind/max_landmarks/world/rank");
    const Partition p = make_partition(kind, max_landmarks, world, rank);
    if (row0) *row0 = p.row0;
    if (row1) *row1 = p.row0 + p.rows;
    if (ld) *ld = p.ldp;
    return ASB_OK;
}

int asb_shard_unique_id(void* out, int bytes) {
    if (!out || bytes < (int)sizeof(ncclUniqueId)) return fail(ASB_ERR_BAD_ARG, "need %zu bytes", sizeof(ncclUniqueId));
    NcclApi* nc = nccl_api();
    if (!nc) return fail(ASB_ERR_CUDA, "libnccl.so.2 could not be loaded: %s", dlerror());
    ncclUniqueId id;
    ASB_NCCL(nc->GetUniqueId(&id));
    std::memcpy(out, &id, sizeof id);
    return ASB_OK;
}

int asb_shard_create(int kind, int quat_mode, const double pose0[10], const asb_noise* cfg, int max_landmarks,
                     int device, int rank, int world, const void* unique_id, asb_shard** out) {
    if (!out) return fail(ASB_ERR_BAD_ARG, "out is NULL");
    *out = nullptr;
    if (kind != ASB_KIND_EKF && kind != ASB_KIND_ROT) return fail(ASB_ERR_BAD_ARG, "kind %d", kind);
    if (quat_mode != ASB_QUAT_AS_WRITTEN && quat_mode != ASB_QUAT_SCALAR_FIRST)
        return fail(ASB_ERR_BAD_ARG, "quat_mode %d", quat_mode);
    if (max_landmarks < 1 || world < 1 || rank < 0 || rank >= world || !unique_id)
        return fail(ASB_ERR_BAD_ARG, "bad max_landmarks/rank/world/unique_id");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(ASB_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(ASB_ERR_BAD_ARG, "device %d of %d", device, ndev);
    NcclApi* nc = nccl_api();
    if (!nc) return fail(ASB_ERR_CUDA, "libnccl.so.2 could not be loaded");
    ASB_CUDA(cudaSetDevice(device));
    asb_shard* s = new (std::nothrow) asb_shard();
    if (!s) return fail(ASB_ERR_BAD_ARG, "out of host memory");
    s->kind = kind; s->quat_mode = quat_mode; s->device = device; s->rank = rank; s->world = world;
    s->D = kind == ASB_KIND_EKF ? 3 : 10;
    s->R = kind == ASB_KIND_EKF ? 3 : 7;
    s->W = CAM + s->D;
    s->nz = to_noise(kind, cfg);
    s->cap_lm = max_landmarks;
    s->ncap = CAM + s->D * max_landmarks;
    s->part = make_partition(kind, max_landmarks, world, rank);
    s->max_obs = std::max(64, std::min(max_landmarks * 2, 4096));
    s->kc_max = kind == ASB_KIND_EKF ? FrontLimits<0>::KC_MAX : FrontLimits<1>::KC_MAX;
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    s->sm_count = prop.multiProcessorCount;
    s->smem_optin = (int)prop.sharedMemPerBlockOptin;
    ncclUniqueId id;
    std::memcpy(&id, unique_id, sizeof id);
    ASB_NCCL(nc->CommInitRank(&s->comm, world, id, rank));
    ASB_CUDA(cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking));
    s->stream = s->own_stream;
    const Partition& pt = s->part;
    const size_t nt_max = CAM + (size_t)s->D * s->kc_max;
    ASB_CUDA(dmalloc(&s->d_state, (size_t)pt.ldp));
    ASB_CUDA(dmalloc(&s->d_P, (size_t)pt.rows * pt.ldp));
    ASB_CUDA(dmalloc(&s->d_PTpart, nt_max * nt_max));
    ASB_CUDA(dmalloc(&s->d_PT, nt_max * nt_max));
    ASB_CUDA(dmalloc(&s->d_At, nt_max * 64));
    ASB_CUDA(dmalloc(&s->d_y, (size_t)64));
    ASB_CUDA(dmalloc(&s->d_T, nt_max));
    ASB_CUDA(dmalloc(&s->d_Vr, (size_t)64 * V_LDB * pt.blocks_per_rank));
    ASB_CUDA(dmalloc(&s->d_Vc, (size_t)64 * V_LDB * pt.blocks_per_rank * world));
    ASB_CUDA(dmalloc(&s->d_dx[0], (size_t)pt.ldp));
    ASB_CUDA(dmalloc(&s->d_dx[1], (size_t)pt.ldp));
    ASB_CUDA(dmalloc(&s->d_blocks, (size_t)s->max_obs * s->R * s->W));
    ASB_CUDA(dmalloc(&s->d_nu, (size_t)s->max_obs * s->R));
    ASB_CUDA(dmalloc(&s->d_status, (size_t)1));
    ASB_CUDA(dmalloc(&s->d_tail, (size_t)2));
    ASB_CUDA(dmalloc(&s->d_idx, (size_t)SH_SLOTS * s->max_obs));
    ASB_CUDA(dmalloc(&s->d_pose, (size_t)SH_SLOTS * s->max_obs * 6));
    ASB_CUDA(dmalloc(&s->d_new, (size_t)SH_SLOTS * s->max_obs));
    ASB_CUDA(cudaMallocHost(&s->h_idx, (size_t)SH_SLOTS * s->max_obs * sizeof(int)));
    ASB_CUDA(cudaMallocHost(&s->h_pose, (size_t)SH_SLOTS * s->max_obs * 6 * sizeof(double)));
    ASB_CUDA(cudaMallocHost(&s->h_new, (size_t)SH_SLOTS * s->max_obs));
    for (auto& e : s->slot_ev) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    // state = pose0; P = init_cam * I10 on the rank that owns rows 0..9 (EKF:46-48)
    const double pose_default[CAM] = {0, 0, 0, 1, 0, 0, 0, 0, 0, 0};
    ASB_CUDA(cudaMemcpy(s->d_state, pose0 ? pose0 : pose_default, CAM * sizeof(double), cudaMemcpyHostToDevice));
    if (pt.row0 == 0) {
        std::vector<double> p0((size_t)CAM * pt.ldp, 0.0);
        for (int i = 0; i < CAM; ++i) p0[(size_t)i * pt.ldp + i] = s->nz.init_cam;
        ASB_CUDA(cudaMemcpy(s->d_P, p0.data(), p0.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    const int big = (int)prop.sharedMemPerBlockOptin - 2048;
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepDual>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_sh_f2<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_sh_f2<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ASB_CUDA(cudaFuncSetAttribute(k_sweep<SweepDual>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<0, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_f1<1, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_sh_f2<0>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaFuncSetAttribute(k_sh_f2<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    ASB_CUDA(cudaStreamSynchronize(s->stream));
    *out = s;
    return ASB_OK;
}

int asb_shard_destroy(asb_shard* s) {
    ASB_CHECK_SHARD(s);
    cudaStreamSynchronize(s->stream);
    if (s->comm) nccl_api()->CommDestroy(s->comm);
    cudaFree(s->d_state); cudaFree(s->d_P); cudaFree(s->d_PTpart); cudaFree(s->d_PT); cudaFree(s->d_At); cudaFree(s->d_y);
    cudaFree(s->d_tail); cudaFree(s->d_T); cudaFree(s->d_Vr); cudaFree(s->d_Vc); cudaFree(s->d_dx[0]); cudaFree(s->d_dx[1]);
    cudaFree(s->d_blocks); cudaFree(s->d_nu); cudaFree(s->d_status); cudaFree(s->d_idx); cudaFree(s->d_pose);
    cudaFree(s->d_new); cudaFree(s->s_idx); cudaFree(s->s_pose); cudaFree(s->s_new);
    cudaFreeHost(s->h_idx); cudaFreeHost(s->h_pose); cudaFreeHost(s->h_new);
    for (auto& e : s->slot_ev) cudaEventDestroy(e);
    cudaStreamDestroy(s->own_stream);
    s->magic = 0;
    delete s;
    return ASB_OK;
}

int asb_shard_set_stream(asb_shard* s, void* cuda_stream) {
    ASB_CHECK_SHARD(s);
    ASB_CUDA(cudaStreamSynchronize(s->stream));
    s->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : s->own_stream;
    return ASB_OK;
}

int asb_shard_synchronize(asb_shard* s) {
    ASB_CHECK_SHARD(s);
    int st = 0;
    ASB_CUDA(cudaMemcpyAsync(&st, s->d_status, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    ASB_CUDA(cudaStreamSynchronize(s->stream));
    if (st != 0) {
        ASB_CUDA(cudaMemsetAsync(s->d_status, 0, sizeof(int), s->stream));
        return fail(st, "innovation covariance S was not positive definite / finite");
    }
    return ASB_OK;
}

// observe (EKF:58-82): every rank is given the same ids / poses; association is replicated.
int asb_shard_observe(asb_shard* s, const int32_t* ids, const double* poses, int k) {
    ASB_CHECK_SHARD(s);
    if (k <= 0 || !ids || !poses) return fail(ASB_ERR_BAD_ARG, "bad ids/poses/k");
    if (k > s->max_obs) return fail(ASB_ERR_CAPACITY, "%d observations > per-frame capacity %d", k, s->max_obs);
    int fresh = 0;
    {
        std::unordered_map<int32_t, int> seen;
        for (int o = 0; o < k; ++o)
            if (!s->map.count(ids[o]) && !seen.count(ids[o])) { seen[ids[o]] = 1; ++fresh; }
    }
    if (s->nlm + fresh > s->cap_lm)
        return fail(ASB_ERR_CAPACITY, "%d landmarks > max_landmarks %d", s->nlm + fresh, s->cap_lm);
    const int n0 = s->n_state();
    const int sl = s->slot;
    s->slot = (s->slot + 1) % SH_SLOTS;
    ASB_CUDA(cudaEventSynchronize(s->slot_ev[sl]));
    const size_t o0 = (size_t)sl * s->max_obs;
    for (int o = 0; o < k; ++o) {
        auto it = s->map.find(ids[o]);
        if (it == s->map.end()) {
            s->h_idx[o0 + o] = s->nlm;
            s->h_new[o0 + o] = 1;
            s->map[ids[o]] = s->nlm++;
            s->id_order.push_back(ids[o]);
        } else {
            s->h_idx[o0 + o] = it->second;
            s->h_new[o0 + o] = 0;
        }
    }
    std::memcpy(s->h_pose + 6 * o0, poses, (size_t)k * 6 * sizeof(double));
    cudaStream_t st = s->stream;
    ASB_CUDA(cudaMemcpyAsync(s->d_idx + o0, s->h_idx + o0, k * sizeof(int), cudaMemcpyHostToDevice, st));
    ASB_CUDA(cudaMemcpyAsync(s->d_pose + 6 * o0, s->h_pose + 6 * o0, (size_t)k * 6 * sizeof(double), cudaMemcpyHostToDevice, st));
    ASB_CUDA(cudaMemcpyAsync(s->d_new + o0, s->h_new + o0, k, cudaMemcpyHostToDevice, st));
    ASB_CUDA(cudaEventRecord(s->slot_ev[sl], st));
    return shard_frame_dyn(s, s->d_idx + o0, s->d_pose + 6 * o0, s->d_new + o0, k, n0, s->n_state());
}

int asb_shard_stream_upload(asb_shard* s, const int32_t* ids, const double* poses, const int64_t* offsets,
                            int n_frames) {
    ASB_CHECK_SHARD(s);
    if (!ids || !poses || !offsets || n_frames < 1) return fail(ASB_ERR_BAD_ARG, "bad stream");
    const int64_t total = offsets[n_frames];
    std::vector<int> idx((size_t)std::max<int64_t>(total, 1));
    std::vector<unsigned char> isnew((size_t)std::max<int64_t>(total, 1), 0);
    std::unordered_map<int32_t, int> sim = s->map;
    int nlm = s->nlm;
    s->frames.clear();
    for (int t = 0; t < n_frames; ++t) {
        ShFrame fr;
        fr.off = offsets[t];
        fr.k = (int)(offsets[t + 1] - offsets[t]);
        if (fr.k > s->max_obs) return fail(ASB_ERR_CAPACITY, "frame %d: %d observations > %d", t, fr.k, s->max_obs);
        fr.n0 = CAM + s->D * nlm;
        for (int64_t o = fr.off; o < fr.off + fr.k; ++o) {
            auto it = sim.find(ids[o]);
            if (it == sim.end()) {
                if (nlm + 1 > s->cap_lm) return fail(ASB_ERR_CAPACITY, "stream needs > %d landmarks", s->cap_lm);
                idx[o] = nlm;
                isnew[o] = 1;
                sim[ids[o]] = nlm++;
                fr.new_ids.push_back(ids[o]);
            } else {
                idx[o] = it->second;
            }
        }
        fr.n1 = CAM + s->D * nlm;
        s->frames.push_back(std::move(fr));
    }
    ASB_CUDA(cudaStreamSynchronize(s->stream));
    cudaFree(s->s_idx); cudaFree(s->s_pose); cudaFree(s->s_new);
    s->s_idx = nullptr; s->s_pose = nullptr; s->s_new = nullptr;
    ASB_CUDA(dmalloc(&s->s_idx, (size_t)total));
    ASB_CUDA(dmalloc(&s->s_pose, (size_t)total * 6));
    ASB_CUDA(dmalloc(&s->s_new, (size_t)total));
    ASB_CUDA(cudaMemcpy(s->s_idx, idx.data(), total * sizeof(int), cudaMemcpyHostToDevice));
    ASB_CUDA(cudaMemcpy(s->s_pose, poses, (size_t)total * 6 * sizeof(double), cudaMemcpyHostToDevice));
    ASB_CUDA(cudaMemcpy(s->s_new, isnew.data(), total, cudaMemcpyHostToDevice));
    s->next_frame = 0;
    return ASB_OK;
}

int asb_shard_stream_run(asb_shard* s, int t0, int t1) {
    ASB_CHECK_SHARD(s);
    if (s->frames.empty()) return fail(ASB_ERR_NO_STREAM, "stream_run before stream_upload");
    if (t0 != s->next_frame || t1 < t0 || t1 > (int)s->frames.size())
        return fail(ASB_ERR_BAD_ARG, "frames must run in order: next is %d, got [%d,%d)", s->next_frame, t0, t1);
    std::vector<ShChunk> chunks;
    for (int t = t0; t < t1; ++t) {
        const ShFrame& fr = s->frames[t];
        for (int32_t id : fr.new_ids) {
            s->map[id] = s->nlm++;
            s->id_order.push_back(id);
        }
        if (fr.k == 0) continue;
        append_chunks(s, chunks, s->s_idx + fr.off, s->s_pose + 6 * fr.off, s->s_new + fr.off, fr.k, fr.n0, fr.n1);
    }
    s->next_frame = t1;
    return shard_chunks_dyn(s, chunks);
}

int asb_shard_get_dims(asb_shard* s, int* n_landmarks, int* n_state) {
    ASB_CHECK_SHARD(s);
    if (n_landmarks) *n_landmarks = s->nlm;
    if (n_state) *n_state = s->n_state();
    return ASB_OK;
}

int asb_shard_get_state(asb_shard* s, double* out, int n_state) {
    ASB_CHECK_SHARD(s);
    if (!out || n_state != s->n_state()) return fail(ASB_ERR_BAD_ARG, "n_state %d != %d", n_state, s->n_state());
    int rc = asb_shard_synchronize(s);
    ASB_CUDA(cudaMemcpy(out, s->d_state, n_state * sizeof(double), cudaMemcpyDeviceToHost));
    return rc;
}

// This rank's live rows of the covariance: rows [row0, min(row1, n)) x n, row-major into out
// (capacity in doubles); *n_rows = how many rows were written (0 when the map has not reached row0).
int asb_shard_get_cov_rows(asb_shard* s, double* out, long long capacity, int* n_rows) {
    ASB_CHECK_SHARD(s);
    const int n = s->n_state();
    const int rows = std::max(0, std::min(s->part.row0 + s->part.rows, n) - s->part.row0);
    if (n_rows) *n_rows = rows;
    if (rows == 0) return asb_shard_synchronize(s);
    if (!out || capacity < (long long)rows * n) return fail(ASB_ERR_BAD_ARG, "capacity %lld < %lld", capacity, (long long)rows * n);
    int rc = asb_shard_synchronize(s);
    ASB_CUDA(cudaMemcpy2D(out, (size_t)n * sizeof(double), s->d_P, (size_t)s->part.ldp * sizeof(double),
                          (size_t)n * sizeof(double), rows, cudaMemcpyDeviceToHost));
    return rc;
}

int asb_shard_get_landmark_ids(asb_shard* s, int32_t* ids, int capacity) {
    ASB_CHECK_SHARD(s);
    if (!ids || capacity < s->nlm) return fail(ASB_ERR_BAD_ARG, "capacity %d < %d", capacity, s->nlm);
    std::memcpy(ids, s->id_order.data(), s->nlm * sizeof(int32_t));
    return ASB_OK;
}

}  // extern "C"
