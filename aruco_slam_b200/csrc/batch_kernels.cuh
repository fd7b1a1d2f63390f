// Batched engine: many independent small filters (BASELINE configs C2 / C5), one CTA per
// sequence, the whole filter — state, covariance (packed lower triangle), id table —
// resident in shared memory for all frames of a run; observations are read once from HBM.
//
// Per frame (EKF.observe, extended_kalman_filter.py:58-82):
//   association of ids on the device (first-seen order, EKF:75-78, bit-exact integers),
//   augmentation (EKF:239-290), predict (EKF:95-105), then the update (EKF:107-156)
//   processed in chunks of CK markers: with R = r I block-diagonal and H evaluated once
//   at the pre-update state, chunk-sequential conditioning with the innovation corrected
//   by H delta is algebraically the batch update (DESIGN.md "chunked update").
//   Each chunk: G = H P (one thread per state column), S = G H^T + R, Cholesky of [S; nu^T] by ONE
//   warp (a row per lane in registers, the column of L exchanged through shared memory) while the other
//   warps evaluate the next chunk's measurement model, V = L^-1 G and delta += V^T y (one thread per
//   column), P -= V^T V on the tile-blocked lower triangle (one thread per 4 x 8 register tile).
// The SINGLE instantiation runs the same frame for ONE asb_filter with a small map (filter_single.cu,
// common.cuh SmallFrame): its dense covariance in, the same phases, its dense covariance out.
#pragma once
#include <cstdint>

#include "../../include/aruco_slam_b200.h"
#include "model.cuh"

namespace asb {

struct BatchParams {
    double* states;        // n_seq x state_stride
    double* covs;          // n_seq x cov_stride, packed lower triangle, row-major
    int* nlm;              // n_seq
    int* lm_ids;           // n_seq x max_lm, ids in index order
    int* status;           // n_seq
    const int* ids;        // packed observations
    const double* poses;   // packed (.,6)
    const long long* offsets;  // n_seq * n_frames + 1
    long long state_stride, cov_stride;
    int n_seq, n_frames, t0, t1, max_lm, max_obs, as_written;
    int seq0, seq1;        // this launch handles sequences [seq0, seq1)
    int obs_seqs;          // distinct observation streams: sequence s reads stream s % obs_seqs (= n_seq unless tiled)
    Noise nz;
    unsigned long long* prof;   // 10 phase accumulators (clocks) of CTA 0, may be null
    // SINGLE instantiation (one frame of ONE asb_filter whose map is small enough for an SM, filter_single.cu):
    // the covariance is that filter's dense array (lower triangle valid, leading dimension ld), the host has
    // already associated the ids (landmark index + first-sighting flag per observation), the frame's k poses
    // are at `poses`; raw innovations / measurement blocks / dx are left for asb_get_last_update
    const int* one_idx; const unsigned char* one_new;
    int one_k, one_nlm0, one_nlm1, ld;
    double *nu_out, *blocks_out, *dx_out;
    double* one_state_host; int* one_status_host;   // page-locked host mirror of the posterior state + status (may be null)
};

__host__ __device__ __forceinline__ int tri(int i) { return (i * (i + 1)) >> 1; }

// Shared-memory layout of the covariance ("tile-blocked"): the lower triangle as 8 x 8 tiles,
// tile (I, J <= I) at tile index tri(I) + J, 64 doubles; diagonal tiles are kept full (both
// halves).  Inside a tile the 32 16-byte chunks (row r, column pair q -> chunk 4 r + q) are
// XOR-swizzled with the tile index, so threads that each own a different (half) tile hit
// different banks with the same logical chunk.  HBM keeps the packed lower triangle.
// The panels (G, then V) keep every 8-column group in GW = 10 doubles: four consecutive groups
// then start in four different 16-byte banks, which is what the 128-bit loads of the rank-m
// update need (with dense rows groups J and J + 2 collide).
__host__ __device__ __forceinline__ int tiled_doubles(int ncap) { return tri((ncap + 7) >> 3) * 64; }
// row stride of the panels: whole 8-column tiles
constexpr int GW = 10;
__host__ __device__ __forceinline__ int batch_npad(int ncap) { return GW * ((ncap + 7) >> 3) + 2; }
__host__ __device__ __forceinline__ int gcol(int j) { return j + ((j >> 3) << 1); }   // column j inside a panel row
__device__ __forceinline__ int tchunk(int tile, int chunk) {
    // low bits mixed with 2 tile + (chunk >= 16): the 8 lanes of a quarter-warp — 4 tiles x 2 half tiles —
    // land in 8 different 16-byte banks (bit 4 of the chunk is preserved, so this is a bijection)
    return (tile << 6) + ((chunk ^ ((2 * tile + (chunk >> 4)) & 15)) << 1);
}
__device__ __forceinline__ int tidx(int i, int j) {   // i >= j
    return tchunk(tri(i >> 3) + (j >> 3), ((i & 7) << 2) + ((j & 7) >> 1)) + (j & 1);
}
// symmetric access into the packed lower triangle
__device__ __forceinline__ double psym(const double* P, int i, int j) {
    return j <= i ? P[tri(i) + j] : P[tri(j) + i];
}

template <int KIND, int CK, int NT, bool SINGLE = false, bool PROF = false>
__global__ void __launch_bounds__(NT, (NT <= 128 && KIND == 0 && !SINGLE) ? 3 : 1)
k_batch(BatchParams p) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D;
    constexpr int MC = R * CK;           // max rows per chunk
    constexpr int BLK = (R * W + 1) & ~1;   // doubles per observation block in shared memory (even: 16-byte aligned)
    constexpr int LDS_S = (MC + 1) | 1;  // leading dim of S (+ appended nu row)
    __shared__ __align__(16) double s_col[64];   // single-warp Cholesky: column j of L, two buffers
    extern __shared__ __align__(16) unsigned char smraw[];
    const int tid = threadIdx.x;
    const int ncap = CAM + D * p.max_lm;
    const int npad = batch_npad(ncap);

    // ---- shared memory carve-up -----------------------------------------------------
    double* s_state = reinterpret_cast<double*>(smraw);       // ncap
    double* s_delta = s_state + npad;                         // ncap
    double* s_G = s_delta + npad;                             // MC x npad
    // measurement-model blocks and raw innovations of TWO chunks: the next chunk's model is evaluated by
    // the idle warps while warp 0 factors S (the state it is evaluated at does not change inside a frame)
    double* s_blk0 = s_G + (size_t)MC * npad;                 // 2 x (CK x R x W)
    double* s_nu0 = s_blk0 + 2 * CK * BLK;                    // 2 x MC (padded)
    constexpr int NU_LD = (MC + 3) & ~3;
    double* s_S = s_nu0 + 2 * NU_LD;                          // (MC+1) x LDS_S
    double* s_dinv = s_S + (MC + 1) * LDS_S;                  // MC
    double* s_pose = s_dinv + ((MC + 3) & ~3);                // max_obs x 6
    int* s_ids = reinterpret_cast<int*>(s_pose + (size_t)p.max_obs * 6);   // max_obs
    int* s_idx = s_ids + p.max_obs;                           // max_obs
    int* s_new = s_idx + p.max_obs;                           // max_obs
    int* s_tab = s_new + p.max_obs;                           // max_lm
    int* s_misc = s_tab + p.max_lm;                           // [0]=nlm [1]=bad
    double* s_P = reinterpret_cast<double*>(
        (reinterpret_cast<uintptr_t>(s_misc + 4) + 15) & ~uintptr_t(15));  // packed, if it fits

    // phase profile of CTA 0 (clock64 deltas accumulated per phase in registers), measurement aid
    // (PROF instantiation only - asb_batch_profile switches a handle to it: the ten 64-bit accumulators cost the
    // production kernel, which sits at its register cap, twenty registers and a spill slot each)
    [[maybe_unused]] long long prof_t = PROF ? clock64() : 0, prof_acc[PROF ? 10 : 1] = {};
    const bool prof_on = PROF && p.prof && blockIdx.x == 0 && tid == 0;
#define BPROF(slot)                                  \
    do {                                             \
        if constexpr (PROF) if (prof_on) {           \
            const long long now_ = clock64();        \
            prof_acc[slot] += now_ - prof_t;         \
            prof_t = now_;                           \
        }                                            \
    } while (0)
    for (int seq = p.seq0 + blockIdx.x; seq < p.seq1; seq += gridDim.x) {
        double* g_state = p.states + (long long)seq * p.state_stride;
        double* g_P = p.covs + (long long)seq * p.cov_stride;
        int* g_tab = p.lm_ids + (long long)seq * p.max_lm;
        double* P = s_P;   // the covariance lives in shared memory (filters that do not fit run on the wide engine)
        auto pget = [&](int i, int j) -> double {   // symmetric read
            const int a = max(i, j), b = min(i, j);
            return P[tidx(a, b)];
        };
        auto pset = [&](int i, int j, double v) {   // i >= j; diagonal tiles hold both halves
            P[tidx(i, j)] = v;
            if ((i >> 3) == (j >> 3)) P[tchunk(tri(i >> 3) + (j >> 3), ((j & 7) << 2) + ((i & 7) >> 1)) + (i & 1)] = v;
        };
        int nlm = SINGLE ? p.one_nlm0 : p.nlm[seq];
        int n = CAM + D * nlm;
        for (int i = tid; i < n; i += NT) s_state[i] = g_state[i];
        if constexpr (!SINGLE)
            for (int i = tid; i < nlm; i += NT) s_tab[i] = g_tab[i];
        {
            for (int e = tid; e < tiled_doubles(ncap); e += NT) s_P[e] = 0.0;
            __syncthreads();
            if constexpr (SINGLE) {
                // dense rows; eight independent loads per thread in flight (a load -> store -> next-load chain
                // costs an L2 round trip per element: 28 us for N = 110)
                const int nn = n * n;
                for (int base = 0; base < nn; base += 8 * NT) {
                    double v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int e = base + u * NT + tid, i = e / n, j = e - i * n;
                        v[u] = (e < nn && j <= i) ? __ldcg(g_P + (long long)i * p.ld + j) : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int e = base + u * NT + tid, i = e / n, j = e - i * n;
                        if (e < nn && j <= i) pset(i, j, v[u]);
                    }
                }
            } else {
                for (int i = tid; i < n; i += NT)
                    for (int j = 0; j <= i; ++j) pset(i, j, g_P[tri(i) + j]);
            }
        }
        if (tid == 0) s_misc[1] = 0;
        __syncthreads();

        for (int t = p.t0; t < p.t1; ++t) {
            long long ob = 0;
            int k = p.one_k;
            if constexpr (!SINGLE) {
                const long long orec = (long long)(seq % p.obs_seqs) * p.n_frames + t;
                ob = p.offsets[orec];
                const long long ob1 = p.offsets[orec + 1];
                k = (int)(ob1 - ob);
                // next frame's observations towards L2 now (its count is not known yet: a window of max_obs, clipped
                // to the stream): by the time this frame's update is done they are an L2 hit, not a DRAM round trip
                if (tid == 0 && t + 1 < p.t1) {
                    const long long last = p.offsets[(long long)p.obs_seqs * p.n_frames];
                    const long long w = min((long long)p.max_obs, last - ob1);
                    if (w > 0) {
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.poses + 6 * ob1),
                                     "r"((uint32_t)(w * 48)) : "memory");
                        const long long i0 = ob1 & ~3ll;
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.ids + i0),
                                     "r"((uint32_t)(((ob1 + w - i0) * 4 + 15) & ~15ll)) : "memory");
                    }
                }
            }
            if (k == 0) continue;  // no detection: no predict either (base_filter.py:197-204)
            if (k > p.max_obs) {
                if (tid == 0) s_misc[1] = ASB_ERR_CAPACITY;
                break;
            }
            // ---- load the frame, associate ids ---------------------------------------
            for (int e = tid; e < 6 * k; e += NT) s_pose[e] = p.poses[6 * ob + e];
            int any_unseen = 1;
            if constexpr (SINGLE) {   // associated on the host (first-seen order, the same rule as below)
                for (int o = tid; o < k; o += NT) { s_idx[o] = p.one_idx[o]; s_new[o] = p.one_new[o]; }
                if (tid == 0) s_misc[0] = p.one_nlm1;
            } else {
                for (int o = tid; o < k; o += NT) s_ids[o] = p.ids[ob + o];
                __syncthreads();
                int unseen = 0;
                // eight lanes per observation walk the id table with stride 8 (ids are unique in the table: at most one
                // lane finds it), combined with three shuffles - a quarter of the serial probes of one thread per id
                for (int o0 = 0; o0 < k; o0 += NT / 8) {
                    const int o = o0 + (tid >> 3);
                    const int id = o < k ? s_ids[o] : -1;
                    int found = -1;
                    if (o < k)
                        for (int i = tid & 7; i < nlm; i += 8)
                            if (s_tab[i] == id) found = i;
                    found = max(found, __shfl_xor_sync(0xffffffffu, found, 1));
                    found = max(found, __shfl_xor_sync(0xffffffffu, found, 2));
                    found = max(found, __shfl_xor_sync(0xffffffffu, found, 4));
                    if (o < k && (tid & 7) == 0) { s_idx[o] = found; s_new[o] = 0; }
                    unseen |= (o < k && found < 0);
                }
                // (the common frame sees known markers only: then the serial pass below and its barrier are skipped)
                any_unseen = __syncthreads_or(unseen);
                if (any_unseen && tid == 0) {  // unseen ids, in frame order; duplicates bind to the first
                    int cur = nlm;
                    for (int o = 0; o < k; ++o) {
                        if (s_idx[o] >= 0) continue;
                        const int id = s_ids[o];
                        int found = -1;
                        for (int i = nlm; i < cur; ++i)
                            if (s_tab[i] == id) { found = i; break; }
                        if (found < 0) {
                            if (cur >= p.max_lm) { s_misc[1] = ASB_ERR_CAPACITY; found = 0; }
                            else { s_tab[cur] = id; found = cur++; s_new[o] = 1; }
                        }
                        s_idx[o] = found;
                    }
                    s_misc[0] = cur;
                }
            }
            if (SINGLE || any_unseen) __syncthreads();
            BPROF(0);
            const int nlm1 = (SINGLE || any_unseen) ? s_misc[0] : nlm;
            if (s_misc[1] == ASB_ERR_CAPACITY) break;
            const int n1 = CAM + D * nlm1;
            // ---- augmentation: new landmark states, block-diagonal covariance growth -
            if (nlm1 > nlm) {
                for (int o = tid; o < k; o += NT) {
                    if (!s_new[o]) continue;
                    double lm[D];
                    new_landmark<KIND>(s_state, s_pose + 6 * o, lm);
#pragma unroll
                    for (int c = 0; c < D; ++c) s_state[CAM + D * s_idx[o] + c] = lm[c];
                }
                for (int e = tri(n) + tid; e < tri(n1); e += NT) {   // new rows n..n1-1, all columns <= row
                    int i = (int)((sqrtf(8.0f * e + 1.0f) - 1.0f) * 0.5f);
                    while (tri(i + 1) <= e) ++i;
                    while (tri(i) > e) --i;
                    const int j = e - tri(i);
                    pset(i, j, i == j ? p.nz.init_lm : 0.0);
                }
                __syncthreads();
            }
            nlm = nlm1;
            n = n1;
            // ---- predict: P += diag(Q); delta = 0 -------------------------------------
            for (int i = tid; i < n; i += NT) {
                pset(i, i, pget(i, i) + process_noise(p.nz, i));
                s_delta[i] = 0.0;
            }
            {   // the panel's padding columns n .. 8 ceil(n / 8) stay zero for every chunk of this frame:
                // G and V only ever write columns < n (the rank-m update reads whole 8-column tiles)
                const int npadc = 8 * ((n + 7) >> 3) - n;
                for (int e = tid; e < MC * npadc; e += NT) s_G[(e / npadc) * npad + gcol(n + e % npadc)] = 0.0;
            }
            // (no barrier: the first chunk's model evaluation below touches neither P, delta nor the panel; the barrier
            // after it covers both)

            BPROF(1);
            // ---- update, CK markers at a time -----------------------------------------
            // measurement model of observation c0 + o at the pre-update state -> buffer `which`
            // (blocks + raw innovation; the correction nu' = nu - H delta is applied where S is built)
            auto eval_model = [&](int which, int c0m, int o) {
                double lm[D], lnu[R];
                const int li = CAM + D * s_idx[c0m + o];
#pragma unroll
                for (int c = 0; c < D; ++c) lm[c] = s_state[li + c];
                observation_model<KIND>(s_state, lm, s_pose + 6 * (c0m + o), lnu, s_blk0 + (which * CK + o) * BLK);
#pragma unroll
                for (int r = 0; r < R; ++r) s_nu0[which * NU_LD + R * o + r] = lnu[r];
                if constexpr (SINGLE) {
                    const double* blk = s_blk0 + (which * CK + o) * BLK;
#pragma unroll
                    for (int r = 0; r < R; ++r) p.nu_out[R * (c0m + o) + r] = lnu[r];
                    for (int c = 0; c < R * W; ++c) p.blocks_out[(size_t)(c0m + o) * R * W + c] = blk[c];
                }
            };
            for (int o = tid; o < min(CK, k); o += NT) eval_model(0, 0, o);
            __syncthreads();
            BPROF(2);
            for (int c0 = 0; c0 < k; c0 += CK) {
                const int kc = min(CK, k - c0);
                const int mc = R * kc;
                const int which = (c0 / CK) & 1;
                const double* s_blk = s_blk0 + which * CK * BLK;
                const double* s_nu = s_nu0 + which * NU_LD;
                const int kc_next = min(CK, k - c0 - CK);   // <= 0: this is the last chunk
                // G = H P  (mc x n): one thread per state column; the 10 camera rows of P are
                // read once per column, then D landmark rows per observation
                for (int j = tid; j < n; j += NT) {
                    double pc[CAM];
                    if (j >= 8) {
                        // P[j][0..9] = row j & 7 of tiles (j >> 3, 0) and (j >> 3, 1): five 16-byte chunks with
                        // one tile index instead of ten swizzled scalar accesses (diagonal tiles hold both halves)
                        const int t0 = tri(j >> 3), r4 = (j & 7) << 2;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const double2 v = *reinterpret_cast<const double2*>(P + tchunk(t0, r4 + q));
                            pc[2 * q] = v.x; pc[2 * q + 1] = v.y;
                        }
                        const double2 v = *reinterpret_cast<const double2*>(P + tchunk(t0 + 1, r4));
                        pc[8] = v.x; pc[9] = v.y;
                    } else {
#pragma unroll
                        for (int c = 0; c < CAM; ++c) pc[c] = pget(c, j);
                    }
                    for (int o = 0; o < kc; ++o) {
                        const int li = CAM + D * s_idx[c0 + o];
                        double pl[D], acc[R];
#pragma unroll
                        for (int c = 0; c < D; ++c) pl[c] = pget(li + c, j);
                        // the observation's R x W block as 128-bit broadcast loads (blocks start 16-byte aligned:
                        // stride BLK), half the shared-memory instructions of one 64-bit load per coefficient
                        if constexpr (BLK <= 40) {
                            double hb[BLK];
                            const double2* blk2 = reinterpret_cast<const double2*>(s_blk + o * BLK);
#pragma unroll
                            for (int e = 0; e < BLK / 2; ++e) {
                                const double2 v = blk2[e];
                                hb[2 * e] = v.x; hb[2 * e + 1] = v.y;
                            }
#pragma unroll
                            for (int r = 0; r < R; ++r) {
                                double sv = 0.0;
#pragma unroll
                                for (int c = 0; c < CAM; ++c) sv += hb[r * W + c] * pc[c];
#pragma unroll
                                for (int c = 0; c < D; ++c) sv += hb[r * W + CAM + c] * pl[c];
                                acc[r] = sv;
                            }
                        } else {   // EKF_Rotations: 140 coefficients do not fit in registers
                            const double* blk = s_blk + o * BLK;
#pragma unroll
                            for (int r = 0; r < R; ++r) {
                                double sv = 0.0;
#pragma unroll
                                for (int c = 0; c < CAM; ++c) sv += blk[r * W + c] * pc[c];
#pragma unroll
                                for (int c = 0; c < D; ++c) sv += blk[r * W + CAM + c] * pl[c];
                                acc[r] = sv;
                            }
                        }
#pragma unroll
                        for (int r = 0; r < R; ++r) s_G[(R * o + r) * npad + gcol(j)] = acc[r];
                    }
                }
                __syncthreads();
                BPROF(3);
                // S = G H^T + R (lower) with nu appended as row mc
                // tasks: the tri(mc) lower-triangle entries, then the mc entries of the nu row (one round of the CTA
                // for mc = 12: 90 tasks; indexing the full square left a third of the threads with nothing to do)
                for (int e = tid; e < tri(mc) + mc; e += NT) {
                    int a, b;
                    if (e < tri(mc)) {
                        a = (int)((sqrtf(8.0f * e + 1.0f) - 1.0f) * 0.5f);
                        while (tri(a + 1) <= e) ++a;
                        while (tri(a) > e) --a;
                        b = e - tri(a);
                    } else {
                        a = mc;
                        b = e - tri(mc);
                    }
                    double s;
                    if (a < mc) {
                        const int o2 = b / R, ib = b - o2 * R;
                        const double* hb = s_blk + o2 * BLK + ib * W;
                        const double* g = s_G + a * npad;
                        const int li = CAM + D * s_idx[c0 + o2];
                        s = (a == b) ? p.nz.r : 0.0;
#pragma unroll
                        for (int c = 0; c < CAM; ++c) s += g[gcol(c)] * hb[c];
#pragma unroll
                        for (int c = 0; c < D; ++c) s += g[gcol(li + c)] * hb[CAM + c];
                    } else {
                        s = s_nu[b];
                        if (c0 > 0) {   // nu' = nu - H delta (delta = the correction accumulated by the chunks before)
                            const int o2 = b / R, ib = b - o2 * R;
                            const double* hb = s_blk + o2 * BLK + ib * W;
                            const int li = CAM + D * s_idx[c0 + o2];
#pragma unroll
                            for (int c = 0; c < CAM; ++c) s -= hb[c] * s_delta[c];
#pragma unroll
                            for (int c = 0; c < D; ++c) s -= hb[CAM + c] * s_delta[li + c];
                        }
                    }
                    s_S[a * LDS_S + b] = s;
                }
                __syncthreads();
                BPROF(4);
                bool bad = false;
                static_assert(MC + 1 <= 32, "the single-warp Cholesky holds one row of [S; nu^T] per lane");
                {
                    // Cholesky of [S; nu^T] by ONE warp, lane i holding row i in registers; the pivot is
                    // broadcast with a shuffle, column j of L through a 32-double shared buffer (no CTA
                    // barrier inside the factorisation)
                    if (tid < 32) {
                        const int i = tid;
                        double a[MC], dv[MC];
#pragma unroll
                        for (int c = 0; c < MC; ++c)
                            a[c] = (i <= mc && c < mc && (c <= i)) ? s_S[i * LDS_S + c] : 0.0;
#pragma unroll
                        for (int j = 0; j < MC; ++j) {
                            if (j < mc) {
                                const double pv = __shfl_sync(0xffffffffu, a[j], j);
                                if (!(pv > 0.0) || !(pv < 1e300)) bad = true;
                                const double rs = rsqrt(pv);
                                dv[j] = rs;
                                const double lij = a[j] * rs;
                                a[j] = lij;
                                // column j of L to every lane through shared memory: one store and ceil((MC - j) / 2)
                                // 128-bit broadcast loads instead of 2 (MC - 1 - j) shuffle instructions per pivot
                                double* col = s_col + 32 * (j & 1);
                                col[i] = lij;
                                __syncwarp();
#pragma unroll
                                for (int c = (j + 1) & ~1; c < MC; c += 2) {
                                    const double2 l2 = *reinterpret_cast<const double2*>(col + c);
                                    if (c > j) a[c] -= lij * l2.x;
                                    if (c + 1 < MC) a[c + 1] -= lij * l2.y;
                                }
                            }
                        }
#pragma unroll
                        for (int c = 0; c < MC; ++c) {
                            if (c < mc && c <= i && i <= mc) s_S[i * LDS_S + c] = a[c];
                            if (i == 0 && c < mc) s_dinv[c] = dv[c];
                        }
                    } else if (tid - 32 < kc_next) {
                        eval_model(which ^ 1, c0 + CK, tid - 32);   // warps 1..: the next chunk's model, meanwhile
                    }
                    __syncthreads();
                }
                BPROF(5);
                if (bad && tid == 0) s_misc[1] = ASB_ERR_NOT_SPD;
                // V = L^-1 G (in place), delta += V^T y   (y = appended row of S): one thread per
                // column, the column in registers, forward substitution fully unrolled
                for (int j = tid; j < n; j += NT) {
                    double v[MC];
#pragma unroll
                    for (int a = 0; a < MC; ++a) v[a] = a < mc ? s_G[a * npad + gcol(j)] : 0.0;
                    double dsum = 0.0;
#pragma unroll
                    for (int a = 0; a < MC; ++a) {
                        if (a < mc) {
                            const double* la = s_S + a * LDS_S;
                            double t = v[a];
#pragma unroll
                            for (int b = 0; b < a; ++b) t -= la[b] * v[b];
                            t *= s_dinv[a];
                            v[a] = t;
                            s_G[a * npad + gcol(j)] = t;
                            dsum += t * s_S[mc * LDS_S + a];
                        }
                    }
                    s_delta[j] += dsum;
                }
                __syncthreads();
                BPROF(6);
                // P -= V^T V.  Shared-memory covariance: one thread per half 8 x 8 tile (4 rows x 8
                // columns in registers; 4 + 8 panel values per rank-1 term feed 32 FMAs — the 4 x 4
                // blocks of the HBM path re-read V twice as often and were shared-memory-bound).
                // [mma.sync m8n8k4 was tried: one warp issues a DMMA only every ~64 clocks, so
                // with 4 warps per CTA it lost to the FMA pipe.]
                {
                    const int nt8 = (n + 7) >> 3;
                    // (columns n..8 nt8 of the panel are zero: cleared once per frame, never written)
                    for (int task = tid; task < 2 * tri(nt8); task += NT) {
                        const int tile = task >> 1, h = task & 1;
                        int I = (int)((sqrtf(8.0f * tile + 1.0f) - 1.0f) * 0.5f);
                        while (tri(I + 1) <= tile) ++I;
                        while (tri(I) > tile) --I;
                        const int J = tile - tri(I);
                        double acc[4][8];
#pragma unroll
                        for (int r = 0; r < 4; ++r)
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const double2 v = *reinterpret_cast<const double2*>(P + tchunk(tile, 4 * (4 * h + r) + q));
                                acc[r][2 * q] = v.x; acc[r][2 * q + 1] = v.y;
                            }
                        const double* xb = s_G + GW * I + 4 * h;
                        const double* yb = s_G + GW * J;
#pragma unroll 2
                        for (int a = 0; a < mc; ++a) {
                            const double2* xp = reinterpret_cast<const double2*>(xb + a * npad);
                            const double2* yp = reinterpret_cast<const double2*>(yb + a * npad);
                            const double2 x0 = xp[0], x1 = xp[1], y0 = yp[0], y1 = yp[1], y2 = yp[2], y3 = yp[3];
                            const double x[4] = {x0.x, x0.y, x1.x, x1.y};
                            const double y[8] = {y0.x, y0.y, y1.x, y1.y, y2.x, y2.y, y3.x, y3.y};
#pragma unroll
                            for (int r = 0; r < 4; ++r)
#pragma unroll
                                for (int c = 0; c < 8; ++c) acc[r][c] -= x[r] * y[c];
                        }
#pragma unroll
                        for (int r = 0; r < 4; ++r)
#pragma unroll
                            for (int q = 0; q < 4; ++q)
                                *reinterpret_cast<double2*>(P + tchunk(tile, 4 * (4 * h + r) + q)) = make_double2(acc[r][2 * q], acc[r][2 * q + 1]);
                    }
                }
                __syncthreads();
                BPROF(8);
            }  // chunks

            // ---- state update + quaternion reset --------------------------------------
            if constexpr (SINGLE)
                for (int i = tid; i < n; i += NT) p.dx_out[i] = s_delta[i];
            // No barrier here: the serial camera reset (two normalisations) runs on the last thread while the others
            // already fetch and associate the next frame's observations (neither touches the state or delta); the
            // barrier after the association orders it before the next model evaluation / predict.
            if (tid == NT - 1) camera_reset(s_state, s_delta, p.as_written);
            for (int l = tid; l < nlm; l += NT)
                landmark_update<KIND>(s_state + CAM + D * l, s_delta + CAM + D * l);
            BPROF(9);
        }  // frames
        __syncthreads();

        // ---- write back -----------------------------------------------------------------
        for (int i = tid; i < n; i += NT) g_state[i] = s_state[i];
        if constexpr (SINGLE) {
            if (p.one_state_host) {   // the caller reads the pose from here after an event wait: no copy operation
                for (int i = tid; i < n; i += NT) p.one_state_host[i] = s_state[i];
                if (tid == 0) *p.one_status_host = s_misc[1];
            }
            for (int i = tid >> 5; i < n; i += NT / 32)
                for (int j = tid & 31; j <= i; j += 32) g_P[(long long)i * p.ld + j] = s_P[tidx(i, j)];
            if (tid == 0 && s_misc[1] != 0) p.status[0] = s_misc[1];
        } else {
            for (int i = tid; i < nlm; i += NT) g_tab[i] = s_tab[i];
            for (int i = tid; i < n; i += NT)
                for (int j = 0; j <= i; ++j) g_P[tri(i) + j] = s_P[tidx(i, j)];
            if (tid == 0) {
                p.nlm[seq] = nlm;
                if (s_misc[1] != 0) p.status[seq] = s_misc[1];
            }
        }
        __syncthreads();
    }
    if constexpr (PROF)
        if (prof_on)
            for (int i = 0; i < 10; ++i) p.prof[i] += (unsigned long long)prof_acc[i];
#undef BPROF
}

// Constructor state of every sequence: state = pose0, P = init_cam * I10 (EKF:46-48), empty map.
__global__ void k_batch_init(double* __restrict__ states, double* __restrict__ covs, int* __restrict__ nlm,
                             int* __restrict__ status, long long state_stride, long long cov_stride,
                             int n_seq, const double* __restrict__ pose0, double init_cam, int tiled) {
    const int seq = blockIdx.x;
    if (seq >= n_seq) return;
    double* st = states + (long long)seq * state_stride;
    double* cv = covs + (long long)seq * cov_stride;
    for (int i = threadIdx.x; i < CAM; i += blockDim.x) st[i] = pose0[i];
    if (tiled) {   // wide engine: 8 x 8 tiles (wide_kernels.cuh), the whole buffer was zeroed by the caller
        for (int i = threadIdx.x; i < CAM; i += blockDim.x)
            cv[((long long)(tri(i >> 3) + (i >> 3)) << 6) + ((i & 7) << 3) + (i & 7)] = init_cam;
        if (threadIdx.x == 0) { nlm[seq] = 0; status[seq] = 0; }
        return;
    }
    for (int e = threadIdx.x; e < tri(CAM); e += blockDim.x) cv[e] = 0.0;
    __syncthreads();
    for (int i = threadIdx.x; i < CAM; i += blockDim.x) cv[tri(i) + i] = init_cam;
    if (threadIdx.x == 0) { nlm[seq] = 0; status[seq] = 0; }
}

// Expand one packed covariance into a dense row-major n x n matrix.
__global__ void k_unpack(const double* __restrict__ packed, int n, double* __restrict__ dense) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)n * n) return;
    const int i = (int)(e / n), j = (int)(e % n);
    dense[e] = psym(packed, i, j);
}

}  // namespace asb
