// Kernels of the single-filter engine: one filter whose covariance P lives in HBM
// (BASELINE config C3: 2000 markers, N = 6010, P = 289 MB; also C1 and any N).
//
// Per observe() (extended_kalman_filter.py:58-82):
//   k_prep    augmentation (add_marker, EKF:239-290) + predict (EKF:95-105) + h / H / nu
//             (parse_poses / landmark_dh, EKF:158-237) — H is never materialised: only
//             the r x (10+d) non-zero block of each observation exists.
//   k_gather  G = H P              (m x N)  from the 10 + k d touched rows of P
//   k_chol    S = G H^T + R, S = L L^T, y = L^-1 nu          (EKF:124-129)
//   k_solve   V = L^-1 G, dx = V^T y  (= K nu, EKF:130-131)
//   k_apply   state update + multiplicative quaternion reset  (EKF:134-152, ROT:143-177)
//   k_sweep   P <- P - V^T V  (= (I - K H) P, EKF:155-156), the HBM-bound rank-m sweep
#pragma once
#include <cstdint>

#include "../../include/aruco_slam_b200.h"
#include "model.cuh"
#include "sweep.cuh"

namespace asb {

constexpr int PAD = 128;       // leading dimension / row padding of P and V (doubles)
constexpr int M_MAX = 96;      // max (padded) measurement rows per chunk (sweep panels: 96 x 268 x 8 B = 206 KB)

// ------------------------------------------------------------------------------------
// k_prep
// ------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256)
k_prep(double* __restrict__ state, double* __restrict__ P, int ld, int n0, int n1, Noise nz,
       const int* __restrict__ idx, const double* __restrict__ poses,
       const unsigned char* __restrict__ isnew, int k, double* __restrict__ blocks,
       double* __restrict__ nu, int do_predict) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D;
    // ---- CTA 0: landmark initialisation, then the measurement model -------------------
    if (blockIdx.x == 0) {
        __shared__ double cam[CAM];
        if (threadIdx.x < CAM) cam[threadIdx.x] = state[threadIdx.x];
        __syncthreads();
        for (int o = threadIdx.x; o < k; o += blockDim.x) {
            if (isnew[o]) {
                double lm[D];
                new_landmark<KIND>(cam, poses + 6 * o, lm);
                double* dst = state + CAM + D * idx[o];
#pragma unroll
                for (int c = 0; c < D; ++c) dst[c] = lm[c];
            }
        }
        __syncthreads();
        for (int o = threadIdx.x; o < k; o += blockDim.x) {
            double lm[D];
            const double* src = state + CAM + D * idx[o];
#pragma unroll
            for (int c = 0; c < D; ++c) lm[c] = src[c];
            double lnu[R], blk[R * W];
            observation_model<KIND>(cam, lm, poses + 6 * o, lnu, blk);
#pragma unroll
            for (int r = 0; r < R; ++r) nu[R * o + r] = lnu[r];
            double* bo = blocks + (size_t)o * R * W;
#pragma unroll
            for (int c = 0; c < R * W; ++c) bo[c] = blk[c];
        }
    }
    // ---- all CTAs: covariance augmentation (block-diagonal growth) and P += diag(Q) ---
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    if (do_predict) {
        for (long long i = tid; i < n0; i += nth) P[i * ld + i] += process_noise(nz, (int)i);
    }
    const int dn = n1 - n0;
    if (dn > 0) {
        // new rows [n0,n1) x [0,n1)
        const long long cnt_rows = (long long)dn * n1;
        for (long long e = tid; e < cnt_rows; e += nth) {
            const int i = n0 + (int)(e / n1), j = (int)(e % n1);
            P[(long long)i * ld + j] = (i == j) ? nz.init_lm + (do_predict ? nz.q_lm : 0.0) : 0.0;
        }
        // new columns [0,n0) x [n0,n1)
        const long long cnt_cols = (long long)n0 * dn;
        for (long long e = tid; e < cnt_cols; e += nth) {
            const int i = (int)(e / dn), j = n0 + (int)(e % dn);
            P[(long long)i * ld + j] = 0.0;
        }
    }
}

// add_marker with an explicit diagonal (EKF:280-281): d values on the new diagonal block.
__global__ void k_set_diag(double* P, int ld, int start, int d, const double* vals) {
    int c = threadIdx.x;
    if (c < d) P[(long long)(start + c) * ld + start + c] = vals[c];
}

// ------------------------------------------------------------------------------------
// k_gather: G[(o,a)][j] = sum_c Hc_o[a][c] P[c][j] + sum_c Hl_o[a][c] P[10 + D idx_o + c][j]
// grid (ceil(N/128), ceil(kc/GATHER_OBS)); rows of P are read coalesced along j.
// ------------------------------------------------------------------------------------
constexpr int GATHER_OBS = 4;

template <int KIND>
__global__ void __launch_bounds__(128)
k_gather(const double* __restrict__ P, int ld, int n, const int* __restrict__ idx,
         const double* __restrict__ blocks, int kc, int m4, double* __restrict__ G) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D;
    __shared__ double sb[GATHER_OBS * R * W];
    __shared__ int sidx[GATHER_OBS];
    const int o0 = blockIdx.y * GATHER_OBS;
    const int no = min(GATHER_OBS, kc - o0);
    for (int e = threadIdx.x; e < no * R * W; e += blockDim.x) sb[e] = blocks[(size_t)o0 * R * W + e];
    if (threadIdx.x < no) sidx[threadIdx.x] = idx[o0 + threadIdx.x];
    __syncthreads();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ld) return;
    const bool live = j < n;
    double pc[CAM];
#pragma unroll
    for (int c = 0; c < CAM; ++c) pc[c] = live ? P[(long long)c * ld + j] : 0.0;
    for (int o = 0; o < no; ++o) {
        double pl[D];
        const long long base = (long long)(CAM + D * sidx[o]) * ld + j;
#pragma unroll
        for (int c = 0; c < D; ++c) pl[c] = live ? P[base + (long long)c * ld] : 0.0;
        const double* b = sb + o * R * W;
#pragma unroll
        for (int a = 0; a < R; ++a) {
            double s = 0.0;
#pragma unroll
            for (int c = 0; c < CAM; ++c) s += b[a * W + c] * pc[c];
#pragma unroll
            for (int c = 0; c < D; ++c) s += b[a * W + CAM + c] * pl[c];
            G[(long long)(R * (o0 + o) + a) * ld + j] = s;
        }
    }
    // zero the padding rows m..m4 (last y-slice only)
    if (blockIdx.y == gridDim.y - 1)
        for (int a = R * kc; a < m4; ++a) G[(long long)a * ld + j] = 0.0;
}

// ------------------------------------------------------------------------------------
// k_chol: one CTA.  Builds the lower triangle of S = G H^T + R in shared memory with nu
// appended as an extra row, and runs a blocked (NB = rows per marker) Cholesky on it; the
// appended row comes out as y^T = (L^-1 nu)^T.  Each thread factors the NB x NB diagonal
// block redundantly in registers (no serial thread), then solves its own panel row.
// Outputs: L (m4 x m4, identity-padded), 1/diag(L), y, status flag.
// ------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256)
k_chol(const double* __restrict__ G, int ld, const int* __restrict__ idx,
       const double* __restrict__ blocks, const double* __restrict__ nu, int kc, int m4,
       double r_var, const double* __restrict__ delta, int use_delta,
       double* __restrict__ Lg, double* __restrict__ invd, double* __restrict__ y,
       int* __restrict__ status) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D, NB = R;
    extern __shared__ double sm[];
    const int m = R * kc;
    const int lds = (m + 1) | 1;
    double* A = sm;  // (m+1) x lds
    const int tid = threadIdx.x, nt = blockDim.x;

    // build lower triangle + appended innovation row
    const int total = (m + 1) * m;
    for (int e = tid; e < total; e += nt) {
        const int a = e / m, b = e % m;
        if (a < m && b > a) continue;
        const int ob = b / R, ib = b % R;
        const double* hb = blocks + (size_t)ob * R * W + ib * W;
        double s;
        if (a < m) {
            const double* g = G + (long long)a * ld;
            const double* gl = g + CAM + D * idx[ob];
            s = (a == b) ? r_var : 0.0;
#pragma unroll
            for (int c = 0; c < CAM; ++c) s += g[c] * hb[c];
#pragma unroll
            for (int c = 0; c < D; ++c) s += gl[c] * hb[CAM + c];
        } else {
            s = nu[b];
            if (use_delta) {  // chunked update: nu' = nu - H delta
                const double* dl = delta + CAM + D * idx[ob];
#pragma unroll
                for (int c = 0; c < CAM; ++c) s -= hb[c] * delta[c];
#pragma unroll
                for (int c = 0; c < D; ++c) s -= hb[CAM + c] * dl[c];
            }
        }
        A[a * lds + b] = s;
    }
    __syncthreads();

    bool bad = false;
    for (int jb = 0; jb < m; jb += NB) {
        // diagonal block, factored redundantly by every thread
        double dblk[NB][NB], dinv[NB];
#pragma unroll
        for (int r = 0; r < NB; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) dblk[r][c] = A[(jb + r) * lds + jb + c];
#pragma unroll
        for (int c = 0; c < NB; ++c) {
            double p = dblk[c][c];
#pragma unroll
            for (int t = 0; t < c; ++t) p -= dblk[c][t] * dblk[c][t];
            if (!(p > 0.0) || !(p < 1e300)) bad = true;
            const double rs = rsqrt(p);
            dinv[c] = rs;
            dblk[c][c] = p * rs;
#pragma unroll
            for (int r = c + 1; r < NB; ++r) {
                double v = dblk[r][c];
#pragma unroll
                for (int t = 0; t < c; ++t) v -= dblk[r][t] * dblk[c][t];
                dblk[r][c] = v * rs;
            }
        }
        // panel: rows below the block (including the appended row m): x L_D^T = a
        for (int i = jb + NB + tid; i <= m; i += nt) {
            double x[NB];
#pragma unroll
            for (int c = 0; c < NB; ++c) {
                double v = A[i * lds + jb + c];
#pragma unroll
                for (int t = 0; t < c; ++t) v -= x[t] * dblk[c][t];
                x[c] = v * dinv[c];
            }
#pragma unroll
            for (int c = 0; c < NB; ++c) A[i * lds + jb + c] = x[c];
        }
#pragma unroll
        for (int r = 0; r < NB; ++r) {
            if (tid == r) {
#pragma unroll
                for (int c = 0; c <= r; ++c) A[(jb + r) * lds + jb + c] = dblk[r][c];
                invd[jb + r] = dinv[r];
            }
        }
        __syncthreads();
        // trailing update of the lower triangle (and the appended row)
        const int t0 = jb + NB;
        const int rem = m - t0;  // trailing size (rows t0..m-1), plus appended row m
        if (rem > 0) {
            const int cnt = (rem + 1) * rem;
            for (int e = tid; e < cnt; e += nt) {
                const int i = t0 + e / rem, kk = t0 + e % rem;
                if (i < m && kk > i) continue;
                double s = A[i * lds + kk];
#pragma unroll
                for (int c = 0; c < NB; ++c) s -= A[i * lds + jb + c] * A[kk * lds + jb + c];
                A[i * lds + kk] = s;
            }
        }
        __syncthreads();
    }
    if (bad) atomicExch(status, ASB_ERR_NOT_SPD);

    // write out L (identity padded to m4), y
    for (int e = tid; e < m4 * m4; e += nt) {
        const int a = e / m4, b = e % m4;
        double v = 0.0;
        if (a < m && b <= a) v = A[a * lds + b];
        else if (a >= m && a == b) v = 1.0;
        Lg[e] = v;
    }
    for (int a = tid; a < m4; a += nt) {
        y[a] = a < m ? A[m * lds + a] : 0.0;
        if (a >= m) invd[a] = 1.0;
    }
}

// ------------------------------------------------------------------------------------
// k_solve: V = L^-1 G, written in the tiled panel layout (blocked forward substitution, 4 rows at a time, one
// thread per column) and dx (+)= V^T y.
// ------------------------------------------------------------------------------------
constexpr int SOLVE_COLS = 64;

__global__ void __launch_bounds__(SOLVE_COLS)
k_solve(const double* __restrict__ G, double* __restrict__ Vt, int ld, int n, int m4,
        const double* __restrict__ Lg, const double* __restrict__ invd, const double* __restrict__ y,
        double* __restrict__ dx, int accumulate) {
    extern __shared__ double sm[];
    double* Ls = sm;                       // m4 x m4
    double* Vs = Ls + m4 * m4;             // m4 x SOLVE_COLS
    double* ds = Vs + m4 * SOLVE_COLS;     // m4
    double* ys = ds + m4;                  // m4
    const int tid = threadIdx.x;
    for (int e = tid; e < m4 * m4; e += SOLVE_COLS) Ls[e] = Lg[e];
    for (int e = tid; e < m4; e += SOLVE_COLS) { ds[e] = invd[e]; ys[e] = y[e]; }
    __syncthreads();
    const int j = blockIdx.x * SOLVE_COLS + tid;
    if (j >= ld) return;
    double acc = 0.0;
    for (int a0 = 0; a0 < m4; a0 += 4) {
        double v0 = G[(long long)(a0 + 0) * ld + j], v1 = G[(long long)(a0 + 1) * ld + j];
        double v2 = G[(long long)(a0 + 2) * ld + j], v3 = G[(long long)(a0 + 3) * ld + j];
        const double* l0 = Ls + (a0 + 0) * m4;
        const double* l1 = Ls + (a0 + 1) * m4;
        const double* l2 = Ls + (a0 + 2) * m4;
        const double* l3 = Ls + (a0 + 3) * m4;
        for (int b = 0; b < a0; ++b) {
            const double vb = Vs[b * SOLVE_COLS + tid];
            v0 -= l0[b] * vb; v1 -= l1[b] * vb; v2 -= l2[b] * vb; v3 -= l3[b] * vb;
        }
        v0 *= ds[a0];
        v1 = (v1 - l1[a0] * v0) * ds[a0 + 1];
        v2 = (v2 - l2[a0] * v0 - l2[a0 + 1] * v1) * ds[a0 + 2];
        v3 = (v3 - l3[a0] * v0 - l3[a0 + 1] * v1 - l3[a0 + 2] * v2) * ds[a0 + 3];
        Vs[(a0 + 0) * SOLVE_COLS + tid] = v0; Vs[(a0 + 1) * SOLVE_COLS + tid] = v1;
        Vs[(a0 + 2) * SOLVE_COLS + tid] = v2; Vs[(a0 + 3) * SOLVE_COLS + tid] = v3;
        double* vt = Vt + v_index(m4, a0, j);  // tiled panel layout the sweep's bulk copies expect (sweep.cuh)
        vt[0] = v0; vt[V_LDB] = v1; vt[2 * V_LDB] = v2; vt[3 * V_LDB] = v3;
        acc += v0 * ys[a0] + v1 * ys[a0 + 1] + v2 * ys[a0 + 2] + v3 * ys[a0 + 3];
    }
    if (j < n) dx[j] = accumulate ? dx[j] + acc : acc;
}

// ------------------------------------------------------------------------------------
// k_apply: camera additive part + multiplicative quaternion reset, landmarks.
// ------------------------------------------------------------------------------------
template <int KIND>
__global__ void k_apply(double* __restrict__ state, const double* __restrict__ dx, int nlm,
                        int as_written) {
    constexpr int D = Dims<KIND>::D;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) {
        double cam[CAM], d[CAM];
#pragma unroll
        for (int c = 0; c < CAM; ++c) { cam[c] = state[c]; d[c] = dx[c]; }
        camera_reset(cam, d, as_written);
#pragma unroll
        for (int c = 0; c < CAM; ++c) state[c] = cam[c];
    }
    if (t < nlm) {
        double lm[D], d[D];
        double* s = state + CAM + D * t;
        const double* dd = dx + CAM + D * t;
#pragma unroll
        for (int c = 0; c < D; ++c) { lm[c] = s[c]; d[c] = dd[c]; }
        landmark_update<KIND>(lm, d);
#pragma unroll
        for (int c = 0; c < D; ++c) s[c] = lm[c];
    }
}

}  // namespace asb
