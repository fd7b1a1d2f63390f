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

namespace asb {

constexpr int SW_TI = 128;     // sweep tile rows
constexpr int SW_TJ = 64;      // sweep tile cols
constexpr int SW_THREADS = 256;
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
// k_solve: V = L^-1 G in place (blocked forward substitution, 4 rows at a time, one
// thread per column) and dx (+)= V^T y.
// ------------------------------------------------------------------------------------
constexpr int SOLVE_COLS = 64;

__global__ void __launch_bounds__(SOLVE_COLS)
k_solve(double* __restrict__ G, int ld, int n, int m4, const double* __restrict__ Lg,
        const double* __restrict__ invd, const double* __restrict__ y, double* __restrict__ dx,
        int accumulate) {
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
        G[(long long)(a0 + 0) * ld + j] = v0; G[(long long)(a0 + 1) * ld + j] = v1;
        G[(long long)(a0 + 2) * ld + j] = v2; G[(long long)(a0 + 3) * ld + j] = v3;
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

// ------------------------------------------------------------------------------------
// k_sweep: P <- P - V^T V on lower-triangle tiles, mirrored into the upper triangle.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
// TMA bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// FP64 tensor-core MMA (SASS: DMMA.8x8x4): D(8x8) += A(8x4) B(4x8).  Lane (g = lane/4,
// t = lane%4) holds A[g][t], B[t][g] and the accumulator pair C[g][2t], C[g][2t+1].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// Tile (I, J) covers rows [128 I, 128 I + 128) x cols [64 J, 64 J + 64); the tiles with
// 64 J <= 128 I + 127 are enumerated t = I (I + 1) + J, J in [0, 2 I + 2).
//
// Persistent kernel, one CTA per SM, 8 warps = 4 (rows) x 2 (cols); a warp owns a 32 x 32
// block of the tile as 4 x 4 DMMA accumulator tiles.  Each CTA walks a contiguous chunk of
// the tile list (row-block major), so the 128-column panel VI stays in shared memory while
// the 64-column panel VJ is double-buffered: the bulk copies (TMA, SASS UBLKCP) of tile
// t+1's VJ fly during tile t's MMAs.  The P sub-tile goes global -> registers in the
// accumulator-fragment layout at the start of the tile and is only consumed after the
// k-loop (out = P - V_I^T V_J), so its HBM latency hides behind the tile's own math; the
// result is stored to (i, j) and mirrored to (j, i).
// Panel row strides (SW_LDI, SW_LDJ) are = 4 mod 16 doubles: the fragment loads
// (address = (4 kk + t) * stride + g) are then bank-conflict free.
constexpr int SW_LDI = SW_TI + 4;
constexpr int SW_LDJ = SW_TJ + 4;

__host__ __device__ inline size_t sweep_smem_bytes(int m4) {
    return (size_t)m4 * (SW_LDI + 2 * SW_LDJ) * sizeof(double);
}

__global__ void __launch_bounds__(SW_THREADS, 1)
k_sweep(double* __restrict__ P, int ld, const double* __restrict__ V, int m4, int ntiles) {
    extern __shared__ __align__(128) unsigned char smraw[];
    double* VI = reinterpret_cast<double*>(smraw);   // m4 x SW_LDI
    double* VJ0 = VI + (size_t)m4 * SW_LDI;          // 2 x (m4 x SW_LDJ)
    __shared__ __align__(8) uint64_t bar[3];         // [0] VI, [1] VJ buf 0, [2] VJ buf 1
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    const int wi = warp >> 1, wj = warp & 1;
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        mbar_init(&bar[2], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int t_begin = (int)((long long)ntiles * blockIdx.x / gridDim.x);
    const int t_end = (int)((long long)ntiles * (blockIdx.x + 1) / gridDim.x);
    if (t_begin >= t_end) return;
    int I = (int)((sqrt(4.0 * t_begin + 1.0) - 1.0) * 0.5);
    while ((I + 1) * (I + 2) <= t_begin) ++I;
    while (I * (I + 1) > t_begin) --I;
    int J = t_begin - I * (I + 1);

    auto load_vi = [&](int i0) {  // warp 0
        if (lane == 0) mbar_expect_tx(&bar[0], (uint32_t)(m4 * SW_TI * sizeof(double)));
        __syncwarp();
        for (int a = lane; a < m4; a += 32)
            bulk_g2s(VI + a * SW_LDI, V + (long long)a * ld + i0, SW_TI * sizeof(double), &bar[0]);
    };
    auto load_vj = [&](int buf, int j0) {  // warp 0
        double* dst = VJ0 + (size_t)buf * m4 * SW_LDJ;
        if (lane == 0) mbar_expect_tx(&bar[1 + buf], (uint32_t)(m4 * SW_TJ * sizeof(double)));
        __syncwarp();
        for (int a = lane; a < m4; a += 32)
            bulk_g2s(dst + a * SW_LDJ, V + (long long)a * ld + j0, SW_TJ * sizeof(double), &bar[1 + buf]);
    };

    uint32_t ph_i = 0, ph_j = 0;  // ph_j: bit b = phase of VJ buffer b
    bool vi_pending = true;
    if (warp == 0) {
        load_vi(I * SW_TI);
        load_vj(0, J * SW_TJ);
    }
    int buf = 0;
    for (int t = t_begin; t < t_end; ++t, buf ^= 1) {
        const int i0 = I * SW_TI, j0 = J * SW_TJ;
        // next tile
        int In = I, Jn = J + 1;
        if (Jn == 2 * I + 2) { In = I + 1; Jn = 0; }
        const bool has_next = t + 1 < t_end;
        if (warp == 0 && has_next) load_vj(buf ^ 1, Jn * SW_TJ);  // buffer freed by the last __syncthreads

        // P sub-tile -> registers (accumulator-fragment layout)
        const int ri = i0 + 32 * wi + g, cj = j0 + 32 * wj + 2 * t4;
        double2 pf[4][4];
#pragma unroll
        for (int rm = 0; rm < 4; ++rm)
#pragma unroll
            for (int rn = 0; rn < 4; ++rn)
                pf[rm][rn] = *reinterpret_cast<const double2*>(P + (long long)(ri + 8 * rm) * ld + cj + 8 * rn);

        double acc[4][4][2];
#pragma unroll
        for (int rm = 0; rm < 4; ++rm)
#pragma unroll
            for (int rn = 0; rn < 4; ++rn) acc[rm][rn][0] = acc[rm][rn][1] = 0.0;

        if (vi_pending) {
            mbar_wait(&bar[0], ph_i);
            ph_i ^= 1;
            vi_pending = false;
        }
        mbar_wait(&bar[1 + buf], (ph_j >> buf) & 1u);
        ph_j ^= 1u << buf;

        const double* va = VI + t4 * SW_LDI + 32 * wi + g;
        const double* vb = VJ0 + (size_t)buf * m4 * SW_LDJ + t4 * SW_LDJ + 32 * wj + g;
#pragma unroll 3
        for (int kk = 0; kk < m4; kk += 4) {
            double a[4], b[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                a[r] = va[kk * SW_LDI + 8 * r];
                b[r] = vb[kk * SW_LDJ + 8 * r];
            }
#pragma unroll
            for (int rm = 0; rm < 4; ++rm)
#pragma unroll
                for (int rn = 0; rn < 4; ++rn) dmma884(acc[rm][rn][0], acc[rm][rn][1], a[rm], b[rn]);
        }

        const bool interior = i0 >= j0 + SW_TJ;  // strictly below the diagonal
        if (interior) {
#pragma unroll
            for (int rm = 0; rm < 4; ++rm)
#pragma unroll
                for (int rn = 0; rn < 4; ++rn) {
                    const double o0 = pf[rm][rn].x - acc[rm][rn][0], o1 = pf[rm][rn].y - acc[rm][rn][1];
                    const int i = ri + 8 * rm, j = cj + 8 * rn;
                    *reinterpret_cast<double2*>(P + (long long)i * ld + j) = make_double2(o0, o1);
                    P[(long long)j * ld + i] = o0;
                    P[(long long)(j + 1) * ld + i] = o1;
                }
        } else {
#pragma unroll
            for (int rm = 0; rm < 4; ++rm)
#pragma unroll
                for (int rn = 0; rn < 4; ++rn) {
                    const double o[2] = {pf[rm][rn].x - acc[rm][rn][0], pf[rm][rn].y - acc[rm][rn][1]};
                    const int i = ri + 8 * rm;
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int j = cj + 8 * rn + e;
                        if (i >= j) {
                            P[(long long)i * ld + j] = o[e];
                            if (i > j) P[(long long)j * ld + i] = o[e];
                        }
                    }
                }
        }
        __syncthreads();  // every warp is done with VJ[buf] (and VI) before they are overwritten
        if (has_next && In != I) {
            if (warp == 0) load_vi(In * SW_TI);
            vi_pending = true;
        }
        I = In;
        J = Jn;
    }
}

}  // namespace asb
