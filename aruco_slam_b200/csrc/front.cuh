// k_front: everything of one update chunk that precedes the covariance sweep, in ONE launch.
//
//   measurement model      nu = z - h, Jacobian blocks     (parse_poses / landmark_dh, EKF:158-237)
//   new-landmark states    add_marker                      (EKF:264-274, ROT:300-319)
//   covariance growth      P <- blockdiag(P, init_lm I)    (EKF:279-290)
//   predict                P + diag(Q), applied ON THE FLY (EKF:95-105; the sweep writes it back)
//   S = H P H^T + R = L L^T, y = L^-1 nu                   (EKF:124-129)
//   V = L^-1 H P, dx (+)= V^T y  (= K nu)                  (EKF:130-131)
//
// Grid: one CTA per 64-column block of the V panel.  Every CTA redundantly evaluates the
// model, gathers the |T| x |T| sub-matrix of P over the touched state indices T (camera
// + observed landmarks), builds S from it and factors it — S does not depend on the full
// G = H P, so nothing serialises on a single CTA and there is no inter-kernel hand-off of
// L / y through HBM.  All HBM reads of the kernel (P[T,T] and the |T| rows of this block's
// 64 columns) are issued up front as cp.async copies and fly while the model is evaluated.
// The Cholesky runs on S with nu and the identity appended as extra rows, which yields
// y = L^-1 nu and W^T = L^-T for free; V = W G is then a register-tiled product instead
// of a serial substitution.  Each CTA writes its block of the tiled panel (sweep.cuh), dx.
//
// Hazards: the kernel never writes a location another CTA reads.  P entries in the rows /
// columns being created by this frame's augmentation (index >= n0) are never read from
// memory (their values are known constants); P + diag(Q) is formed in registers and only
// written back by the sweep; the state is read-only here except for CTA 0 storing new
// landmarks, which every CTA recomputes instead of reading; dx ping-pongs between buffers.
#pragma once
#include "model.cuh"
#include "sweep.cuh"

namespace asb {

constexpr int FRONT_THREADS = 256;
constexpr int FRONT_LDG = V_BLK + 8;   // row stride of the CTA's G / V slice in shared memory

struct FrontArgs {
    double* state;               // read; CTA 0 stores new-landmark entries
    double* P;                   // read (+ augmentation stores on the first chunk)
    int ld, n0, n1;              // state dims before / after this frame's augmentation
    Noise nz;
    const int* idx;              // landmark index of each observation of the chunk
    const double* poses;         // (kc,6)
    const unsigned char* isnew;  // observation creates its landmark (first chunk semantics)
    int kc, m4;
    int first_chunk;             // do augmentation stores
    int add_q;                   // predict not yet folded into P: use P + diag(Q)
    int use_delta;               // nu' = nu - H dx_in   (chunks after the first)
    const double* dx_in;
    double* dx_out;
    double* Vr;                  // tiled panel in V_I column order (sweep.cuh)
    double* Vc;                  // tiled panel in V_J column order
    double* blocks_out;          // (kc, R, W) for asb_get_last_update, may be null
    double* nu_out;              // (R kc)
    int* status;
    int dbg;                     // experiments only (ASB_FRONT_DBG): 1 skip trailing, 2 skip panel, 4 skip diag, 8 skip extract
    unsigned long long* prof;    // phase timestamps of CTA 0 (globaltimer ns), 8 slots, may be null
};

__device__ __forceinline__ int ftri(int i) { return (i * (i + 1)) >> 1; }
__device__ __forceinline__ void front_clk(const FrontArgs& p, int slot, bool on) {
    if (p.prof && on && blockIdx.x == 0 && threadIdx.x == 0) p.prof[slot] = (unsigned long long)clock64();
}
__device__ __forceinline__ void front_stamp(const FrontArgs& p, int slot) {
    if (p.prof && blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.prof[slot] = t;
    }
}

// Chunk limits (rows per chunk m = R kc): shared memory holds PT (nT^2), GT (m x nT), the
// augmented S ((2m+1) x (m|1)), the staged rows (nT x 64) and the model blocks.
template <int KIND> struct FrontLimits;
template <> struct FrontLimits<0> { static constexpr int KC_MAX = 21; };   // m = 63 (<= 8 column tiles, 16 row tiles)
template <> struct FrontLimits<1> { static constexpr int KC_MAX = 8; };    // m = 56, 212 KB

template <int KIND>
struct FrontSmem {
    static constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D;
    static __host__ __device__ size_t region1(int kc, int m4) {   // PT, later the G slice
        const size_t nT = CAM + D * kc, pt = nT * nT;
        (void)m4;
        return (pt + 1) & ~(size_t)1;   // keep 16-byte alignment behind it
    }
    static __host__ __device__ size_t s_rows(int kc) {            // S with nu appended, padded to even
        const size_t m = R * kc;
        return (size_t)128 * (m | 1);   // rows 0..m: [S; nu^T], rows 64..127: G^T of the block
    }
    static __host__ __device__ size_t gt_rows(int kc) {           // GT = H PT, row stride nT|1
        const size_t m = R * kc, nT = CAM + D * kc;
        return (m * (nT | 1) + 1) & ~(size_t)1;
    }
    static __host__ __device__ size_t region3(int kc, int m4) {   // staged rows, later the V slice
        (void)m4;
        return (size_t)(CAM + D * kc) * V_BLK;
    }
    static __host__ __device__ size_t bytes(int kc, int m4) {
        const int m = R * kc, nT = CAM + D * kc, lds = m | 1;
        size_t dbl = region1(kc, m4) + s_rows(kc) + gt_rows(kc) + region3(kc, m4) + (size_t)kc * R * W +
                     (size_t)kc * D + 2 * (size_t)m4 + 16;
        return dbl * sizeof(double) + (size_t)(nT + 2 * kc + 8) * sizeof(int);
    }
};

__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// Blocked right-looking Cholesky of the augmented matrix [S; nu^T; X^T] held in s_S (row stride lds):
// rows 0..m-1: S -> L;  row m: nu -> y^T = (L^-1 nu)^T;  rows GROW+jj (jj < 64): X[:,jj]^T -> (L^-1 X[:,jj])^T.
// 256 threads.  Returns true if a pivot was not positive (S not SPD).
constexpr int GROW = 64;
template <int NB>
__device__ __forceinline__ bool front_cholesky(double* s_S, int lds, int m, int dbg) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- phase 6: blocked right-looking Cholesky of the augmented matrix [S; nu^T; G^T] -------
    // rows 0..m-1: S -> L;  row m: nu -> y^T = (L^-1 nu)^T;  rows GROW+jj: G[:,jj]^T -> V[:,jj]^T.
    // NB = rows per marker.  The trailing matrix lives in REGISTERS as 8 x 8 DMMA accumulator
    // tiles — warp w owns row tile w of the S part and row tile w of the G part, all 8 column
    // tiles — and is updated with mma.sync m8n8k4 (k = NB padded to 4); shared memory only
    // carries the current block column, written out before the panel solve.  (With the trailing
    // matrix in shared memory the update was shared-memory-bandwidth-bound.)
    const int g8 = lane >> 2, t4 = lane & 3;
    const int r0 = 8 * warp + g8, r1 = GROW + 8 * warp + g8;   // this lane's row in the S / G tile
    double C[2][8][2];
#pragma unroll
    for (int ct = 0; ct < 8; ++ct)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int c = 8 * ct + 2 * t4 + e;
            C[0][ct][e] = (c < m && r0 <= m && (r0 == m || c <= r0)) ? s_S[r0 * lds + c] : 0.0;
            C[1][ct][e] = c < m ? s_S[r1 * lds + c] : 0.0;
        }
    bool bad = false;
    for (int jb = 0; jb < m; jb += NB) {
        const int t0 = jb + NB;
        if (jb > 0 && !(dbg & 8)) {  // block column jb .. jb+NB-1 of the trailing matrix: registers -> shared memory
            const int ct_lo = jb >> 3, ct_hi = (jb + NB - 1) >> 3;   // the (at most two) column tiles involved
#pragma unroll
            for (int ct = 0; ct < 8; ++ct) {
                if (ct < ct_lo || ct > ct_hi) continue;               // warp-uniform: registers stay statically indexed
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int c = 8 * ct + 2 * t4 + e;
                    if (c >= jb && c < t0) {
                        if (r0 >= jb && r0 <= m) s_S[r0 * lds + c] = C[0][ct][e];
                        s_S[r1 * lds + c] = C[1][ct][e];
                    }
                }
            }
            __syncthreads();
        }
        if (tid < 128) {
            // warps 0-3: factor the NB x NB diagonal block (redundantly per thread, in registers),
            // then one row of the panel each: S rows t0..m (thread < 64), G rows (thread >= 64)
            double dblk[NB][NB], dinv[NB];
#pragma unroll
            for (int r = 0; r < NB; ++r)
#pragma unroll
                for (int c = 0; c <= r; ++c) dblk[r][c] = s_S[(jb + r) * lds + jb + c];
            if (dbg & 4) {
#pragma unroll
                for (int c = 0; c < NB; ++c) dinv[c] = 1.0;
            } else if constexpr (NB == 3) {
                // closed form with the three reciprocal square roots independent of each other
                // (leading principal minors): the serial pivot chain is the critical path of the step
                const double a00 = dblk[0][0], a10 = dblk[1][0], a11 = dblk[1][1];
                const double a20 = dblk[2][0], a21 = dblk[2][1], a22 = dblk[2][2];
                const double det2 = a00 * a11 - a10 * a10;
                const double q3 = a00 * a21 * a21 - 2.0 * a10 * a20 * a21 + a11 * a20 * a20;
                const double det3 = a22 * det2 - q3;
                if (!(a00 > 0.0) || !(det2 > 0.0) || !(det3 > 0.0) || !(a00 < 1e300) || !(det3 < 1e300)) bad = true;
                const double rs0 = rsqrt(a00), rd2 = rsqrt(det2), rd3 = rsqrt(det3);
                const double l00 = a00 * rs0, sd2 = det2 * rd2;   // sqrt(a00), sqrt(det2)
                dinv[0] = rs0;
                dinv[1] = rd2 * l00;          // 1 / sqrt(det2 / a00)
                dinv[2] = rd3 * sd2;          // 1 / sqrt(det3 / det2)
                dblk[1][0] = a10 * rs0;
                dblk[2][0] = a20 * rs0;
                dblk[2][1] = (a21 - dblk[2][0] * dblk[1][0]) * dinv[1];
            } else {
#pragma unroll
                for (int c = 0; c < NB; ++c) {
                    double pv = dblk[c][c];
#pragma unroll
                    for (int t = 0; t < c; ++t) pv -= dblk[c][t] * dblk[c][t];
                    if (!(pv > 0.0) || !(pv < 1e300)) bad = true;
                    const double rs = rsqrt(pv);
                    dinv[c] = rs;
#pragma unroll
                    for (int r = c + 1; r < NB; ++r) {
                        double v = dblk[r][c];
#pragma unroll
                        for (int t = 0; t < c; ++t) v -= dblk[r][t] * dblk[c][t];
                        dblk[r][c] = v * rs;
                    }
                }
            }
            // (the factored diagonal block itself is never read again: only y and V are used)
            const int i = tid < 64 ? t0 + tid : GROW + (tid - 64);
            if ((tid >= 64 || i <= m) && !(dbg & 2)) {
                double x[NB];
#pragma unroll
                for (int c = 0; c < NB; ++c) x[c] = s_S[i * lds + jb + c];
#pragma unroll
                for (int c = 0; c < NB; ++c) {
                    double v = x[c];
#pragma unroll
                    for (int t = 0; t < c; ++t) v -= x[t] * dblk[c][t];
                    x[c] = v * dinv[c];
                }
#pragma unroll
                for (int c = 0; c < NB; ++c) s_S[i * lds + jb + c] = x[c];
            }
        }
        __syncthreads();
        // trailing update C -= L_panel L_panel^T on the register tiles (FP64 tensor pipe)
        if (t0 < m && !(dbg & 1)) {
            constexpr int KS = (NB + 3) / 4;
            const bool act0 = (8 * warp + 7 >= t0) && (8 * warp <= m);   // warp-uniform
            double a0[KS], a1[KS];
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                const int t = 4 * ks + t4;
                a1[ks] = t < NB ? -s_S[r1 * lds + jb + t] : 0.0;
                a0[ks] = (act0 && t < NB && r0 >= t0 && r0 <= m) ? -s_S[r0 * lds + jb + t] : 0.0;
            }
            const int ct0 = t0 >> 3;
#pragma unroll
            for (int ct = 0; ct < 8; ++ct) {
                if (ct < ct0 || 8 * ct >= m) continue;     // no active column in this tile
                const int kk = 8 * ct + g8;
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) {
                    const int t = 4 * ks + t4;
                    const double bv = (t < NB && kk >= t0 && kk < m) ? s_S[kk * lds + jb + t] : 0.0;
                    dmma884(C[1][ct][0], C[1][ct][1], a1[ks], bv);
                    if (act0 && ct <= warp) dmma884(C[0][ct][0], C[0][ct][1], a0[ks], bv);
                }
            }
        }
    }
    __syncthreads();
    return bad;
}

template <int KIND>
__global__ void __launch_bounds__(FRONT_THREADS, 1)
k_front(FrontArgs p) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D, NB = R;
    constexpr int NT = FRONT_THREADS;
    extern __shared__ __align__(16) unsigned char fsm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kc = p.kc, m = R * kc, m4 = p.m4, nT = CAM + D * kc;
    const int lds = m | 1;
    const int ld = p.ld, n0 = p.n0, n1 = p.n1;
    const int j0 = blockIdx.x * V_BLK;

    // ---- shared memory carve-up ---------------------------------------------------------
    double* s_PT = reinterpret_cast<double*>(fsm);                       // nT x nT (dense)
    double* s_S = s_PT + FrontSmem<KIND>::region1(kc, m4);               // 128 x lds: [S; nu^T] and G^T at row 64
    double* s_GT = s_S + FrontSmem<KIND>::s_rows(kc);                    // m x ldt
    double* s_stage = s_GT + FrontSmem<KIND>::gt_rows(kc);                           // nT x 64: P_eff[T, this block's columns]
    double* s_blk = s_stage + FrontSmem<KIND>::region3(kc, m4);          // kc x R x W
    double* s_lm = s_blk + (size_t)kc * R * W;                           // kc x D
    double* s_dinv = s_lm + (size_t)kc * D;                              // m4
    double* s_y = s_dinv + m4;                                           // m4
    double* s_cam = s_y + m4;                                            // 10 (+6 pad)
    int* s_T = reinterpret_cast<int*>(s_cam + 16);                       // nT touched state indices
    int* s_idx = s_T + nT;                                               // kc
    int* s_new = s_idx + kc;                                             // kc

    front_stamp(p, 0);
    // ---- phase 0: touched index set T ------------------------------------------------------
    if (tid < CAM) { s_cam[tid] = p.state[tid]; s_T[tid] = tid; }
    for (int e = tid; e < D * kc; e += NT) {
        const int o = e / D, c = e - o * D;
        const int ix = p.idx[o];
        s_T[CAM + e] = CAM + D * ix + c;
        if (c == 0) { s_idx[o] = ix; s_new[o] = p.isnew ? p.isnew[o] : 0; }
    }
    __syncthreads();

    // ---- phase 1: every HBM read of the kernel, asynchronously -----------------------------
    // P entries in rows / columns >= n0 (this frame's augmentation) are known constants and
    // are never read from memory.  diag(Q) is added after the copies land (phase 3).
    for (int a = warp; a < nT; a += NT / 32) {
        const int i = s_T[a];
        const double* prow = p.P + (long long)i * ld;
        for (int b = lane; b < nT; b += 32) {  // PT[a][b] = P[T[a]][T[b]]
            const int j = s_T[b];
            if (i < n0 && j < n0) cp_async8(s_PT + a * nT + b, prow + j);
            else s_PT[a * nT + b] = (i == j) ? p.nz.init_lm : 0.0;
        }
        const int j = j0 + 2 * lane;          // stage[a][0..64) = P[T[a]][j0..j0+64)
        double* dst = s_stage + a * V_BLK + 2 * lane;
        if (i < n0 && j + 1 < n0) {
            cp_async16(dst, prow + j);
        } else {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (i < n0 && j + q < n0) cp_async8(dst + q, prow + j + q);
                else dst[q] = (i == j + q && i < n1) ? p.nz.init_lm : 0.0;
            }
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    front_stamp(p, 1);

    // ---- phase 2 (while the copies fly): measurement model, 
    for (int o = tid; o < kc; o += NT) {
        double lm[D];
        const int li = CAM + D * s_idx[o];
        if (s_new[o]) {
            new_landmark<KIND>(s_cam, p.poses + 6 * o, lm);
            if (blockIdx.x == 0) {
#pragma unroll
                for (int c = 0; c < D; ++c) p.state[li + c] = lm[c];
            }
        } else {
#pragma unroll
            for (int c = 0; c < D; ++c) lm[c] = p.state[li + c];
        }
#pragma unroll
        for (int c = 0; c < D; ++c) s_lm[o * D + c] = lm[c];
    }
    __syncthreads();
    for (int o = tid; o < kc; o += NT) {
        double lm[D];
        int src = o;
        if (!s_new[o])  // a repeated id whose landmark was created earlier in this same chunk
            for (int o2 = 0; o2 < kc; ++o2)
                if (s_new[o2] && s_idx[o2] == s_idx[o]) { src = o2; break; }
#pragma unroll
        for (int c = 0; c < D; ++c) lm[c] = s_lm[src * D + c];
        double lnu[R], blk[R * W];
        observation_model<KIND>(s_cam, lm, p.poses + 6 * o, lnu, blk);
        if (blockIdx.x == 0 && p.nu_out) {
#pragma unroll
            for (int r = 0; r < R; ++r) p.nu_out[R * o + r] = lnu[r];
#pragma unroll
            for (int c = 0; c < R * W; ++c) p.blocks_out[(size_t)o * R * W + c] = blk[c];
        }
        if (p.use_delta) {  // chunked update: nu' = nu - H delta
            const int li = CAM + D * s_idx[o];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                double s = lnu[r];
#pragma unroll
                for (int c = 0; c < CAM; ++c) s -= blk[r * W + c] * p.dx_in[c];
#pragma unroll
                for (int c = 0; c < D; ++c) s -= blk[r * W + CAM + c] * p.dx_in[li + c];
                lnu[r] = s;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) s_S[m * lds + R * o + r] = lnu[r];
#pragma unroll
        for (int c = 0; c < R * W; ++c) s_blk[o * R * W + c] = blk[c];
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    front_stamp(p, 2);

    // ---- phase 3: predict folded in — diag(Q) on the staged copies --------------------------
    if (p.add_q) {
        for (int a = tid; a < nT; a += NT) {
            const int i = s_T[a];
            const double q = process_noise(p.nz, i);
            s_PT[a * nT + a] += q;
            if (i >= j0 && i < j0 + V_BLK) s_stage[a * V_BLK + (i - j0)] += q;
        }
        for (int e = tid; e < kc * kc; e += NT) {   // a repeated id: the same state index twice in T
            const int o1 = e / kc, o2 = e - o1 * kc;
            if (o1 != o2 && s_idx[o1] == s_idx[o2]) {
#pragma unroll
                for (int c = 0; c < D; ++c)
                    s_PT[(CAM + D * o1 + c) * nT + CAM + D * o2 + c] += p.nz.q_lm;
            }
        }
        __syncthreads();
    }

    // ---- phase 4: S = H PT H^T + R in two small products with warp-uniform observations, so
    //      the Jacobian block of the observation sits in registers and lanes read consecutive
    //      columns (stage a) / rows (stage b) of shared memory.
    const int ldt = nT | 1;
    for (int o = warp; o < kc; o += NT / 32) {   // (a) GT[rows of o][tb] = H_o PT[T_o, tb]
        double hb[R * W];
#pragma unroll
        for (int c = 0; c < R * W; ++c) hb[c] = s_blk[(size_t)o * R * W + c];
        for (int tb = lane; tb < nT; tb += 32) {
            double acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.0;
#pragma unroll
            for (int c = 0; c < W; ++c) {
                const double pv = s_PT[(c < CAM ? c : CAM + D * o + (c - CAM)) * nT + tb];
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] += hb[r * W + c] * pv;
            }
#pragma unroll
            for (int r = 0; r < R; ++r) s_GT[(R * o + r) * ldt + tb] = acc[r];
        }
    }
    __syncthreads();
    for (int ob = warp; ob < kc; ob += NT / 32) {   // (b) S[a][rows of ob] = GT[a][T_ob] H_ob^T, a >= R ob
        double hb[R * W];
#pragma unroll
        for (int c = 0; c < R * W; ++c) hb[c] = s_blk[(size_t)ob * R * W + c];
        for (int a = R * ob + lane; a < m; a += 32) {
            const double* g = s_GT + a * ldt;
            double gv[W];
#pragma unroll
            for (int c = 0; c < W; ++c) gv[c] = g[c < CAM ? c : CAM + D * ob + (c - CAM)];
#pragma unroll
            for (int r2 = 0; r2 < R; ++r2) {
                const int b = R * ob + r2;
                double sv = (a == b) ? p.nz.r : 0.0;
#pragma unroll
                for (int c = 0; c < W; ++c) sv += gv[c] * hb[r2 * W + c];
                if (b <= a) s_S[a * lds + b] = sv;
            }
        }
    }
    __syncthreads();
    front_stamp(p, 3);

    // ---- phase 5: G^T for this block's 64 columns (G = H P_eff[T, block]) appended to S as rows
    //      GROW..GROW+63: the appended-row trick of the Cholesky below turns row GROW+jj into
    //      (L^-1 G[:, jj])^T = V[:, jj]^T, so V needs no triangular solve of its own.
    for (int e = tid; e < kc * V_BLK; e += NT) {
        const int o = e / V_BLK, jj = e - o * V_BLK;
        const double* blk = s_blk + (size_t)o * R * W;
        const double* sl = s_stage + (size_t)(CAM + D * o) * V_BLK + jj;
        double acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.0;
#pragma unroll
        for (int c = 0; c < W; ++c) {
            const double pv = c < CAM ? s_stage[c * V_BLK + jj] : sl[(c - CAM) * V_BLK];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] += blk[r * W + c] * pv;
        }
#pragma unroll
        for (int r = 0; r < R; ++r) s_S[(GROW + jj) * lds + R * o + r] = acc[r];
    }
    __syncthreads();
    front_stamp(p, 4);

    // ---- phase 6: Cholesky of [S; nu^T; G^T] (front_cholesky above)
    const bool bad = front_cholesky<NB>(s_S, lds, m, p.dbg);
    if (bad && blockIdx.x == 0 && tid == 0) atomicExch(p.status, ASB_ERR_NOT_SPD);
    front_stamp(p, 5);
    front_stamp(p, 6);
    // dx (+)= V^T y, and this CTA's block of the tiled panel (sweep.cuh): m4 rows x V_LDB.
    // V[a][jj] = s_S[(GROW + jj) * lds + a], y[a] = s_S[m * lds + a]
    if (tid < V_BLK) {
        double dsum = 0.0;
        const double* vrow = s_S + (size_t)(GROW + tid) * lds;
        const double* yrow = s_S + (size_t)m * lds;
        for (int a = 0; a < m; ++a) dsum += vrow[a] * yrow[a];
        const int j = j0 + tid;
        if (j < n1) p.dx_out[j] = (p.use_delta ? p.dx_in[j] : 0.0) + dsum;
    }
    {
        // both column orders of the block (sweep.cuh): position of column jj is v_pos_row / v_pos_col
        double* vr = p.Vr + (size_t)blockIdx.x * m4 * V_LDB;
        double* vc = p.Vc + (size_t)blockIdx.x * m4 * V_LDB;
        for (int e = tid; e < m4 * V_BLK; e += NT) {
            const int a = e / V_BLK, pos = e - a * V_BLK;
            // inverse maps: which column lands at `pos`
            const int cr = (pos & ~31) + 4 * (pos & 7) + ((pos & 31) >> 3);
            const int cc = (pos & ~15) + 4 * ((pos & 7) >> 1) + 2 * ((pos & 15) >> 3) + (pos & 1);
            vr[a * V_LDB + pos] = a < m ? s_S[(GROW + cr) * lds + a] : 0.0;   // rows m..m4: identity padding of L
            vc[a * V_LDB + pos] = a < m ? s_S[(GROW + cc) * lds + a] : 0.0;
        }
    }

    // ---- phase 8: covariance growth for this block's columns (first chunk only) ----------
    if (p.first_chunk && n1 > n0) {
        const int ja = j0, jb = min(j0 + V_BLK, n1);
        if (ja < jb) {
            const int wcols = jb - ja;  // new rows [n0, n1) x [ja, jb)
            for (int e = tid; e < (n1 - n0) * wcols; e += NT) {
                const int i = n0 + e / wcols, j = ja + e % wcols;
                p.P[(long long)i * ld + j] = (i == j) ? p.nz.init_lm : 0.0;
            }
            const int jc = max(ja, n0);  // old rows [0, n0) x ([ja, jb) ∩ [n0, n1))
            if (jc < jb) {
                const int w2 = jb - jc;
                for (int e = tid; e < n0 * w2; e += NT) {
                    const int i = e / w2, j = jc + e % w2;
                    p.P[(long long)i * ld + j] = 0.0;
                }
            }
        }
    }
    front_stamp(p, 7);
}

}  // namespace asb
