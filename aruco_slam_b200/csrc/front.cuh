// Shared pieces of the single-filter front end (pipe.cuh: k_f1 / k_f2): argument block, phase stamps,
// chunk limits, cp.async helpers and the blocked Cholesky of [S; nu^T; I] whose appended rows yield
// y = L^-1 nu and L^-1 (EKF:124-131).  (The one-kernel front end k_front that these were written for was
// replaced by the pipelined pair and removed in round 2.)
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

}  // namespace asb
