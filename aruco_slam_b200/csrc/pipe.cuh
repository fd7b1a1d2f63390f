// Pipelined front end: k_front (front.cuh) split into the part that does not need the
// finished covariance (k_f1) and the part that does (k_f2), so that k_f1 of update t+1 runs
// WHILE the sweep of update t is still streaming the covariance through HBM.
//
//   main stream   ... k_f2(t) -> k_sweep(t): Pa -> Pb  -> k_f2(t+1) -> k_sweep(t+1): Pb -> Pa ...
//   side stream        \-> k_f1(t+1) ----------------------/
//
// k_f1 (ONE CTA): measurement model, S = H P H^T + R, Cholesky, y = L^-1 nu and
//   A = L^-1 H_T (m x |T|, T = touched state indices).  The |T| x |T| block of P it needs
//   is read from the sweep's INPUT buffer Pa (complete since the sweep before) and brought up
//   to date with the panel of update t:  P[T,T] = Pa[T,T] + Q_t[T,T] - V_t[:,T]^T V_t[:,T]
//   ("lazy" operand) — the covariance is ping-ponged between two buffers by the sweep so
//   that Pa stays intact while Pb is being written.
// k_f2 (one CTA per 64 state columns, on the critical path): V[:, block] = A P[T, block]
//   (one small dense product), dx (+)= V^T y, the two tiled copies of the panel for the sweep,
//   covariance growth for new landmarks; the last CTA to finish applies the state update
//   (EKF:134-152 / ROT:143-177) on the final chunk of a frame.
//
// Per frame the critical path is k_f2 + k_sweep instead of k_front + k_sweep.
#pragma once
#include "front.cuh"

namespace asb {

struct F1Args {
    double* state;               // read; new-landmark entries are stored
    const double* P;             // covariance to read: the running sweep's INPUT buffer when lazy
    const double* PT_dense;      // row-sharded engine (filter_shard.cu): P[T,T] (nT x nT, lower triangle valid) summed
                                 // over the row owners by NCCL; P is not read at all when this is set
    int ld, n0, n1;              // state dims before / after this frame's augmentation
    Noise nz;
    const int* idx;              // landmark index of each observation of the chunk
    const double* poses;         // (kc,6)
    const unsigned char* isnew;
    int kc, m4;
    int first_chunk;
    int add_q;                   // predict not yet folded into P: use P + diag(Q)
    int use_delta;               // nu' = nu - H dx_in   (chunks after the first)
    const double* dx_in;
    int lazy;                    // P is one sweep behind: subtract Vp^T Vp (and add Q if lazy_q)
    const double* Vp;            // panel of that sweep by state column: column j at Vp + F1_LDT j, m4p values
    int m4p, lazy_q;
    const double* Vc_tiled;      // row-sharded engine: that panel in the tiled V_J layout (sweep.cuh) instead of Vp
    double* At;                  // out (nT, m4): A^T, A = L^-1 H_T (rows a >= m are zero)
    double* y;                   // out (m4): L^-1 nu (zero padded)
    int* T;                      // out (nT): touched state indices
    double* blocks_out;          // (kc, R, W) for asb_get_last_update
    double* nu_out;              // (R kc)
    int* status;
    int dbg;
    unsigned long long* prof;    // phase timestamps (globaltimer ns), 8 slots, may be null
};

constexpr int F1_LDT = 64;          // doubles per state column of the column-major panel copy (global)
constexpr int F1_LDK = 66;          // ... and of its gathered columns in shared memory (66 / 2 odd: conflict-free 128-bit loads)

template <int KIND>
struct F1Smem {
    static constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D;
    static __host__ __device__ size_t rows(int kc) {   // [S; nu^T; I] for the Cholesky, earlier the gathered panel (nT x F1_LDK)
        const size_t m = R * kc, nT = CAM + D * kc, a = (size_t)128 * (m | 1), b = (size_t)F1_LDK * nT;
        return ((a > b ? a : b) + 1) & ~(size_t)1;
    }
    static __host__ __device__ size_t bytes(int kc, int m4) {
        const size_t nT = CAM + D * kc;
        size_t dbl = ((nT * nT + 1) & ~(size_t)1) + rows(kc) + (size_t)kc * R * W + (size_t)kc * D + (size_t)m4 + 16;
        return dbl * sizeof(double) + (nT + 2 * kc + 8) * sizeof(int);
    }
};

// EXCL = false: <= 128 registers, the CTA must fit on an SM next to ONE resident sweep CTA (122 registers x
// 256 threads; per-frame calls, side stream).  EXCL = true: the CTA has an SM to itself (stream mode,
// row-sharded engine): no register cap, and EKF chunks with an even number of markers factor S two
// markers (6 rows) per block step.
template <int KIND, bool EXCL>
__global__ void __launch_bounds__(FRONT_THREADS, EXCL ? 1 : 2)
k_f1(F1Args p) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D, NB = R;
    constexpr int NT = FRONT_THREADS;
    extern __shared__ __align__(16) unsigned char fsm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kc = p.kc, m = R * kc, m4 = p.m4, nT = CAM + D * kc;
    const int lds = m | 1;
    const int ld = p.ld, n0 = p.n0;
    FrontArgs stampargs{};       // front_stamp only looks at prof
    stampargs.prof = p.prof;

    double* s_PT = reinterpret_cast<double*>(fsm);                       // nT x nT (dense)
    double* s_S = s_PT + (((size_t)nT * nT + 1) & ~(size_t)1);           // 128 x lds; before that the gathered panel
    double* s_blk = s_S + F1Smem<KIND>::rows(kc);                        // kc x R x W
    double* s_lm = s_blk + (size_t)kc * R * W;                           // kc x D
    double* s_nu = s_lm + (size_t)kc * D;                                // m4
    double* s_cam = s_nu + m4;                                           // 10 (+6 pad)
    int* s_T = reinterpret_cast<int*>(s_cam + 16);                       // nT
    int* s_idx = s_T + nT;                                               // kc
    int* s_new = s_idx + kc;                                             // kc
    double* s_Vg = s_S;                                                  // nT x F1_LDK: Vp[:, T] by column
    __shared__ __align__(8) uint64_t s_bar;

    if (tid == 0) {
        mbar_init(&s_bar, nT);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // stream mode (programmatic dependent launches): resident while the previous k_f2 drains; once
    // that has completed, release the sweep that is queued behind this kernel
    pdl_wait();
    pdl_trigger();
    front_stamp(stampargs, 0);
    // ---- touched index set T ----------------------------------------------------------------
    if (tid < CAM) { s_cam[tid] = p.state[tid]; s_T[tid] = tid; p.T[tid] = tid; }
    for (int e = tid; e < D * kc; e += NT) {
        const int o = e / D, c = e - o * D;
        const int ix = p.idx[o];
        s_T[CAM + e] = CAM + D * ix + c;
        p.T[CAM + e] = CAM + D * ix + c;
        if (c == 0) { s_idx[o] = ix; s_new[o] = p.isnew ? p.isnew[o] : 0; }
    }
    __syncthreads();

    // ---- all global reads, asynchronously: the lower triangle of P[T,T] (cp.async) and the
    //      previous panel's columns T (one bulk copy, SASS UBLKCP, per column) -------------------
    // Entries in rows / columns >= n0 (this frame's augmentation) are known constants.
    if (p.lazy && p.Vc_tiled) {
        for (int e = tid; e < nT * p.m4p; e += NT) {
            const int col = e / p.m4p, a = e - col * p.m4p;
            const int i = s_T[col];
            s_Vg[(size_t)col * F1_LDK + a] =
                i < n0 ? p.Vc_tiled[((size_t)(i / V_BLK) * p.m4p + a) * V_LDB + v_pos_col(i % V_BLK)] : 0.0;
        }
    } else if (p.lazy && tid < nT) {
        const int i = s_T[tid];
        const uint32_t bytes = i < n0 ? (uint32_t)(p.m4p * sizeof(double)) : 0u;
        mbar_expect_tx(&s_bar, bytes);
        if (bytes) bulk_g2s(s_Vg + (size_t)tid * F1_LDK, p.Vp + (size_t)i * F1_LDT, bytes, &s_bar);
        else for (int k = 0; k < p.m4p; ++k) s_Vg[(size_t)tid * F1_LDK + k] = 0.0;
    }
    for (int a = warp; a < nT; a += NT / 32) {
        const int i = s_T[a];
        const double* prow = p.P + (long long)i * ld;
        for (int b = lane; b <= a; b += 32) {
            const int j = s_T[b];
            // the covariance may hold only its lower triangle (SweepExtra::lower_only): read P[max][min]
            if (i < n0 && j < n0)
                cp_async8(s_PT + a * nT + b, p.PT_dense ? p.PT_dense + a * nT + b
                                                        : (j <= i ? prow + j : p.P + (long long)j * ld + i));
            else s_PT[a * nT + b] = (i == j) ? p.nz.init_lm : 0.0;
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    front_stamp(stampargs, 1);

    // ---- measurement model while the copies fly ----------------------------------------------
    for (int o = tid; o < kc; o += NT) {
        double lm[D];
        const int li = CAM + D * s_idx[o];
        if (s_new[o]) {
            new_landmark<KIND>(s_cam, p.poses + 6 * o, lm);
#pragma unroll
            for (int c = 0; c < D; ++c) p.state[li + c] = lm[c];
        } else {
#pragma unroll
            for (int c = 0; c < D; ++c) lm[c] = p.state[li + c];
        }
#pragma unroll
        for (int c = 0; c < D; ++c) s_lm[o * D + c] = lm[c];
    }
    for (int a = m + tid; a < m4; a += NT) s_nu[a] = 0.0;
    __syncthreads();
    for (int o = tid; o < kc; o += NT) {
        double lm[D];
        int src = o;
        if (!s_new[o])  // a repeated id whose landmark was created earlier in this same chunk
            for (int o2 = 0; o2 < kc; ++o2)
                if (s_new[o2] && s_idx[o2] == s_idx[o]) { src = o2; break; }
#pragma unroll
        for (int c = 0; c < D; ++c) lm[c] = s_lm[src * D + c];
        // the dense block goes straight to shared memory: as a local array its R x W values (140 for
        // EKF_Rotations) do not fit the register file and the model ran out of local memory
        double lnu[R];
        double* blk = s_blk + (size_t)o * R * W;
        observation_model<KIND>(s_cam, lm, p.poses + 6 * o, lnu, blk);
        if (p.nu_out) {
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
        for (int r = 0; r < R; ++r) s_nu[R * o + r] = lnu[r];
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    front_stamp(stampargs, 2);

    // ---- lazy operand: bring P[T,T] up to date with the sweep that is still running ---------
    // Lower triangle only, then mirrored.  Thread tile: rows {ta + G q}, columns {tb + G q'} (q, q' < 4,
    // G = ceil(nT / 4)): lanes read consecutive rows of the gathered panel.
    if (p.lazy) {
        if (!p.Vc_tiled) mbar_wait(&s_bar, 0);
        const int G = (nT + 3) >> 2;
        const int kp = (p.m4p + 1) >> 1;   // pairs of k (m4p is a multiple of 4)
        for (int task = tid; task < ftri(G); task += NT) {
            int ta = (int)((sqrtf(8.0f * task + 1.0f) - 1.0f) * 0.5f);
            while (ftri(ta + 1) <= task) ++ta;
            while (ftri(ta) > task) --ta;
            const int tb = task - ftri(ta);
            double acc[4][4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int r = 0; r < 4; ++r) acc[q][r] = 0.0;
            const double* pa[4];
            const double* pb[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                pa[q] = s_Vg + (size_t)min(ta + G * q, nT - 1) * F1_LDK;
                pb[q] = s_Vg + (size_t)min(tb + G * q, nT - 1) * F1_LDK;
            }
            for (int k2 = 0; k2 < kp; ++k2) {
                double2 va[4], vb[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    va[q] = *reinterpret_cast<const double2*>(pa[q] + 2 * k2);
                    vb[q] = *reinterpret_cast<const double2*>(pb[q] + 2 * k2);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int r = 0; r < 4; ++r) acc[q][r] += va[q].x * vb[r].x + va[q].y * vb[r].y;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int a = ta + G * q, b = tb + G * r;
                    if (a >= nT || b >= nT || (ta == tb && r > q)) continue;
                    const int i = s_T[a], j = s_T[b];
                    const int hi = max(a, b), lo = min(a, b);
                    double v = s_PT[hi * nT + lo];
                    if (i < n0 && j < n0) {
                        v -= acc[q][r];
                        if (p.lazy_q && i == j) v += process_noise(p.nz, i);
                    }
                    s_PT[hi * nT + lo] = v;
                    s_PT[lo * nT + hi] = v;
                }
        }
    } else {
        for (int e = tid; e < nT * nT; e += NT) {   // mirror the gathered lower triangle
            const int a = e / nT, b = e - a * nT;
            if (b > a) s_PT[e] = s_PT[b * nT + a];
        }
    }
    __syncthreads();
    front_stamp(stampargs, 6);

    // ---- predict folded in: diag(Q) -----------------------------------------------------------
    if (p.add_q) {
        for (int e = tid; e < nT * nT; e += NT) {   // same state index (also a repeated id: twice in T)
            const int a = e / nT, b = e - a * nT;
            if (s_T[a] == s_T[b]) s_PT[e] += process_noise(p.nz, s_T[a]);
        }
    }
    // [S; nu^T; I]: nu row and identity rows now (the panel copy in s_S is dead)
    for (int e = tid; e < 64 * lds; e += NT) {
        const int jj = e / lds, c = e - jj * lds;
        s_S[(GROW + jj) * lds + c] = (jj == c && c < m) ? 1.0 : 0.0;
    }
    for (int a = tid; a < m; a += NT) s_S[m * lds + a] = s_nu[a];
    __syncthreads();
    front_stamp(stampargs, 7);

    // ---- S = H PT H^T + R, one task per (observation pair, row): no m x nT intermediate --------
    {
        const int ntask = ftri(kc) * R;
        for (int task = tid; task < ntask; task += NT) {
            const int pr = task / R, r = task - pr * R;
            int oa = (int)((sqrtf(8.0f * pr + 1.0f) - 1.0f) * 0.5f);
            while (ftri(oa + 1) <= pr) ++oa;
            while (ftri(oa) > pr) --oa;
            const int ob = pr - ftri(oa);
            const double* ha = s_blk + (size_t)(oa * R + r) * W;
            const double* hb = s_blk + (size_t)ob * R * W;
            double tmp[W];
#pragma unroll
            for (int c = 0; c < W; ++c) tmp[c] = 0.0;
#pragma unroll
            for (int c1 = 0; c1 < W; ++c1) {
                const double h = ha[c1];
                const double* prow = s_PT + (size_t)(c1 < CAM ? c1 : CAM + D * oa + (c1 - CAM)) * nT;
#pragma unroll
                for (int c = 0; c < W; ++c) tmp[c] += h * prow[c < CAM ? c : CAM + D * ob + (c - CAM)];
            }
            const int a = R * oa + r;
#pragma unroll
            for (int r2 = 0; r2 < R; ++r2) {
                const int b = R * ob + r2;
                double sv = (a == b) ? p.nz.r : 0.0;
#pragma unroll
                for (int c = 0; c < W; ++c) sv += tmp[c] * hb[r2 * W + c];
                if (b <= a) s_S[a * lds + b] = sv;
            }
        }
    }
    __syncthreads();
    front_stamp(stampargs, 3);

    // ---- Cholesky of [S; nu^T; I]: row m -> y^T, row GROW+b -> column b of L^-1 ------------------
    // EKF with an even number of markers: two markers (6 rows) per block step halves the number of serial
    // steps (each costs two CTA barriers and a shared-memory round trip of the block column)
    bool bad;
    if constexpr (KIND == 0 && EXCL) {
        bad = ((kc & 1) == 0 && !(p.dbg & 16)) ? front_cholesky<2 * NB>(s_S, lds, m, p.dbg)
                                               : front_cholesky<NB>(s_S, lds, m, p.dbg);
    } else {
        bad = front_cholesky<NB>(s_S, lds, m, p.dbg);
    }
    if (bad && tid == 0) atomicExch(p.status, ASB_ERR_NOT_SPD);
    front_stamp(stampargs, 4);

    // ---- outputs: y, A^T = (L^-1 H_T)^T ----------------------------------------------------------
    // Linv[a][b] = s_S[(GROW + b) * lds + a] (zero for b > a); H_T row b = block row b of s_blk
    for (int a = tid; a < m4; a += NT) p.y[a] = a < m ? s_S[m * lds + a] : 0.0;
    for (int e = tid; e < nT * m4; e += NT) {
        const int tb = e / m4, a = e - tb * m4;
        double s = 0.0;
        if (a < m) {
            if (tb < CAM) {   // camera columns: every row of H contributes; four partial sums shorten the chain
                double s1 = 0.0, s2 = 0.0, s3 = 0.0;
                int b = 0;
                for (; b + 3 <= a; b += 4) {
                    s += s_S[(GROW + b) * lds + a] * s_blk[(size_t)b * W + tb];
                    s1 += s_S[(GROW + b + 1) * lds + a] * s_blk[(size_t)(b + 1) * W + tb];
                    s2 += s_S[(GROW + b + 2) * lds + a] * s_blk[(size_t)(b + 2) * W + tb];
                    s3 += s_S[(GROW + b + 3) * lds + a] * s_blk[(size_t)(b + 3) * W + tb];
                }
                for (; b <= a; ++b) s += s_S[(GROW + b) * lds + a] * s_blk[(size_t)b * W + tb];
                s = (s + s1) + (s2 + s3);
            } else {
                const int o = (tb - CAM) / D, c = (tb - CAM) - o * D;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int b = R * o + r;
                    if (b <= a) s += s_S[(GROW + b) * lds + a] * s_blk[(size_t)b * W + CAM + c];
                }
            }
        }
        p.At[e] = s;
    }
    front_stamp(stampargs, 5);
}

// ------------------------------------------------------------------------------------------
struct F2Args {
    double* state;               // updated by the last CTA when apply_kind >= 0
    double* P;                   // current covariance: rows T are read; growth stores on the first chunk
    int ld, n0, n1;
    Noise nz;
    int kc, m4;
    int first_chunk, add_q, use_delta;
    const double* At;            // (nT, m4) from k_f1
    const double* y;             // (m4)
    const int* T;                // (nT)
    const double* dx_in;
    double* dx_out;
    double* Vr;                  // tiled panel in V_I column order (sweep.cuh)
    double* Vc;                  // tiled panel in V_J column order
    double* Vt;                  // panel by state column (F1_LDT doubles each) for the next k_f1's lazy operand
    int lower_only;              // P holds only its lower triangle: row T[a] right of the diagonal is read from column T[a]
    int apply_kind;              // -1: none (more chunks follow), 0: EKF, 1: EKF_Rotations
    int as_written, nlm;
    unsigned int* counter;       // CTAs done (self-resetting)
    unsigned long long* prof;    // 4 slots, may be null
};

constexpr int F2_THREADS = 256;
constexpr int F2_COLS = 32;           // state columns per CTA: half a 64-column block of the sweep's panels
constexpr int F2_LDB = F2_COLS + 4;   // row stride of the staged P rows  (= 4 mod 16: conflict-free DMMA B fragments)
constexpr int F2_LDA = 64 + 4;        // row stride of A^T in shared memory (= 4 mod 16: conflict-free DMMA A fragments)
constexpr int F2_LDV = F2_COLS + 2;   // row stride of the V slice in shared memory

template <int KIND>
struct F2Smem {
    static constexpr int D = Dims<KIND>::D;
    static __host__ __device__ size_t bytes(int kc, int m4) {
        const size_t nT = CAM + D * kc, kp = (nT + 3) & ~(size_t)3;
        size_t dbl = kp * F2_LDB + kp * F2_LDA + (size_t)64 * F2_LDV + 64 + 8 * F2_COLS;
        return dbl * sizeof(double) + (nT + 4) * sizeof(int);
    }
};

template <int KIND>
__global__ void __launch_bounds__(F2_THREADS, 2)
k_f2(F2Args p) {
    constexpr int D = Dims<KIND>::D;
    constexpr int NT = F2_THREADS;
    extern __shared__ __align__(16) unsigned char fsm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kc = p.kc, m4 = p.m4, nT = CAM + D * kc, kp = (nT + 3) & ~3;
    const int ld = p.ld, n0 = p.n0, n1 = p.n1;
    const int j0 = blockIdx.x * F2_COLS;
    const int vblock = blockIdx.x >> 1, half = blockIdx.x & 1;   // 64-column panel block and which half of it
    FrontArgs stampargs{};
    stampargs.prof = p.prof;

    double* s_stage = reinterpret_cast<double*>(fsm);        // kp x F2_LDB: P_eff[T, block] (rows >= nT zero)
    double* s_At = s_stage + (size_t)kp * F2_LDB;            // kp x F2_LDA: A^T (columns >= m4 and rows >= nT zero)
    double* s_V = s_At + (size_t)kp * F2_LDA;                // 64 x F2_LDV: V[:, block]
    double* s_y = s_V + (size_t)64 * F2_LDV;                 // 64
    double* s_red = s_y + 64;                                // 8 x F2_COLS partial sums of V^T y
    int* s_T = reinterpret_cast<int*>(s_red + 8 * F2_COLS);  // nT
    __shared__ int s_last;

    pdl_wait();      // stream mode: made resident while the sweep before it retired
    pdl_trigger();   // the next k_f1 may take one of the SMs this grid leaves empty; it waits for this grid's end
    front_stamp(stampargs, 0);
    // EKF: the state entries this CTA will update at the very end (their load latency is off the tail)
    double state_pre = 0.0;
    if (KIND == 0 && p.apply_kind == 0 && tid < F2_COLS && j0 + tid < n1) state_pre = p.state[j0 + tid];
    for (int a = tid; a < nT; a += NT) s_T[a] = p.T[a];
    for (int a = tid; a < 64; a += NT) s_y[a] = a < m4 ? p.y[a] : 0.0;
    {   // A^T rows (m4 is a multiple of 4: 16-byte pieces never straddle a row)
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
    // stage[a][0..32) = P_eff[T[a]][j0..j0+32); entries >= n0 are this frame's constants.
    // Half a warp per row: 16 lanes x 16 bytes.
    for (int a = 2 * warp + (lane >> 4); a < nT; a += 2 * (NT / 32)) {
        const int i = s_T[a];
        const double* prow = p.P + (long long)i * ld;
        const int c = 2 * (lane & 15);
        const int j = j0 + c;
        double* dst = s_stage + (size_t)a * F2_LDB + c;
        if (i < n0 && j + 1 < n0 && (!p.lower_only || j + 1 <= i)) {
            cp_async16(dst, prow + j);
        } else {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int jj = j + q;
                if (i < n0 && jj < n0)
                    cp_async8(dst + q, (!p.lower_only || jj <= i) ? prow + jj : p.P + (long long)jj * ld + i);
                else dst[q] = (i == jj && i < n1) ? p.nz.init_lm : 0.0;
            }
        }
    }
    cp_async_wait_all();
    __syncthreads();
    if (p.add_q) {   // predict folded in: diag(Q) on the staged rows
        for (int a = tid; a < nT; a += NT) {
            const int i = s_T[a];
            if (i >= j0 && i < j0 + F2_COLS) s_stage[(size_t)a * F2_LDB + (i - j0)] += process_noise(p.nz, i);
        }
        __syncthreads();
    }
    front_stamp(stampargs, 1);

    // V[:, block] = A stage on the FP64 tensor pipe (DMMA.8x8x4): warp w owns rows 8 w .. 8 w + 7 and all
    // four 8-column tiles; lane (g, t) holds V[8 w + g][8 ct + 2 t .. + 1].
    {
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
    front_stamp(stampargs, 2);

    // panel by state column for the next k_f1's lazy operand: 4 consecutive rows per 256-bit store
    // (null: that k_f1 gathers its columns from the tiled V_J panel instead)
    if (p.Vt) {
        const int q4 = m4 >> 2;
        for (int e = tid; e < F2_COLS * q4; e += NT) {
            const int c = e / q4, q = e - c * q4;
            const double* v = s_V + (size_t)(4 * q) * F2_LDV + c;
            st256(p.Vt + (size_t)(j0 + c) * F1_LDT + 4 * q, v[0], v[F2_LDV], v[2 * F2_LDV], v[3 * F2_LDV]);
        }
    }
    // dx (+)= V^T y; EKF (additive landmark states): each CTA applies its own columns, CTA 0 the camera
    __shared__ double s_dcam[CAM];
    {
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
        if (j < n1) {
            const double d = (p.use_delta ? p.dx_in[j] : 0.0) + dsum;
            p.dx_out[j] = d;
            if (KIND == 0 && p.apply_kind == 0) {
                if (j >= CAM) p.state[j] = state_pre + d;
                else s_dcam[j] = d;
            }
        }
    }
    if (KIND == 0 && p.apply_kind == 0 && blockIdx.x == 0) {
        __syncthreads();
        if (tid == 0) {
            double cam[CAM], d[CAM];
#pragma unroll
            for (int c = 0; c < CAM; ++c) { cam[c] = p.state[c]; d[c] = s_dcam[c]; }
            camera_reset(cam, d, p.as_written);
#pragma unroll
            for (int c = 0; c < CAM; ++c) p.state[c] = cam[c];
        }
    }
    {
        // both column orders of this half block (sweep.cuh): position `pos` of the 64-column block holds
        // column cr (V_I order) / cc (V_J order); both maps stay inside a 32-column half
        double* vr = p.Vr + (size_t)vblock * m4 * V_LDB;
        double* vc = p.Vc + (size_t)vblock * m4 * V_LDB;
        for (int e = tid; e < m4 * F2_COLS; e += NT) {
            const int a = e / F2_COLS, pos = F2_COLS * half + (e - a * F2_COLS);
            const int cr = (pos & ~31) + 4 * (pos & 7) + ((pos & 31) >> 3);   // inverse maps: which column lands at `pos`
            const int cc = (pos & ~15) + 4 * ((pos & 7) >> 1) + 2 * ((pos & 15) >> 3) + (pos & 1);
            vr[a * V_LDB + pos] = s_V[(size_t)a * F2_LDV + (cr - F2_COLS * half)];
            vc[a * V_LDB + pos] = s_V[(size_t)a * F2_LDV + (cc - F2_COLS * half)];
        }
    }

    // covariance growth for this block's columns (first chunk only; EKF:279-290)
    if (p.first_chunk && n1 > n0) {
        const int ja = j0, jb = min(j0 + F2_COLS, n1);
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

    // EKF_Rotations (landmark quaternions straddle column blocks): the last CTA to get here applies
    // the state update with the complete dx
    if (KIND == 1 && p.apply_kind >= 0) {
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = (atomicAdd(p.counter, 1u) == gridDim.x - 1) ? 1 : 0;
        __syncthreads();
        if (s_last) {
            __threadfence();
            apply_state(p.state, p.dx_out, p.apply_kind, p.as_written, p.nlm, tid, NT);
            if (tid == 0) *p.counter = 0u;
        }
    }
    front_stamp(stampargs, 3);
}

}  // namespace asb
