// Wide batched engine: many independent filters whose covariance does NOT fit one CTA's shared
// memory (EKF_Rotations at 32 / 64 markers: N = 330 / 650, P 0.46 / 1.7 MB per sequence; BASELINE
// configs C2-ROT / C5-ROT).  Frame-synchronous: for frame t every sequence advances together through
// four kernels per update chunk (<= WCK = 16 markers, m <= 112 rows), the covariance staying in HBM:
//
//   k_wide_model   one CTA per sequence: pending state update + quaternion reset (ROT:143-177),
//                  id association / augmentation in frame order (ROT:66-90, 275-335), measurement
//                  model (ROT:337-427 closed form), S = H P[T,T] H^T + R (ROT:128-133) straight from
//                  the touched 20 x 20 blocks of P
//   k_wide_chol    one CTA per sequence: blocked Cholesky of S on the FP64 tensor pipe,
//                  y = L^-1 nu', the 8 x 8 inverse diagonal blocks of L
//   k_wide_panel   one CTA per (sequence, 64 state columns): G = H P[T, cols] (H never materialised),
//                  V = L^-1 G by blocked forward substitution with DMMA, delta += V^T y
//   k_wide_sweep   one CTA per (sequence, 64-row strip of the lower triangle): P <- P + Q - V^T V
//                  (= (I - K H) P, ROT:180-181), V_I resident, V_J slabs by TMA bulk copy on an mbarrier
//                  ring fed by a producer warp, DMMA
//
// K is never formed (K nu = V^T y, K H P = V^T V).  Chunks of one frame condition sequentially with
// the innovation corrected by H delta (DESIGN.md "chunked update"); predict (P += diag Q, ROT:103-113)
// is folded into the first chunk.  Same model code (model.cuh) as the other engines.
//
// Covariance layout in HBM ("tiled"): lower triangle as 8 x 8 tiles, tile (I8, J8 <= I8) at
// tri(I8) + J8, 64 doubles row-major, diagonal tiles full.  A warp's DMMA accumulator fragment of
// one tile is then one contiguous 512-byte access (lane l owns doubles 2l, 2l+1).
#pragma once
#include <cstdint>

#include "batch_kernels.cuh"
#include "sweep.cuh"

namespace asb {

constexpr int WCK = 16;         // markers per update chunk
constexpr int W_VLD = 68;       // doubles per V panel row of a 64-column block (= 4 mod 16: conflict-free fragments)
// measured at the C2-ROT / C5-ROT shapes (slab rows, stages): (16,4) 654.6 / 399.3 ms, (8,8) 685, (24,3) 645.9,
// (32,2) 641.1 / 390.7, (40,2) 637.9 / 388.4: fewer, larger slabs (less mbarrier traffic per k-step) win
#ifndef W_SLAB_ROWS
#define W_SLAB_ROWS 40
#define W_STAGES 2
#endif
constexpr int W_SLAB = W_SLAB_ROWS;   // V rows per pipeline stage of the sweep
constexpr int W_NST = W_STAGES;       // stages

template <int KIND> struct WideDims {
    static constexpr int R = Dims<KIND>::R, D = Dims<KIND>::D, W = CAM + D;
    static constexpr int MC = R * WCK;                 // rows per chunk
    static constexpr int M8 = (MC + 7) & ~7;           // padded to whole 8-row blocks
    static constexpr int NB = M8 / 8;
    static constexpr int TW = CAM + D * WCK;           // touched state columns per chunk
    static constexpr int LROW = (NB - 1) * 64 > 12 * 68 ? (NB - 1) * 64 : 12 * 68;   // k_wide_panel's L row buffer
    static constexpr int HLS = 16 + 8 * ((D + 7) / 8) + 4;   // padded block row: [cam 16 | lm 8 ceil(D/8)] + 4 (= 4 mod 16)
};

struct WideParams {
    double* states;
    double* covs;              // tiled layout, cov_stride doubles per sequence
    int* nlm;
    int* lm_ids;
    int* status;
    const int* ids;
    const double* poses;
    const long long* offsets;
    long long state_stride, cov_stride;
    int n_seq, n_frames, max_lm, max_obs, as_written;
    int t, round;
    int obs_seqs;              // distinct observation streams: sequence s reads stream s % obs_seqs
    int seq0;                  // this launch covers sequences seq0 + blockIdx (sequence groups run on separate streams)
    Noise nz;
    // per-sequence scratch
    int* meta;                 // 8 ints: k, nchunks, chunk size, n, pending, m8 (0 = idle this round), kc, c0
    int* idx;                  // max_obs landmark indices of the current frame
    double* delta;             // state_stride
    double* H;                 // WCK * R * W   measurement blocks of the current chunk
    double* Hp;                // WCK * 8 * HLS the same blocks zero-padded to fragment order (k_wide_model's S assembly)
    double* L;                 // tri(NB) * 64  Cholesky factor, 8 x 8 tiles
    double* Dinv;              // NB * 64       inverse diagonal blocks
    double* y;                 // M8
    double* V;                 // nb64 * M8 * W_VLD panel, 64-column blocks
    long long v_stride;        // doubles per sequence of V
    int nb64;                  // 64-column blocks at capacity
};

__device__ __forceinline__ void w_cp_async16(void* smem, const void* gmem) {   // 16-byte LDGSTS, L2 only
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ long long wt_off(int i, int j) {   // (i >> 3) >= (j >> 3)
    return ((long long)(tri(i >> 3) + (j >> 3)) << 6) + ((i & 7) << 3) + (j & 7);
}
__device__ __forceinline__ double wt_sym(const double* __restrict__ P, int i, int j) {
    return (i >> 3) >= (j >> 3) ? P[wt_off(i, j)] : P[wt_off(j, i)];
}

// ---- k_wide_model: association, augmentation, measurement model, S -> HBM (high occupancy: the serial
//      model evaluation and the ragged S assembly of one sequence hide behind five other CTAs) --------
template <int KIND>
__global__ void __launch_bounds__(128, 6) k_wide_model(WideParams p) {
    using WD = WideDims<KIND>;
    constexpr int D = WD::D, R = WD::R, W = WD::W, M8 = WD::M8, NB = WD::NB, NT = 128;
    extern __shared__ __align__(16) unsigned char smraw[];
    constexpr int HLS = WD::HLS;
    double* s_Pc = reinterpret_cast<double*>(smraw);          // CAM x TW: P[cam, touched columns]
    double* s_pose = s_Pc + CAM * WD::TW;                     // max(max_obs, WCK) x 6
    const int cap_obs = max(p.max_obs, WCK);
    int* s_ids = reinterpret_cast<int*>(s_pose + (size_t)cap_obs * 6);
    int* s_idx = s_ids + cap_obs;
    int* s_new = s_idx + cap_obs;
    int* s_tab = s_new + cap_obs;                             // max_lm
    int* s_misc = s_tab + p.max_lm;                           // [0] nlm after, [1] status, [2] bad
    const int tid = threadIdx.x;
    const int seq = p.seq0 + blockIdx.x;
    int* meta = p.meta + (long long)seq * 8;
    double* state = p.states + (long long)seq * p.state_stride;
    double* P = p.covs + (long long)seq * p.cov_stride;
    double* delta = p.delta + (long long)seq * p.state_stride;
    int* g_idx = p.idx + (long long)seq * p.max_obs;
    const long long orec = (long long)(seq % p.obs_seqs) * p.n_frames + p.t;
    const long long ob = p.offsets[orec];
    int k, cs, n;

    if (p.round == 0) {
        k = (int)(p.offsets[orec + 1] - ob);
        const int st = p.status[seq];
        if (k == 0 || k > p.max_obs || st == ASB_ERR_CAPACITY) {   // no detection: no predict either (BASE:197-204)
            if (tid == 0) {
                meta[1] = 0; meta[5] = 0;
                if (k > p.max_obs) p.status[seq] = ASB_ERR_CAPACITY;
            }
            return;
        }
        int nlm = p.nlm[seq];
        // pending state update of the previous frame (its delta is complete: all panels ran)
        if (meta[4]) {
            if (tid == 0) {
                double cam[CAM], d[CAM];
#pragma unroll
                for (int c = 0; c < CAM; ++c) { cam[c] = state[c]; d[c] = delta[c]; }
                camera_reset(cam, d, p.as_written);
#pragma unroll
                for (int c = 0; c < CAM; ++c) state[c] = cam[c];
            }
            for (int l = tid; l < nlm; l += NT) {
                double lm[D], d[D];
#pragma unroll
                for (int c = 0; c < D; ++c) { lm[c] = state[CAM + D * l + c]; d[c] = delta[CAM + D * l + c]; }
                landmark_update<KIND>(lm, d);
#pragma unroll
                for (int c = 0; c < D; ++c) state[CAM + D * l + c] = lm[c];
            }
        }
        for (int o = tid; o < k; o += NT) s_ids[o] = p.ids[ob + o];
        for (int e = tid; e < 6 * k; e += NT) s_pose[e] = p.poses[6 * ob + e];
        for (int i = tid; i < nlm; i += NT) s_tab[i] = p.lm_ids[(long long)seq * p.max_lm + i];
        if (tid == 0) s_misc[1] = 0;
        __syncthreads();
        for (int o = tid; o < k; o += NT) {
            const int id = s_ids[o];
            int found = -1;
            for (int i = 0; i < nlm; ++i)
                if (s_tab[i] == id) { found = i; break; }
            s_idx[o] = found;
            s_new[o] = 0;
        }
        __syncthreads();
        if (tid == 0) {   // unseen ids, in frame order; duplicates bind to the first (EKF:75-78)
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
        __syncthreads();
        const int nlm1 = s_misc[0];
        if (s_misc[1] == ASB_ERR_CAPACITY) {
            if (tid == 0) { p.status[seq] = ASB_ERR_CAPACITY; meta[1] = 0; meta[5] = 0; meta[4] = 0; }
            return;
        }
        n = CAM + D * nlm1;
        if (nlm1 > nlm) {   // augmentation: new landmark states, block-diagonal covariance growth (rows/cols are zero)
            for (int o = tid; o < k; o += NT) {
                if (!s_new[o]) continue;
                double cam[CAM], lm[D];
#pragma unroll
                for (int c = 0; c < CAM; ++c) cam[c] = state[c];
                new_landmark<KIND>(cam, s_pose + 6 * o, lm);
#pragma unroll
                for (int c = 0; c < D; ++c) state[CAM + D * s_idx[o] + c] = lm[c];
            }
            for (int i = CAM + D * nlm + tid; i < n; i += NT) P[wt_off(i, i)] = p.nz.init_lm;
            for (int i = nlm + tid; i < nlm1; i += NT) p.lm_ids[(long long)seq * p.max_lm + i] = s_tab[i];
        }
        for (int i = tid; i < n; i += NT) delta[i] = 0.0;
        for (int o = tid; o < k; o += NT) g_idx[o] = s_idx[o];
        const int nchunks = (k + WCK - 1) / WCK;
        cs = (k + nchunks - 1) / nchunks;
        if (tid == 0) {
            p.nlm[seq] = nlm1;
            meta[0] = k; meta[1] = nchunks; meta[2] = cs; meta[3] = n; meta[4] = 1;
        }
        __syncthreads();   // state / P diagonal / delta writes of this CTA are visible to its own reads below
    } else {
        if (p.round >= meta[1]) {
            if (tid == 0) meta[5] = 0;
            return;
        }
        k = meta[0]; cs = meta[2]; n = meta[3];
        const int c0r = p.round * cs, kcr = min(cs, k - c0r);
        for (int o = tid; o < kcr; o += NT) s_idx[c0r + o] = g_idx[c0r + o];
        for (int e = tid; e < 6 * kcr; e += NT) s_pose[6 * c0r + e] = p.poses[6 * (ob + c0r) + e];
        __syncthreads();
    }
    const int c0 = p.round * cs, kc = min(cs, k - c0), m = R * kc, m8 = (m + 7) & ~7;
    const bool add_q = p.round == 0;
    double* gH = p.H + (long long)seq * (WCK * R * W);
    double* gS = p.L + (long long)seq * (tri(NB) * 64);       // S tiles (k_wide_chol factors them in place)
    double* gnu = p.y + (long long)seq * M8;
    // ---- measurement model at the frame's pre-update state; nu' = nu - H delta -------------------
    for (int o = tid; o < kc; o += NT) {
        double cam[CAM], lm[D], lnu[R];
        double* blk = gH + o * R * W;
#pragma unroll
        for (int c = 0; c < CAM; ++c) cam[c] = state[c];
        const int li = CAM + D * s_idx[c0 + o];
#pragma unroll
        for (int c = 0; c < D; ++c) lm[c] = state[li + c];
        observation_model<KIND>(cam, lm, s_pose + 6 * (c0 + o), lnu, blk);
        if (p.round > 0) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                double s = lnu[r];
#pragma unroll
                for (int c = 0; c < CAM; ++c) s -= blk[r * W + c] * delta[c];
#pragma unroll
                for (int c = 0; c < D; ++c) s -= blk[r * W + CAM + c] * delta[li + c];
                lnu[r] = s;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) gnu[R * o + r] = lnu[r];
    }
    for (int a = m + tid; a < m8; a += NT) gnu[a] = 0.0;
    // P[cam, touched columns] (+ Q on the diagonal in the first round) -> shared
    const int tw = CAM + D * kc;
    for (int e = tid; e < CAM * tw; e += NT) {
        const int c = e / tw, sl = e - c * tw;
        const int col = sl < CAM ? sl : CAM + D * s_idx[c0 + (sl - CAM) / D] + (sl - CAM) % D;
        double v = wt_sym(P, c, col);
        if (add_q && c == col) v += process_noise(p.nz, c);
        s_Pc[c * WD::TW + sl] = v;
    }
    __syncthreads();   // the blocks in gH (written by this CTA's model threads) are complete
    // measurement blocks in fragment-friendly padded form: s_H[o][g < 8][HLS], columns 0..9 camera, 16..16+D-1
    // landmark, everything else (rows >= R included) zero.  Kept in HBM scratch (read back through L1), not in
    // shared memory: 37 KB less per CTA is two more CTAs per SM for this latency-bound kernel
    double* s_H = p.Hp + (long long)seq * (WCK * 8 * HLS);
    for (int e = tid; e < kc * 8 * HLS; e += NT) {
        const int o = e / (8 * HLS), rem = e - o * (8 * HLS), gg = rem / HLS, c = rem - gg * HLS;
        double v = 0.0;
        if (gg < R) {
            if (c < CAM) v = gH[o * R * W + gg * W + c];
            else if (c >= 16 && c < 16 + D) v = gH[o * R * W + gg * W + CAM + (c - 16)];
        }
        s_H[e] = v;
    }
    __syncthreads();
    // ---- S = H P[T,T] H^T + R (lower triangle) on the FP64 tensor pipe, one warp per observation pair
    //      (o1 >= o2):  M (8 x [16 | 8 LN]) = H1 ([12 | 4 LK] padded) x P[T1, T2];  S_blk = M x H2^T.
    //      P[T1, T2]: camera rows / columns from the shared copy, the D x D landmark block from HBM. ----
    {
        const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t4 = lane & 3;
        constexpr int LK = (D + 3) / 4, LN = (D + 7) / 8, NTL = 2 + LN;
        // the D x D landmark block P[l1.., l2..] of a pair comes from HBM in B-fragment order, one pair ahead
        auto pair_of = [&](int pr, int& o1, int& o2) {
            o1 = (int)((sqrtf(8.0f * pr + 1.0f) - 1.0f) * 0.5f);
            while (tri(o1 + 1) <= pr) ++o1;
            while (tri(o1) > pr) --o1;
            o2 = pr - tri(o1);
        };
        auto load_lm = [&](int pr, double (&blk)[LK][LN]) {
            int o1, o2;
            pair_of(pr, o1, o2);
            const int l1 = CAM + D * s_idx[c0 + o1], l2 = CAM + D * s_idx[c0 + o2];
#pragma unroll
            for (int kk = 0; kk < LK; ++kk)
#pragma unroll
                for (int nt = 0; nt < LN; ++nt) {
                    const int k = 4 * kk + t4, nn = 8 * nt + g;
                    double v = 0.0;
                    if (k < D && nn < D) {
                        v = wt_sym(P, l1 + k, l2 + nn);
                        if (add_q && l1 + k == l2 + nn) v += p.nz.q_lm;
                    }
                    blk[kk][nt] = v;
                }
        };
        double nxt[LK][LN];
        if (warp < tri(kc)) load_lm(warp, nxt);
        for (int pr = warp; pr < tri(kc); pr += NT / 32) {
            double cur[LK][LN];
#pragma unroll
            for (int kk = 0; kk < LK; ++kk)
#pragma unroll
                for (int nt = 0; nt < LN; ++nt) cur[kk][nt] = nxt[kk][nt];
            if (pr + NT / 32 < tri(kc)) load_lm(pr + NT / 32, nxt);
            int o1, o2;
            pair_of(pr, o1, o2);
            const int sl1 = CAM + D * o1, sl2 = CAM + D * o2;
            const double* h1 = s_H + o1 * 8 * HLS + g * HLS;
            const double* h2 = s_H + o2 * 8 * HLS + g * HLS;
            double mc[NTL][2];
#pragma unroll
            for (int nt = 0; nt < NTL; ++nt) mc[nt][0] = mc[nt][1] = 0.0;
            // camera rows of P[T1, T2] (k = 0..11, rows 10, 11 zero)
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) {
                const int k = 4 * kk + t4;
                const double a = h1[k];
                const double* pc = s_Pc + (k < CAM ? k : 0) * WD::TW;
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int nn = 8 * nt + g;
                    dmma884(mc[nt][0], mc[nt][1], a, (k < CAM && nn < CAM) ? pc[nn] : 0.0);
                }
#pragma unroll
                for (int nt = 0; nt < LN; ++nt) {
                    const int nn = 8 * nt + g;
                    dmma884(mc[2 + nt][0], mc[2 + nt][1], a, (k < CAM && nn < D) ? pc[sl2 + nn] : 0.0);
                }
            }
            // landmark rows (k' = 0..4 LK - 1): P[l1 + k', cam] by symmetry from the shared copy, P[l1 + k', l2 + n'] from HBM
#pragma unroll
            for (int kk = 0; kk < LK; ++kk) {
                const int k = 4 * kk + t4;
                const double a = h1[16 + k];
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int nn = 8 * nt + g;
                    dmma884(mc[nt][0], mc[nt][1], a, (k < D && nn < CAM) ? s_Pc[nn * WD::TW + sl1 + k] : 0.0);
                }
#pragma unroll
                for (int nt = 0; nt < LN; ++nt) dmma884(mc[2 + nt][0], mc[2 + nt][1], a, cur[kk][nt]);
            }
            // S_blk = M H2^T: M from accumulator to A-fragment layout with shuffles; k over the columns of M
            double s0 = 0.0, s1 = 0.0;
#pragma unroll
            for (int kq = 0; kq < 3 + LK; ++kq) {
                // column of M: camera part columns 4 kq + t4 (tiles 0, 1), landmark part 16 + 4 (kq - 3) + t4
                const int col = kq < 3 ? 4 * kq + t4 : 16 + 4 * (kq - 3) + t4;
                const int nt = kq < 3 ? (kq >> 1) : 2 + ((kq - 3) >> 1);
                const int cc = col & 7;
                const int src = (lane & ~3) | (cc >> 1);
                const double v0 = __shfl_sync(0xffffffffu, mc[nt][0], src), v1 = __shfl_sync(0xffffffffu, mc[nt][1], src);
                dmma884(s0, s1, (cc & 1) ? v1 : v0, h2[col]);
            }
            const int ra = R * o1 + g;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int cb = 2 * t4 + e, rb = R * o2 + cb;
                if (g < R && cb < R && rb <= ra) gS[wt_off(ra, rb)] = (e ? s1 : s0) + (ra == rb ? p.nz.r : 0.0);
            }
        }
    }
    if (tid == 0) { meta[5] = m8; meta[6] = kc; meta[7] = c0; }
}

// ---- k_wide_chol: S = L L^T in place (8 x 8 blocks, trailing updates on the FP64 tensor pipe), y = L^-1 nu',
//      the inverse diagonal blocks.  One CTA of 4 warps per sequence, S in shared memory. -----------------
template <int KIND>
__global__ void __launch_bounds__(128, 4) k_wide_chol(WideParams p) {
    using WD = WideDims<KIND>;
    constexpr int M8 = WD::M8, NB = WD::NB, NT = 128, NW = NT / 32, R = WD::R;
    extern __shared__ __align__(16) unsigned char smraw[];
    double* s_S = reinterpret_cast<double*>(smraw);           // tri(NB) tiles x 64
    double* s_nu = s_S + tri(NB) * 64;                        // M8
    double* s_dv = s_nu + M8;                                 // 64
    __shared__ int s_bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seq = p.seq0 + blockIdx.x;
    const int* meta = p.meta + (long long)seq * 8;
    const int m8 = meta[5];
    if (m8 == 0) return;
    const int m = R * meta[6], nb = m8 >> 3;
    double* gL = p.L + (long long)seq * (tri(NB) * 64);
    double* gD = p.Dinv + (long long)seq * (NB * 64);
    double* gy = p.y + (long long)seq * M8;
    for (int t = warp; t < tri(nb); t += NW) {
        // tile t = (ib, jb): padding rows (>= m) are identity, the strictly upper part of diagonal tiles is
        // never read.  One coalesced 512-byte read per tile.
        int ib = (int)((sqrtf(8.0f * t + 1.0f) - 1.0f) * 0.5f);
        while (tri(ib + 1) <= t) ++ib;
        while (tri(ib) > t) --ib;
        const int jb = t - tri(ib);
        const double2 v = *reinterpret_cast<const double2*>(gL + t * 64 + 2 * lane);
        const int a = 8 * ib + (lane >> 2), b = 8 * jb + 2 * (lane & 3);
        double2 o;
        o.x = a < m ? (b <= a ? v.x : 0.0) : (a == b ? 1.0 : 0.0);
        o.y = a < m ? (b + 1 <= a ? v.y : 0.0) : (a == b + 1 ? 1.0 : 0.0);
        *reinterpret_cast<double2*>(s_S + t * 64 + 2 * lane) = o;
    }
    for (int a = tid; a < m8; a += NT) s_nu[a] = gy[a];
    if (tid == 0) s_bad = 0;
    __syncthreads();
    const int g = lane >> 2, t4 = lane & 3;
    for (int jb = 0; jb < nb; ++jb) {
        double* dt = s_S + (tri(jb) + jb) * 64;
        if (warp == 0) {
            // diagonal block: lane i < 8 holds row i (lower part) in registers, columns broadcast by shuffle
            double a[8], dvr[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) a[c] = (lane < 8 && c <= lane) ? dt[lane * 8 + c] : 0.0;
            bool bad = false;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const double pv = __shfl_sync(0xffffffffu, a[j], j);
                if (!(pv > 0.0) || !(pv < 1e300)) bad = true;
                const double rs = rsqrt(pv);
                dvr[j] = rs;
                const double lij = a[j] * rs;
                a[j] = lij;
#pragma unroll
                for (int c = j + 1; c < 8; ++c) {
                    const double lcj = __shfl_sync(0xffffffffu, lij, c);
                    a[c] -= lij * lcj;
                }
            }
            if (bad && lane == 0) s_bad = 1;
            if (lane < 8) {
#pragma unroll
                for (int c = 0; c < 8; ++c) dt[lane * 8 + c] = c <= lane ? a[c] : 0.0;
            }
            __syncwarp();
            // inverse of the block: lane c < 8 solves L x = e_c (column c of L^-1)
            if (lane < 8) {
                double x[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    double s = (i == lane) ? 1.0 : 0.0;
#pragma unroll
                    for (int q = 0; q < i; ++q) s -= dt[i * 8 + q] * x[q];
                    x[i] = s * dvr[i];
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) { s_dv[i * 8 + lane] = x[i]; gD[jb * 64 + i * 8 + lane] = x[i]; }
            }
        }
        __syncthreads();
        // panel: X_ib = S(ib, jb) Dinv^T (two DMMAs per tile), y_jb = Dinv nu_jb
        for (int ib = jb + 1 + warp; ib < nb; ib += NW) {
            double* tl = s_S + (tri(ib) + jb) * 64;
            double c0v = 0.0, c1v = 0.0;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk)
                dmma884(c0v, c1v, tl[g * 8 + 4 * kk + t4], s_dv[g * 8 + 4 * kk + t4]);
            __syncwarp();
            *reinterpret_cast<double2*>(tl + g * 8 + 2 * t4) = make_double2(c0v, c1v);
        }
        if (warp == NW - 1 && lane < 8) {
            double s = 0.0;
#pragma unroll
            for (int q = 0; q < 8; ++q) s += s_dv[lane * 8 + q] * s_nu[8 * jb + q];
            __syncwarp(0xff);
            s_nu[8 * jb + lane] = s;
        }
        __syncthreads();
        // trailing: S(ib, kb) -= X_ib X_kb^T for jb < kb <= ib; nu_ib -= X_ib y_jb
        const int rem = nb - jb - 1;
        for (int e = warp; e < tri(rem); e += NW) {
            int ir = (int)((sqrtf(8.0f * e + 1.0f) - 1.0f) * 0.5f);
            while (tri(ir + 1) <= e) ++ir;
            while (tri(ir) > e) --ir;
            const int ib = jb + 1 + ir, kb = jb + 1 + (e - tri(ir));
            const double* xi = s_S + (tri(ib) + jb) * 64;
            const double* xk = s_S + (tri(kb) + jb) * 64;
            double* ct = s_S + (tri(ib) + kb) * 64;
            double2 cv = *reinterpret_cast<double2*>(ct + g * 8 + 2 * t4);
#pragma unroll
            for (int kk = 0; kk < 2; ++kk)
                dmma884(cv.x, cv.y, -xi[g * 8 + 4 * kk + t4], xk[g * 8 + 4 * kk + t4]);
            *reinterpret_cast<double2*>(ct + g * 8 + 2 * t4) = cv;
        }
        for (int a = 8 * (jb + 1) + tid; a < m8; a += NT) {
            const double* xr = s_S + (tri(a >> 3) + jb) * 64 + (a & 7) * 8;
            double s = s_nu[a];
#pragma unroll
            for (int q = 0; q < 8; ++q) s -= xr[q] * s_nu[8 * jb + q];
            s_nu[a] = s;
        }
        __syncthreads();
    }
    // Output for k_wide_panel's forward substitution V_b = Dinv_b G_b - sum_{b' < b} (Dinv_b L_bb') V_b': the
    // off-diagonal tiles leave as  Lt(ib, jb) = -Dinv_ib L(ib, jb)  (two DMMAs per tile), so the panel kernel needs
    // no multiply by Dinv AFTER its dependent chain (one shared-memory round trip and two serial MMAs per block row less)
    for (int t = warp; t < tri(nb); t += NW) {
        int ib = (int)((sqrtf(8.0f * t + 1.0f) - 1.0f) * 0.5f);
        while (tri(ib + 1) <= t) ++ib;
        while (tri(ib) > t) --ib;
        const int jb = t - tri(ib);
        if (ib == jb) continue;   // diagonal blocks are not read downstream (Dinv carries them)
        const double* lt = s_S + t * 64;
        double c0v = 0.0, c1v = 0.0;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
            dmma884(c0v, c1v, -__ldcg(gD + ib * 64 + g * 8 + 4 * kk + t4), lt[(4 * kk + t4) * 8 + g]);
        *reinterpret_cast<double2*>(gL + t * 64 + g * 8 + 2 * t4) = make_double2(c0v, c1v);
    }
    for (int a = tid; a < m8; a += NT) gy[a] = s_nu[a];
    if (tid == 0 && s_bad) p.status[seq] = ASB_ERR_NOT_SPD;
}

// ---- k_wide_panel -------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256, 3) k_wide_panel(WideParams p) {
    using WD = WideDims<KIND>;
    constexpr int D = WD::D, R = WD::R, W = WD::W, M8 = WD::M8, NB = WD::NB, NT = 256;
    extern __shared__ __align__(16) unsigned char smraw[];
    const int seq = p.seq0 + blockIdx.y, jb64 = blockIdx.x;
    const int* meta = p.meta + (long long)seq * 8;
    const int m8 = meta[5];
    if (m8 == 0) return;
    const int n = meta[3];
    const int col0 = 64 * jb64;
    if (col0 >= n) return;
    const int kc = meta[6], c0 = meta[7], m = R * kc, nb = m8 >> 3;
    // L rows for phase 2 come through two shared-memory buffers of one block row ((NB - 1) tiles) each: all 8
    // warps of the CTA walk the same Lt row at the same time, and fetched per warp from L2 the factor alone is
    // 8 x 54 KB per CTA - the kernel was L2-bandwidth-bound on it.  Buffer 0 aliases s_Pc (dead after phase 1).
    constexpr int LROW = WD::LROW;                            // doubles per row buffer (>= s_Pc's 12 x W_VLD)
    double* Gs = reinterpret_cast<double*>(smraw);            // m8 x W_VLD
    double* s_L = Gs + M8 * W_VLD;                            // 2 x LROW
    double* s_Pc = s_L;                                       // 12 x W_VLD: P[cam, cols], rows 10, 11 zero (phase 1)
    int* s_li = reinterpret_cast<int*>(s_L + 2 * LROW);       // WCK
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    const double* P = p.covs + (long long)seq * p.cov_stride;
    const double* gH = p.H + (long long)seq * (WCK * R * W);
    const double* gL = p.L + (long long)seq * (tri(NB) * 64);
    const double* gD = p.Dinv + (long long)seq * (NB * 64);
    const bool add_q = p.round == 0;
    for (int o = tid; o < kc; o += NT) s_li[o] = CAM + D * p.idx[(long long)seq * p.max_obs + c0 + o];
    {   // L2 prefetch of the landmark rows of P this CTA will gather in phase 1: observation tid / 16, its (up to) two
        // row tiles x the eight column tiles of this 64-column block, one 512-byte tile per thread
        const int o = tid >> 4, r = (tid >> 3) & 1, c = tid & 7;
        if (o < kc) {
            const int li = CAM + D * __ldg(p.idx + (long long)seq * p.max_obs + c0 + o);
            const int i8 = (li >> 3) + r, j8 = 8 * jb64 + c;
            if (i8 <= ((li + D - 1) >> 3) && 8 * j8 < n) {
                const long long off = i8 >= j8 ? ((long long)(tri(i8) + j8) << 6) : ((long long)(tri(j8) + i8) << 6);
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], 512;" ::"l"(P + off) : "memory");
            }
        }
    }
    const double* gy = p.y + (long long)seq * M8;
    for (int e = tid; e < 12 * 64; e += NT) {
        const int c = e >> 6, col = col0 + (e & 63);
        double v = 0.0;
        if (c < CAM) {
            v = wt_sym(P, c, col);
            if (add_q && c == col) v += process_noise(p.nz, c);
        }
        s_Pc[c * W_VLD + (e & 63)] = v;
    }
    for (int e = m * W_VLD + tid; e < m8 * W_VLD; e += NT) Gs[e] = 0.0;
    __syncthreads();
    // ---- G = H P[T, cols] on the FP64 tensor pipe: per observation an (R <= 8) x (10 + D) x 64 product,
    //      A = the observation's dense block (rows >= R zero), B = the 10 camera rows of P (shared) and
    //      the D landmark rows (straight from HBM in fragment order: H is never materialised) ----------
    constexpr int KL = (D + 3) / 4;
    // the landmark rows of an item come from HBM in B-fragment order, fetched one item ahead of the MMAs
    auto load_item = [&](int it, double (&blk)[2][KL]) {
        const int o = it >> 2, q = it & 3;
        const int li = s_li[o];
#pragma unroll
        for (int nb8 = 0; nb8 < 2; ++nb8) {
            const int col = col0 + 16 * q + 8 * nb8 + g;
#pragma unroll
            for (int kk = 0; kk < KL; ++kk) {
                const int c = 4 * kk + t4;
                double v = 0.0;
                if (c < D) {
                    v = wt_sym(P, li + c, col);
                    if (add_q && li + c == col) v += p.nz.q_lm;
                }
                blk[nb8][kk] = v;
            }
        }
    };
    double bnext[2][KL];
    if (warp < 4 * kc) load_item(warp, bnext);
    for (int it = warp; it < 4 * kc; it += 8) {      // item = (observation, 16 of the 64 columns)
        const int o = it >> 2, q = it & 3;
        const double* hb = gH + o * R * W;
        double ac[3], al[KL], bl[2][KL];
#pragma unroll
        for (int nb8 = 0; nb8 < 2; ++nb8)
#pragma unroll
            for (int kk = 0; kk < KL; ++kk) bl[nb8][kk] = bnext[nb8][kk];
        if (it + 8 < 4 * kc) load_item(it + 8, bnext);
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) ac[kk] = (g < R && 4 * kk + t4 < CAM) ? __ldg(hb + g * W + 4 * kk + t4) : 0.0;
#pragma unroll
        for (int kk = 0; kk < KL; ++kk) al[kk] = (g < R && 4 * kk + t4 < D) ? __ldg(hb + g * W + CAM + 4 * kk + t4) : 0.0;
#pragma unroll
        for (int nb8 = 0; nb8 < 2; ++nb8) {
            const int jc = 16 * q + 8 * nb8;
            double c0v = 0.0, c1v = 0.0;
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) dmma884(c0v, c1v, ac[kk], s_Pc[(4 * kk + t4) * W_VLD + jc + g]);
#pragma unroll
            for (int kk = 0; kk < KL; ++kk) dmma884(c0v, c1v, al[kk], bl[nb8][kk]);
            if (g < R) *reinterpret_cast<double2*>(Gs + (R * o + g) * W_VLD + jc + 2 * t4) = make_double2(c0v, c1v);
        }
    }
    __syncthreads();
    // ---- V = L^-1 G: each warp owns 8 columns for all rows; block row b needs blocks < b ---------
    double* gs = Gs + 8 * warp;
    double* gV = p.V + (long long)seq * p.v_stride + (long long)jb64 * (M8 * W_VLD) + 8 * warp;
    double dx0 = 0.0, dx1 = 0.0;
    // row b of Lt (tiles (b, 0 .. b-1), contiguous in HBM) -> buffer b & 1, 16 bytes per cp.async
    auto fetch_row = [&](int b) {
        const double* src = gL + (long long)tri(b) * 64;
        double* dst = s_L + (b & 1) * LROW;
        for (int e = tid; e < b * 32; e += NT) w_cp_async16(dst + 2 * e, src + 2 * e);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (nb > 1) fetch_row(1);      // buffer 1; buffer 0 (= s_Pc) is free as of the barrier above
    double nd0 = __ldg(gD + g * 8 + t4), nd1 = __ldg(gD + g * 8 + 4 + t4), ny = __ldg(gy + g);
    for (int b = 0; b < nb; ++b) {
        // V_b = Dinv_b G_b + sum_{b' < b} Lt(b, b') V_b'   (Lt = -Dinv L from k_wide_chol): every product takes its B
        // operand straight from shared memory (G_b as written by phase 1, V_b' as written by the rows before), four
        // independent accumulator chains, one shared-memory store per block row
        if (b >= 1) {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncthreads();       // row b has landed for everyone; everyone is done with the other buffer (row b - 1)
            if (b + 1 < nb) fetch_row(b + 1);
        }
        const double d0 = nd0, d1 = nd1, yv = ny;
        if (b + 1 < nb) {   // Dinv fragments and y of the next block row: fetched a row ahead of their use
            nd0 = __ldg(gD + (b + 1) * 64 + g * 8 + t4); nd1 = __ldg(gD + (b + 1) * 64 + g * 8 + 4 + t4);
            ny = __ldg(gy + 8 * (b + 1) + g);
        }
        double v0 = 0.0, v1 = 0.0, e0 = 0.0, e1 = 0.0, f0 = 0.0, f1 = 0.0, h0 = 0.0, h1 = 0.0;
        const double* lrow = s_L + (b & 1) * LROW + g * 8 + t4;
        dmma884(f0, f1, d0, gs[(8 * b + t4) * W_VLD + g]);
        dmma884(h0, h1, d1, gs[(8 * b + 4 + t4) * W_VLD + g]);
        int tb = 0;
#pragma unroll 2
        for (; tb + 2 <= b; tb += 2) {
            const double* gk = gs + (8 * tb + t4) * W_VLD + g;
            dmma884(v0, v1, lrow[tb * 64], gk[0]);
            dmma884(e0, e1, lrow[tb * 64 + 4], gk[4 * W_VLD]);
            dmma884(f0, f1, lrow[tb * 64 + 64], gk[8 * W_VLD]);
            dmma884(h0, h1, lrow[tb * 64 + 68], gk[12 * W_VLD]);
        }
        if (tb < b) {
            const double* gk = gs + (8 * tb + t4) * W_VLD + g;
            dmma884(v0, v1, lrow[tb * 64], gk[0]);
            dmma884(e0, e1, lrow[tb * 64 + 4], gk[4 * W_VLD]);
        }
        v0 += (e0 + f0) + h0; v1 += (e1 + f1) + h1;
        __syncwarp();   // every lane has read G_b before it is overwritten by V_b
        *reinterpret_cast<double2*>(gs + (8 * b + g) * W_VLD + 2 * t4) = make_double2(v0, v1);
        *reinterpret_cast<double2*>(gV + (8 * b + g) * W_VLD + 2 * t4) = make_double2(v0, v1);
        dx0 += v0 * yv; dx1 += v1 * yv;
        __syncwarp();   // V_b visible to the warp before the next block row reads it
    }
#pragma unroll
    for (int sh = 4; sh < 32; sh <<= 1) {
        dx0 += __shfl_xor_sync(0xffffffffu, dx0, sh);
        dx1 += __shfl_xor_sync(0xffffffffu, dx1, sh);
    }
    if (lane < 4) {
        double* delta = p.delta + (long long)seq * p.state_stride;
        const int col = col0 + 8 * warp + 2 * lane;
        if (col < n) delta[col] += dx0;
        if (col + 1 < n) delta[col + 1] += dx1;
    }
}

// ---- k_wide_sweep -------------------------------------------------------------------------------
// CTA = one 64-row strip I of one sequence: walks the tiles (I, J = 0..I) with its own 64 columns of the
// panel (V_I) resident in shared memory and V_J streamed in W_SLAB-row slabs through a W_NST-stage ring
// (TMA bulk copies, full / empty mbarriers) by a dedicated producer warp, so the ring never drains between
// tiles.  8 consumer warps of 32 x 16 (4 x 2 DMMA tiles); the P sub-tiles go HBM -> registers in the
// accumulator-fragment layout at the start of a tile and are consumed after its k-loop.
constexpr int W_SWEEP_THREADS = 320;   // 8 consumer warps, a producer warp, a prefetch warp (9 warps are allocated as 10 anyway)
// "empty" arrival of a consumer warp.  .relaxed: the default .release semantics make the arrive wait for the
// warp's outstanding global loads (the P sub-tiles prefetched for the end of the tile), which serialises what
// the prefetch is there to overlap; the stage's shared-memory reads have already been consumed by issued MMAs.
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
template <int KIND>
__global__ void __launch_bounds__(W_SWEEP_THREADS, 2) k_wide_sweep(WideParams p) {
    using WD = WideDims<KIND>;
    constexpr int M8 = WD::M8;
    extern __shared__ __align__(128) unsigned char smraw[];
    constexpr int NSL = (M8 + W_SLAB - 1) / W_SLAB;             // slabs of a full panel
    __shared__ __align__(8) uint64_t bar_full[W_NST], bar_empty[W_NST], bar_vi[NSL];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // idle work items leave before the barrier set-up (its init fence costs about a microsecond, and in a later
    // chunk round more than half of the CTAs are idle)
    const int seq = p.seq0 + blockIdx.y;
    const int* meta = p.meta + (long long)seq * 8;
    const int m8 = meta[5];
    if (m8 == 0) return;
    const int n = meta[3], nt8 = (n + 7) >> 3, nb64 = (n + 63) >> 6;
    const int I = (int)gridDim.x - 1 - (int)blockIdx.x;        // heavy strips first
    if (I >= nb64) return;
    if (tid == 0) {
        for (int s = 0; s < W_NST; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 8); }
        for (int s = 0; s < NSL; ++s) mbar_init(&bar_vi[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    constexpr int SLAB_D = W_SLAB * W_VLD;
    double* VI = reinterpret_cast<double*>(smraw);               // M8 x W_VLD
    double* ring = VI + M8 * W_VLD;                              // W_NST x W_SLAB x W_VLD
    double* P = p.covs + (long long)seq * p.cov_stride;
    const double* Vb = p.V + (long long)seq * p.v_stride;
    const int nslab = (m8 + W_SLAB - 1) / W_SLAB;
    __syncthreads();
    if (warp == 8) {   // ---- producer: V_I and V_J slabs interleaved, so the first MMAs need two slabs only ----
        if (lane == 0) {
            const double* VIg = Vb + (long long)I * (M8 * W_VLD);
            int cnt = 0;
            for (int J = 0; J <= I; ++J) {
                const double* VJg = Vb + (long long)J * (M8 * W_VLD);
                for (int s = 0; s < nslab; ++s, ++cnt) {
                    const uint32_t bytes = (uint32_t)(min(W_SLAB, m8 - s * W_SLAB) * W_VLD * sizeof(double));
                    if (J == 0) {
                        mbar_expect_tx(&bar_vi[s], bytes);
                        bulk_g2s(VI + (size_t)s * SLAB_D, VIg + (long long)s * SLAB_D, bytes, &bar_vi[s]);
                    }
                    const int st = cnt % W_NST;
                    if (cnt >= W_NST) mbar_wait(&bar_empty[st], (uint32_t)(((cnt / W_NST) - 1) & 1));
                    mbar_expect_tx(&bar_full[st], bytes);
                    bulk_g2s(ring + (size_t)st * SLAB_D, VJg + (long long)s * SLAB_D, bytes, &bar_full[st]);
                }
            }
        }
        return;
    }
    if (warp == 9) {   // ---- L2 prefetch of the strip's P tiles: row i8 of 8 x 8 tiles is one contiguous run ----
        if (lane < 8) {
            const int i8 = 8 * I + lane;
            if (i8 < nt8)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(P + ((long long)tri(i8) << 6)),
                             "r"((uint32_t)((i8 + 1) * 512))
                             : "memory");
        }
        return;
    }
    // ---- consumers ----
    const int g = lane >> 2, t4 = lane & 3, wi = warp >> 2, wj = warp & 3;
    const bool add_q = p.round == 0;
    const double* vi = VI + 32 * wi + g;
    int cnt = 0;
    for (int J = 0; J <= I; ++J) {
        bool valid[4][2];
        double2 pv[4][2];
        double acc[4][2][2];
        bool anyv = false, allv = true;
#pragma unroll
        for (int rm = 0; rm < 4; ++rm)
#pragma unroll
            for (int rn = 0; rn < 2; ++rn) {
                const int i8 = 8 * I + 4 * wi + rm, j8 = 8 * J + 2 * wj + rn;
                valid[rm][rn] = i8 < nt8 && j8 <= i8;
                anyv = anyv || valid[rm][rn];
                allv = allv && valid[rm][rn];
                acc[rm][rn][0] = acc[rm][rn][1] = 0.0;
                pv[rm][rn] = make_double2(0.0, 0.0);
                if (valid[rm][rn])
                    pv[rm][rn] = *reinterpret_cast<const double2*>(P + ((long long)(tri(i8) + j8) << 6) + 2 * lane);
            }
        for (int s = 0; s < nslab; ++s, ++cnt) {
            const int st = cnt % W_NST;
            if (J == 0) mbar_wait(&bar_vi[s], 0);
            mbar_wait(&bar_full[st], (uint32_t)((cnt / W_NST) & 1));
            if (allv) {   // interior warp tile: no per-MMA predicate (each costs a WARPSYNC in SASS)
                const double* vj = ring + (size_t)st * SLAB_D + 16 * wj + g;
                const double* via = vi + (size_t)s * SLAB_D;
                const int ksteps = min(W_SLAB, m8 - s * W_SLAB) >> 2;
#pragma unroll 2
                for (int kk = 0; kk < ksteps; ++kk) {
                    double a[4], b[2];
#pragma unroll
                    for (int rm = 0; rm < 4; ++rm) a[rm] = via[(4 * kk + t4) * W_VLD + 8 * rm];
#pragma unroll
                    for (int rn = 0; rn < 2; ++rn) b[rn] = vj[(4 * kk + t4) * W_VLD + 8 * rn];
#pragma unroll
                    for (int rm = 0; rm < 4; ++rm)
#pragma unroll
                        for (int rn = 0; rn < 2; ++rn) dmma884(acc[rm][rn][0], acc[rm][rn][1], a[rm], b[rn]);
                }
            } else if (anyv) {
                const double* vj = ring + (size_t)st * SLAB_D + 16 * wj + g;
                const double* via = vi + (size_t)s * SLAB_D;
                const int ksteps = min(W_SLAB, m8 - s * W_SLAB) >> 2;
#pragma unroll 2
                for (int kk = 0; kk < ksteps; ++kk) {
                    double a[4], b[2];
#pragma unroll
                    for (int rm = 0; rm < 4; ++rm) a[rm] = via[(4 * kk + t4) * W_VLD + 8 * rm];
#pragma unroll
                    for (int rn = 0; rn < 2; ++rn) b[rn] = vj[(4 * kk + t4) * W_VLD + 8 * rn];
#pragma unroll
                    for (int rm = 0; rm < 4; ++rm)
#pragma unroll
                        for (int rn = 0; rn < 2; ++rn)
                            if (valid[rm][rn]) dmma884(acc[rm][rn][0], acc[rm][rn][1], a[rm], b[rn]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_empty[st]);
        }
#pragma unroll
        for (int rm = 0; rm < 4; ++rm)
#pragma unroll
            for (int rn = 0; rn < 2; ++rn) {
                if (!valid[rm][rn]) continue;
                const int i8 = 8 * I + 4 * wi + rm, j8 = 8 * J + 2 * wj + rn;
                double o0 = pv[rm][rn].x - acc[rm][rn][0], o1 = pv[rm][rn].y - acc[rm][rn][1];
                if (add_q && i8 == j8) {
                    const int row = 8 * i8 + g;
                    if (row < n) {
                        if (g == 2 * t4) o0 += process_noise(p.nz, row);
                        if (g == 2 * t4 + 1) o1 += process_noise(p.nz, row);
                    }
                }
                *reinterpret_cast<double2*>(P + ((long long)(tri(i8) + j8) << 6) + 2 * lane) = make_double2(o0, o1);
            }
    }
}

// Final state update of a run (the pending delta of the last frame).
template <int KIND>
__global__ void k_wide_apply(WideParams p) {
    constexpr int D = Dims<KIND>::D;
    const int seq = p.seq0 + blockIdx.x, tid = threadIdx.x;
    int* meta = p.meta + (long long)seq * 8;
    if (!meta[4]) return;
    double* state = p.states + (long long)seq * p.state_stride;
    const double* delta = p.delta + (long long)seq * p.state_stride;
    const int nlm = p.nlm[seq];
    if (tid == 0) {
        double cam[CAM], d[CAM];
#pragma unroll
        for (int c = 0; c < CAM; ++c) { cam[c] = state[c]; d[c] = delta[c]; }
        camera_reset(cam, d, p.as_written);
#pragma unroll
        for (int c = 0; c < CAM; ++c) state[c] = cam[c];
    }
    for (int l = tid; l < nlm; l += blockDim.x) {
        double lm[D], d[D];
#pragma unroll
        for (int c = 0; c < D; ++c) { lm[c] = state[CAM + D * l + c]; d[c] = delta[CAM + D * l + c]; }
        landmark_update<KIND>(lm, d);
#pragma unroll
        for (int c = 0; c < D; ++c) state[CAM + D * l + c] = lm[c];
    }
    __syncthreads();
    if (tid == 0) meta[4] = 0;
}

// tiled covariance -> dense row-major n x n
__global__ void k_unpack_tiled(const double* __restrict__ tiled, int n, double* __restrict__ dense) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)n * n) return;
    const int i = (int)(e / n), j = (int)(e % n);
    dense[e] = wt_sym(tiled, i, j);
}

}  // namespace asb
