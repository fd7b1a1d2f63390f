// k_sweep: the HBM-bound rank-m covariance sweep  P <- P - V^T V  (= (I - K H) P,
// extended_kalman_filter.py:155-156 / ekf_with_rotations.py:180-181) of the single-filter
// engine.  V = L^-1 H P is m4 x N (row a contiguous over the N state columns, leading
// dimension ld), P is N x N row-major with the same ld.  Only the lower triangle is read;
// every result is stored to (i, j) and mirrored to (j, i), so P stays exactly symmetric.
//
// Algorithmic traffic: one read + one write of P = 16 N^2 bytes per frame;
// actual DRAM traffic ~ 8 N^2 / 2 read + 8 N^2 write.  Flops: N^2 m (triangle only).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "model.cuh"

namespace asb {

// Work folded into the sweep launch so that a frame is two launches (k_front, k_sweep):
//   predict   P += diag(Q) (EKF:95-105) is added on the diagonal tiles as they stream by
//             (k_front already used P + diag(Q) in registers);
//   apply     the state update + multiplicative quaternion reset (EKF:134-152, ROT:143-177)
//             is done by CTA 0 before it starts on its tiles.
struct SweepExtra {
    int n;                 // live state dims (diagonal entries i < n receive Q)
    int add_q;
    Noise nz;
    int apply_kind;        // -1: none, 0: EKF, 1: EKF_Rotations
    int as_written;
    int nlm;
    double* state;
    const double* dx;
};

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

// Layout of V in global memory ("tiled panel"): blocks of V_BLK = 64 state columns; block b is
// m4 rows x V_LDB = 68 doubles (64 values + 4 pad), contiguous, so one bulk copy (TMA, SASS
// UBLKCP) of m4 * 544 bytes brings a whole 64-column panel into shared memory already in
// its bank-conflict-free layout: V_LDB = 4 mod 16, so the DMMA fragment loads
// (address = (4 kk + t) * V_LDB + g) hit 32 distinct banks.  k_solve writes this layout.
constexpr int V_BLK = 64, V_LDB = 68;
__host__ __device__ inline size_t v_index(int m4, int a, int j) {
    return (size_t)(j / V_BLK) * m4 * V_LDB + (size_t)a * V_LDB + (j % V_BLK);
}

// Tile shape: WM x WN warps, each owning (8 RM) x (8 RN) elements as RM x RN DMMA tiles.
template <int WM_, int WN_, int RM_, int RN_, int MINB_>
struct SweepCfg {
    static constexpr int WM = WM_, WN = WN_, RM = RM_, RN = RN_, MINB = MINB_;
    static constexpr int TI = WM * RM * 8, TJ = WN * RN * 8, THREADS = WM * WN * 32;
    static constexpr int RATIO = TI / TJ;             // column tiles per row tile on the diagonal
    static_assert(TJ == V_BLK && TI % V_BLK == 0, "tiles are made of whole V blocks");
    __host__ __device__ static size_t smem_bytes(int m4) {
        return (size_t)m4 * V_LDB * (RATIO + 2) * sizeof(double);
    }
    // lower-triangle tiles of an n x n matrix: row block I has RATIO (I + 1) column tiles
    __host__ __device__ static int num_tiles(int n) {
        const int nI = (n + TI - 1) / TI;
        return RATIO * nI * (nI + 1) / 2;
    }
};
using SweepBig = SweepCfg<4, 2, 4, 4, 1>;    // 128 x 64 tile, 1 CTA / SM (large m4)
using SweepDual = SweepCfg<2, 4, 4, 2, 2>;   // 64 x 64 tile, 2 CTAs / SM: one computes while the other drains

// Persistent kernel.  Each CTA walks a contiguous chunk of the tile list (row-block major),
// so the TI-column panel VI stays in shared memory while the TJ-column panel VJ is
// double-buffered: the bulk copies (TMA, SASS UBLKCP) of tile t+1's VJ fly during tile t's
// MMAs.  The P sub-tile goes global -> registers in the accumulator-fragment layout at the
// start of the tile and is only consumed after the k-loop (out = P - V_I^T V_J), so its HBM
// latency hides behind the tile's own math.
template <class C>
__global__ void __launch_bounds__(C::THREADS, C::MINB)
k_sweep(double* __restrict__ P, int ld, const double* __restrict__ Vt, int m4, int ntiles, int dbg,
        SweepExtra ex) {
    // dbg (asb_bench_sweep only; 0 in the product path): 1 skip MMAs, 2 skip mirror stores,
    // 4 skip direct stores, 8 skip P loads
    constexpr int RM = C::RM, RN = C::RN, TI = C::TI, TJ = C::TJ;
    extern __shared__ __align__(128) unsigned char smraw[];
    const int blk = m4 * V_LDB;                      // doubles per 64-column block of V
    double* VI = reinterpret_cast<double*>(smraw);   // RATIO blocks
    double* VJ0 = VI + (size_t)C::RATIO * blk;       // 2 x 1 block
    __shared__ __align__(8) uint64_t bar[3];         // [0] VI, [1] VJ buf 0, [2] VJ buf 1
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    const int wi = warp / C::WN, wj = warp % C::WN;
    if (blockIdx.x == 0 && ex.apply_kind >= 0) {  // state update (k_apply folded in)
        if (tid == 0) {
            double cam[CAM], d[CAM];
#pragma unroll
            for (int c = 0; c < CAM; ++c) { cam[c] = ex.state[c]; d[c] = ex.dx[c]; }
            camera_reset(cam, d, ex.as_written);
#pragma unroll
            for (int c = 0; c < CAM; ++c) ex.state[c] = cam[c];
        }
        if (ex.apply_kind == 0) {
            for (int e = tid; e < 3 * ex.nlm; e += C::THREADS) ex.state[CAM + e] += ex.dx[CAM + e];
        } else {
            for (int t = tid; t < ex.nlm; t += C::THREADS) {
                double lm[10], d[10];
#pragma unroll
                for (int c = 0; c < 10; ++c) { lm[c] = ex.state[CAM + 10 * t + c]; d[c] = ex.dx[CAM + 10 * t + c]; }
                landmark_update<1>(lm, d);
#pragma unroll
                for (int c = 0; c < 10; ++c) ex.state[CAM + 10 * t + c] = lm[c];
            }
        }
    }
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
    // t = RATIO I (I + 1) / 2 + J
    int I = (int)((sqrt(8.0 * t_begin / C::RATIO + 1.0) - 1.0) * 0.5);
    while (C::RATIO * (I + 1) * (I + 2) / 2 <= t_begin) ++I;
    while (C::RATIO * I * (I + 1) / 2 > t_begin) --I;
    int J = t_begin - C::RATIO * I * (I + 1) / 2;

    const uint32_t blk_bytes = (uint32_t)(blk * sizeof(double));
    auto load_vi = [&](int i0) {  // one thread
        mbar_expect_tx(&bar[0], C::RATIO * blk_bytes);
        bulk_g2s(VI, Vt + (size_t)(i0 / V_BLK) * blk, C::RATIO * blk_bytes, &bar[0]);
    };
    auto load_vj = [&](int buf, int j0) {  // one thread
        mbar_expect_tx(&bar[1 + buf], blk_bytes);
        bulk_g2s(VJ0 + (size_t)buf * blk, Vt + (size_t)(j0 / V_BLK) * blk, blk_bytes, &bar[1 + buf]);
    };

    uint32_t ph_i = 0, ph_j = 0;  // ph_j: bit b = phase of VJ buffer b
    bool vi_pending = true;
    if (tid == 0) {
        load_vi(I * TI);
        load_vj(0, J * TJ);
    }
    int buf = 0;
    for (int t = t_begin; t < t_end; ++t, buf ^= 1) {
        const int i0 = I * TI, j0 = J * TJ;
        int In = I, Jn = J + 1;
        if (Jn == C::RATIO * (I + 1)) { In = I + 1; Jn = 0; }
        const bool has_next = t + 1 < t_end;
        if (tid == 0 && has_next) load_vj(buf ^ 1, Jn * TJ);  // buffer freed by the last __syncthreads

        // P sub-tile -> registers (accumulator-fragment layout)
        const int ri = i0 + 8 * RM * wi + g, cj = j0 + 8 * RN * wj + 2 * t4;
        double2 pf[RM][RN];
#pragma unroll
        for (int rm = 0; rm < RM; ++rm)
#pragma unroll
            for (int rn = 0; rn < RN; ++rn)
                pf[rm][rn] = (dbg & 8) ? make_double2(0.0, 0.0)
                                       : *reinterpret_cast<const double2*>(P + (long long)(ri + 8 * rm) * ld + cj + 8 * rn);

        double acc[RM][RN][2];
#pragma unroll
        for (int rm = 0; rm < RM; ++rm)
#pragma unroll
            for (int rn = 0; rn < RN; ++rn) acc[rm][rn][0] = acc[rm][rn][1] = 0.0;

        if (vi_pending) {
            mbar_wait(&bar[0], ph_i);
            ph_i ^= 1;
            vi_pending = false;
        }
        mbar_wait(&bar[1 + buf], (ph_j >> buf) & 1u);
        ph_j ^= 1u << buf;

        const double* va = VI + (size_t)((8 * RM * wi) / V_BLK) * blk + t4 * V_LDB + (8 * RM * wi) % V_BLK + g;
        const double* vb = VJ0 + (size_t)buf * blk + t4 * V_LDB + 8 * RN * wj + g;
#pragma unroll 3
        for (int kk = 0; kk < ((dbg & 1) ? 4 : m4); kk += 4) {
            double a[RM], b[RN];
#pragma unroll
            for (int r = 0; r < RM; ++r) a[r] = va[kk * V_LDB + 8 * r];
#pragma unroll
            for (int r = 0; r < RN; ++r) b[r] = vb[kk * V_LDB + 8 * r];
#pragma unroll
            for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                for (int rn = 0; rn < RN; ++rn) dmma884(acc[rm][rn][0], acc[rm][rn][1], a[rm], b[rn]);
        }

        const bool interior = i0 >= j0 + TJ;  // strictly below the diagonal
        if (interior) {
#pragma unroll
            for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                for (int rn = 0; rn < RN; ++rn) {
                    const double o0 = pf[rm][rn].x - acc[rm][rn][0], o1 = pf[rm][rn].y - acc[rm][rn][1];
                    const int i = ri + 8 * rm, j = cj + 8 * rn;
                    if (!(dbg & 4)) *reinterpret_cast<double2*>(P + (long long)i * ld + j) = make_double2(o0, o1);
                    if (!(dbg & 2)) {
                        P[(long long)j * ld + i] = o0;
                        P[(long long)(j + 1) * ld + i] = o1;
                    }
                }
        } else {
#pragma unroll
            for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                for (int rn = 0; rn < RN; ++rn) {
                    double o[2] = {pf[rm][rn].x - acc[rm][rn][0], pf[rm][rn].y - acc[rm][rn][1]};
                    const int i = ri + 8 * rm;
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int j = cj + 8 * rn + e;
                        if (i >= j) {
                            if (i == j && ex.add_q && i < ex.n) o[e] += process_noise(ex.nz, i);
                            P[(long long)i * ld + j] = o[e];
                            if (i > j) P[(long long)j * ld + i] = o[e];
                        }
                    }
                }
        }
        __syncthreads();  // every warp is done with VJ[buf] (and VI) before they are overwritten
        if (has_next && In != I) {
            if (tid == 0) load_vi(In * TI);
            vi_pending = true;
        }
        I = In;
        J = Jn;
    }
}

}  // namespace asb
