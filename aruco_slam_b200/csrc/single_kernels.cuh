// Kernels of the single-filter engine: one filter whose covariance P lives in HBM
// (BASELINE config C3: 2000 markers, N = 6010, P = 289 MB; also C1 and any N).
//
// Per observe() (extended_kalman_filter.py:58-82) the engine launches, per chunk of markers,
//   k_front   (front.cuh)  model, S, Cholesky, V = L^-1 H P, dx — with augmentation and
//                          predict folded in
//   k_sweep   (sweep.cuh)  P <- P + diag(Q) - V^T V, and the state update / quaternion reset.
// This file holds what is left for predict() / add_marker() called on their own:
//   k_prep    augmentation (add_marker, EKF:239-290) + predict (EKF:95-105)
#pragma once
#include <cstdint>

#include "../../include/aruco_slam_b200.h"
#include "model.cuh"
#include "sweep.cuh"

namespace asb {

constexpr int PAD = 128;       // leading dimension / row padding of P and V (doubles)
constexpr int M_MAX = 68;      // max (padded) measurement rows per chunk (front.cuh FrontLimits)

// ------------------------------------------------------------------------------------
// k_prep
// ------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256)
k_prep(double* __restrict__ state, double* __restrict__ P, int ld, int n0, int n1, Noise nz,
       const int* __restrict__ idx, const double* __restrict__ poses,
       const unsigned char* __restrict__ isnew, int k, int do_predict) {
    constexpr int D = Dims<KIND>::D;
    // ---- CTA 0: landmark initialisation -------------------------------------------------
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

}  // namespace asb
