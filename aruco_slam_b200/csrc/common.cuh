// Host-side helpers shared by the translation units of libarucoslam_b200.so.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/aruco_slam_b200.h"
#include "model.cuh"

namespace asb {

// thread-local message behind asb_last_error()
void set_error(const char* fmt, ...);
const char* get_error();

inline int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    set_error("%s", buf);
    return code;
}

#define ASB_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess)                                                           \
            return ::asb::fail(ASB_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr, \
                               cudaGetErrorString(_e));                                  \
    } while (0)

inline Noise default_noise(int kind) {
    // extended_kalman_filter.py:21-27 / ekf_with_rotations.py:23-31
    Noise n;
    n.init_cam = 0.1;
    n.init_lm = 0.7;
    n.r = 0.9;
    n.q_cam = kind == ASB_KIND_ROT ? 0.2 : 0.3;
    n.q_err = 0.5;
    n.q_lm = 0.01;
    return n;
}

inline Noise to_noise(int kind, const asb_noise* cfg) {
    if (!cfg) return default_noise(kind);
    Noise n;
    n.init_cam = cfg->init_cam;
    n.init_lm = cfg->init_lm;
    n.r = cfg->r;
    n.q_cam = cfg->q_cam;
    n.q_err = cfg->q_err;
    n.q_lm = cfg->q_lm;
    return n;
}

// One observe() of a single filter whose whole covariance fits an SM's shared memory: ONE launch of the resident
// batched kernel's SINGLE instantiation (batch_kernels.cuh) instead of the pipelined k_f1 / k_f2 / k_sweep.
// Defined in filter_batch.cu (the translation unit that instantiates k_batch), used by filter_single.cu.
struct SmallFrame {
    int kind, as_written;
    double* state; double* P; int ld;           // the filter's dense covariance (lower triangle valid)
    const int* idx; const unsigned char* isnew; const double* poses; int k;   // this frame, on the device
    int nlm0, nlm1;                              // landmarks before / after this frame's first sightings
    Noise nz;
    int* status; double* nu_out; double* blocks_out; double* dx_out;
    double* host_state; int* host_status;        // page-locked mirror the kernel writes the posterior state into (or null)
};
size_t small_frame_smem(int kind, int nlm1, int k);     // dynamic shared memory the launch needs
int small_frame_launch(const SmallFrame& a, cudaStream_t st);

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline size_t round_up16(size_t v) { return (v + 15) & ~size_t(15); }

}  // namespace asb
