// Batched square-marker pose estimation: one thread per corner set, FP64.
// Replaces the per-marker cv2.solvePnP(marker_points, c, K, dist,
// flags=cv2.SOLVEPNP_IPPE_SQUARE) loop of BaseFilter.estimate_pose_of_markers
// (filters/base_filter.py:125-171).  Algorithm: SURVEY.md §8a row A9 —
//   (1) undistortPoints: 5 fixed-point iterations of the inverse Brown-Conrady model in
//       FP64, result ROUNDED TO float32 (OpenCV's output takes the input depth).  Written
//       with explicit round-to-nearest intrinsics so no FMA contraction changes the value
//       that gets rounded;
//   (2) exact homography canonical square -> 4 normalised points (H22 = 1);
//   (3) IPPE: the two rotations from the homography Jacobian at the origin;
//   (4) least-squares translation for each rotation (3x3 normal equations);
//   (5) order by float reprojection error; (6) Rodrigues vector of the better one.
#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace asb {

struct Intrinsics {
    double fx, fy, cx, cy, k1, k2, p1, p2, k3;
};

__device__ __forceinline__ double mul_(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_(double a, double b) { return __dadd_rn(a, b); }

__device__ __forceinline__ void undistort1(float u, float v, const Intrinsics& in, double ifx, double ify,
                                           double& xo, double& yo) {
    const double x0 = mul_(add_((double)u, -in.cx), ifx);
    const double y0 = mul_(add_((double)v, -in.cy), ify);
    double x = x0, y = y0;
#pragma unroll 1
    for (int it = 0; it < 5; ++it) {
        const double r2 = add_(mul_(x, x), mul_(y, y));
        const double den = add_(1.0, mul_(add_(mul_(add_(mul_(in.k3, r2), in.k2), r2), in.k1), r2));
        const double icd = __ddiv_rn(1.0, den);
        if (icd < 0) { x = x0; y = y0; break; }
        const double dx = add_(mul_(mul_(mul_(2.0, in.p1), x), y), mul_(in.p2, add_(r2, mul_(mul_(2.0, x), x))));
        const double dy = add_(mul_(in.p1, add_(r2, mul_(mul_(2.0, y), y))), mul_(mul_(mul_(2.0, in.p2), x), y));
        x = mul_(add_(x0, -dx), icd);
        y = mul_(add_(y0, -dy), icd);
    }
    xo = (double)__double2float_rn(x);
    yo = (double)__double2float_rn(y);
}

struct Pose {
    double r[3][3];
    double t[3];
    float err;
};

__device__ __forceinline__ void solve_translation(Pose& p, const double* x, const double* y, double half) {
    const double ox[4] = {-half, half, half, -half}, oy[4] = {half, half, -half, -half};
    double a02 = 0, a12 = 0, a22 = 0, b0 = 0, b1 = 0, b2 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double rx = p.r[0][0] * ox[i] + p.r[0][1] * oy[i];
        const double ry = p.r[1][0] * ox[i] + p.r[1][1] * oy[i];
        const double rz = p.r[2][0] * ox[i] + p.r[2][1] * oy[i];
        const double a2 = -x[i], bb2 = -y[i];
        a02 += a2; a12 += bb2; a22 += a2 * a2 + bb2 * bb2;
        const double bx = -a2 * rz - rx, by = -bb2 * rz - ry;
        b0 += bx; b1 += by; b2 += a2 * bx + bb2 * by;
    }
    const double n = 4.0;
    const double det_inv = 1.0 / (n * n * a22 - n * a12 * a12 - a02 * n * a02);
    const double s00 = n * a22 - a12 * a12, s01 = a02 * a12, s02 = -a02 * n;
    const double s10 = a12 * a02, s11 = n * a22 - a02 * a02, s12 = -n * a12;
    const double s20 = -n * a02, s21 = -n * a12, s22 = n * n;
    p.t[0] = det_inv * (s00 * b0 + s01 * b1 + s02 * b2);
    p.t[1] = det_inv * (s10 * b0 + s11 * b1 + s12 * b2);
    p.t[2] = det_inv * (s20 * b0 + s21 * b1 + s22 * b2);
}

__device__ __forceinline__ void reproj_error(Pose& p, const double* x, const double* y, double half) {
    const double ox[4] = {-half, half, half, -half}, oy[4] = {half, half, -half, -half};
    float err = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double px = p.r[0][0] * ox[i] + p.r[0][1] * oy[i] + p.t[0];
        const double py = p.r[1][0] * ox[i] + p.r[1][1] * oy[i] + p.t[1];
        const double pz = p.r[2][0] * ox[i] + p.r[2][1] * oy[i] + p.t[2];
        const float dx = (float)(px / pz - x[i]);
        const float dy = (float)(py / pz - y[i]);
        err = __fadd_rn(err, __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
    }
    p.err = err;
}

__device__ __forceinline__ void rodrigues(const double r[3][3], double* out) {
    const double rx = r[2][1] - r[1][2], ry = r[0][2] - r[2][0], rz = r[1][0] - r[0][1];
    const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double c = (r[0][0] + r[1][1] + r[2][2] - 1.0) * 0.5;
    c = fmin(1.0, fmax(-1.0, c));
    const double theta = acos(c);
    if (s < 1e-5) {
        if (c > 0) {
            out[0] = out[1] = out[2] = 0.0;
        } else {
            double t0 = sqrt(fmax((r[0][0] + 1) * 0.5, 0.0));
            double t1 = sqrt(fmax((r[1][1] + 1) * 0.5, 0.0)) * (r[0][1] < 0 ? -1.0 : 1.0);
            double t2 = sqrt(fmax((r[2][2] + 1) * 0.5, 0.0)) * (r[0][2] < 0 ? -1.0 : 1.0);
            if (fabs(t0) < fabs(t1) && fabs(t0) < fabs(t2) && ((r[1][2] > 0) != (t1 * t2 > 0))) t2 = -t2;
            const double sc = theta / sqrt(t0 * t0 + t1 * t1 + t2 * t2);
            out[0] = t0 * sc; out[1] = t1 * sc; out[2] = t2 * sc;
        }
    } else {
        const double vth = theta / (2.0 * s);
        out[0] = rx * vth; out[1] = ry * vth; out[2] = rz * vth;
    }
}

__global__ void __launch_bounds__(128)
k_ippe_square(const float* __restrict__ corners, long long n, Intrinsics in, double half,
              double* __restrict__ out, double* __restrict__ out2, float* __restrict__ errs) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 c01 = __ldg(reinterpret_cast<const float4*>(corners) + 2 * i);
    const float4 c23 = __ldg(reinterpret_cast<const float4*>(corners) + 2 * i + 1);
    const double ifx = __ddiv_rn(1.0, in.fx), ify = __ddiv_rn(1.0, in.fy);
    double x[4], y[4];
    undistort1(c01.x, c01.y, in, ifx, ify, x[0], y[0]);
    undistort1(c01.z, c01.w, in, ifx, ify, x[1], y[1]);
    undistort1(c23.x, c23.y, in, ifx, ify, x[2], y[2]);
    undistort1(c23.z, c23.w, in, ifx, ify, x[3], y[3]);

    // homography: unit square (0,0),(1,0),(1,1),(0,1) <- corners 3,2,1,0, then the affine
    // pre-scale (X,Y) -> (X/(2h) + 1/2, Y/(2h) + 1/2); normalised so that H22 = 1
    const double x0 = x[3], x1 = x[2], x2 = x[1], x3 = x[0];
    const double y0 = y[3], y1 = y[2], y2 = y[1], y3 = y[0];
    const double sx = x0 - x1 + x2 - x3, sy = y0 - y1 + y2 - y3;
    const double dx1 = x1 - x2, dx2 = x3 - x2, dy1 = y1 - y2, dy2 = y3 - y2;
    const double den = dx1 * dy2 - dy1 * dx2;
    const double g = (sx * dy2 - sy * dx2) / den;
    const double hh = (dx1 * sy - dy1 * sx) / den;
    const double a = x1 - x0 + g * x1, b = x3 - x0 + hh * x3, c = x0;
    const double d = y1 - y0 + g * y1, e = y3 - y0 + hh * y3, f = y0;
    const double i2h = 1.0 / (2.0 * half);
    const double h22 = 0.5 * (g + hh) + 1.0;
    const double H00 = a * i2h / h22, H01 = b * i2h / h22, H02 = (0.5 * (a + b) + c) / h22;
    const double H10 = d * i2h / h22, H11 = e * i2h / h22, H12 = (0.5 * (d + e) + f) / h22;
    const double H20 = g * i2h / h22, H21 = hh * i2h / h22;

    // IPPE
    const double p = H02, q = H12;
    const double j00 = H00 - H20 * p, j01 = H01 - H21 * p, j10 = H10 - H20 * q, j11 = H11 - H21 * q;
    const double s = sqrt(p * p + q * q + 1.0);
    const double ax = p / s, ay = q / s, az = 1.0 / s;
    const double w = 1.0 / (1.0 + az);
    double rv[3][3];
    rv[0][0] = 1.0 - ax * ax * w; rv[0][1] = -ax * ay * w;      rv[0][2] = ax;
    rv[1][0] = -ax * ay * w;      rv[1][1] = 1.0 - ay * ay * w; rv[1][2] = ay;
    rv[2][0] = -ax;               rv[2][1] = -ay;               rv[2][2] = az;
    const double b00 = rv[0][0] - p * rv[2][0], b01 = rv[0][1] - p * rv[2][1];
    const double b10 = rv[1][0] - q * rv[2][0], b11 = rv[1][1] - q * rv[2][1];
    const double dtinv = 1.0 / (b00 * b11 - b01 * b10);
    const double bi00 = dtinv * b11, bi01 = -dtinv * b01, bi10 = -dtinv * b10, bi11 = dtinv * b00;
    const double a00 = bi00 * j00 + bi01 * j10, a01 = bi00 * j01 + bi01 * j11;
    const double a10 = bi10 * j00 + bi11 * j10, a11 = bi10 * j01 + bi11 * j11;
    const double ata00 = a00 * a00 + a01 * a01, ata01 = a00 * a10 + a01 * a11, ata11 = a10 * a10 + a11 * a11;
    const double gamma = sqrt(0.5 * (ata00 + ata11 + sqrt((ata00 - ata11) * (ata00 - ata11) + 4.0 * ata01 * ata01)));
    const double r00 = a00 / gamma, r01 = a01 / gamma, r10 = a10 / gamma, r11 = a11 / gamma;
    const double bb0 = sqrt(fmax(1.0 - r00 * r00 - r10 * r10, 0.0));
    double bb1 = sqrt(fmax(1.0 - r01 * r01 - r11 * r11, 0.0));
    if (-r00 * r01 - r10 * r11 < 0) bb1 = -bb1;

    Pose sol[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const double sg = k == 0 ? 1.0 : -1.0;
        const double c0[3] = {r00, r10, sg * bb0}, c1[3] = {r01, r11, sg * bb1};
        const double c2[3] = {c0[1] * c1[2] - c0[2] * c1[1], c0[2] * c1[0] - c0[0] * c1[2],
                              c0[0] * c1[1] - c0[1] * c1[0]};
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            sol[k].r[r][0] = rv[r][0] * c0[0] + rv[r][1] * c0[1] + rv[r][2] * c0[2];
            sol[k].r[r][1] = rv[r][0] * c1[0] + rv[r][1] * c1[1] + rv[r][2] * c1[2];
            sol[k].r[r][2] = rv[r][0] * c2[0] + rv[r][1] * c2[1] + rv[r][2] * c2[2];
        }
        solve_translation(sol[k], x, y, half);
        reproj_error(sol[k], x, y, half);
    }
    const int first = (sol[0].err < sol[1].err) ? 0 : 1;
    double rvec[3];
    rodrigues(sol[first].r, rvec);
    double* o = out + 6 * i;
    o[0] = sol[first].t[0]; o[1] = sol[first].t[1]; o[2] = sol[first].t[2];
    o[3] = rvec[0]; o[4] = rvec[1]; o[5] = rvec[2];
    if (out2) {
        rodrigues(sol[1 - first].r, rvec);
        double* o2 = out2 + 6 * i;
        o2[0] = sol[1 - first].t[0]; o2[1] = sol[1 - first].t[1]; o2[2] = sol[1 - first].t[2];
        o2[3] = rvec[0]; o2[4] = rvec[1]; o2[5] = rvec[2];
    }
    if (errs) {
        errs[2 * i] = sol[first].err;
        errs[2 * i + 1] = sol[1 - first].err;
    }
}

static Intrinsics make_intrinsics(const double km[9], const double dist[5]) {
    Intrinsics in;
    in.fx = km[0]; in.fy = km[4]; in.cx = km[2]; in.cy = km[5];
    in.k1 = dist ? dist[0] : 0; in.k2 = dist ? dist[1] : 0; in.p1 = dist ? dist[2] : 0;
    in.p2 = dist ? dist[3] : 0; in.k3 = dist ? dist[4] : 0;
    return in;
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_ippe_square_device(const float* d_corners, int64_t n, const double camera_matrix[9],
                           const double dist_coeffs[5], double marker_size, double* d_out,
                           double* d_out2, float* d_err, void* cuda_stream) {
    if (n < 0 || !camera_matrix || (n > 0 && (!d_corners || !d_out)))
        return fail(ASB_ERR_BAD_ARG, "bad ippe arguments");
    if (n == 0) return ASB_OK;
    // object points are float32 in the reference (BASE:112-120): half = float32(size/2)
    const double half = (double)(float)(marker_size / 2.0);
    const Intrinsics in = make_intrinsics(camera_matrix, dist_coeffs);
    const int block = 128;
    const long long grid = (n + block - 1) / block;
    k_ippe_square<<<(unsigned)grid, block, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
        d_corners, n, in, half, d_out, d_out2, d_err);
    ASB_CUDA(cudaGetLastError());
    return ASB_OK;
}

// host buffers in, host buffers out; out2 / errs may be null (second solution, float reprojection errors)
static int ippe_host(const float* corners, int64_t n, const double camera_matrix[9], const double dist_coeffs[5],
                     double marker_size, double* out, double* out2, float* errs, int device) {
    if (n < 0 || !camera_matrix || (n > 0 && (!corners || !out))) return fail(ASB_ERR_BAD_ARG, "bad ippe arguments");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(ASB_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (n == 0) return ASB_OK;
    ASB_CUDA(cudaSetDevice(device));
    // Device buffers are kept between calls (grow-only, per device): the per-frame caller
    // (BaseFilter.estimate_pose_of_markers, a few dozen markers) must not pay cudaMalloc /
    // cudaFree round trips per frame.
    struct Cache { float* dc = nullptr; double* dout = nullptr; double* dout2 = nullptr; float* derr = nullptr; int64_t cap = 0;
                   cudaStream_t st[3] = {nullptr, nullptr, nullptr};
                   cudaEvent_t ev_in[64] = {}, ev_k[64] = {}; };
#ifndef ASB_IPPE_CHUNK
#define ASB_IPPE_CHUNK 32768
#endif
#ifndef ASB_IPPE_CHUNK_MAX
#define ASB_IPPE_CHUNK_MAX 32768
#endif
    // slice sizes of the large-batch path (first slice, doubling up to the maximum).  Measured on the B200 box for
    // 1M sets with page-locked buffers (tools/ippe_e2e_exp.py; both copies alone, concurrently: 1.01-1.04 ms):
    // 32768 fixed 1.16-1.20 ms, 16384 doubling to 131072 / 262144 1.20-1.23 ms, 65536 fixed 1.38 ms (3 MB copy-outs
    // alternate with the 2 MB copy-ins instead of overlapping)
    constexpr int64_t IPPE_CHUNK = ASB_IPPE_CHUNK, IPPE_CHUNK_MAX = ASB_IPPE_CHUNK_MAX;
    constexpr int IPPE_EVENTS = 64;
    static std::mutex mu;
    static Cache cache[64];
    std::lock_guard<std::mutex> lock(mu);
    Cache& c = cache[device & 63];
    if (n > c.cap) {
        cudaFree(c.dc); cudaFree(c.dout); cudaFree(c.dout2); cudaFree(c.derr);
        c.dc = nullptr; c.dout = c.dout2 = nullptr; c.derr = nullptr; c.cap = 0;
        const int64_t cap = std::max<int64_t>(n, 4096);
        ASB_CUDA(cudaMalloc(&c.dc, (size_t)cap * 8 * sizeof(float)));
        ASB_CUDA(cudaMalloc(&c.dout, (size_t)cap * 6 * sizeof(double)));
        ASB_CUDA(cudaMalloc(&c.dout2, (size_t)cap * 6 * sizeof(double)));
        ASB_CUDA(cudaMalloc(&c.derr, (size_t)cap * 2 * sizeof(float)));
        c.cap = cap;
    }
    if (n >= 4 * 65536) {
        // Large batches (BASELINE config C4, 1M sets = 32 MB in + 48 MB out) go in slices: the H2D copy of slice
        // i+1, the kernel of slice i and the D2H copy of slice i-1 overlap (separate copy engines per direction).
        // One cudaMemcpyAsync per slice and buffer; with pageable host memory the copies stage through the driver
        // and the overlap is partial.
        if (!c.st[0]) {
            for (auto& st : c.st) ASB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            for (auto& e : c.ev_in) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            for (auto& e : c.ev_k) ASB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        }
        // three streams: copy-in, kernels, copy-out.  With a copy and its kernel on ONE stream the H2D copy of
        // slice i+2 queues behind the D2H copy of slice i and the two directions alternate instead of overlapping.
        cudaStream_t s_in = c.st[0], s_k = c.st[1], s_out = c.st[2];
        int i = 0;
        int64_t cn = 0, chunk = IPPE_CHUNK;
        for (int64_t o = 0; o < n; o += cn, ++i) {
            cn = std::min<int64_t>(chunk, n - o);
            if (i > 0) chunk = std::min<int64_t>(2 * chunk, IPPE_CHUNK_MAX);
            const int e = i % IPPE_EVENTS;
            ASB_CUDA(cudaMemcpyAsync(c.dc + 8 * o, corners + 8 * o, (size_t)cn * 8 * sizeof(float), cudaMemcpyHostToDevice, s_in));
            ASB_CUDA(cudaEventRecord(c.ev_in[e], s_in));
            ASB_CUDA(cudaStreamWaitEvent(s_k, c.ev_in[e], 0));
            int rc = asb_ippe_square_device(c.dc + 8 * o, cn, camera_matrix, dist_coeffs, marker_size, c.dout + 6 * o,
                                            out2 ? c.dout2 + 6 * o : nullptr, errs ? c.derr + 2 * o : nullptr, s_k);
            if (rc != ASB_OK) return rc;
            ASB_CUDA(cudaEventRecord(c.ev_k[e], s_k));
            ASB_CUDA(cudaStreamWaitEvent(s_out, c.ev_k[e], 0));
            ASB_CUDA(cudaMemcpyAsync(out + 6 * o, c.dout + 6 * o, (size_t)cn * 6 * sizeof(double), cudaMemcpyDeviceToHost, s_out));
            if (out2) ASB_CUDA(cudaMemcpyAsync(out2 + 6 * o, c.dout2 + 6 * o, (size_t)cn * 6 * sizeof(double), cudaMemcpyDeviceToHost, s_out));
            if (errs) ASB_CUDA(cudaMemcpyAsync(errs + 2 * o, c.derr + 2 * o, (size_t)cn * 2 * sizeof(float), cudaMemcpyDeviceToHost, s_out));
        }
        ASB_CUDA(cudaStreamSynchronize(s_out));   // every copy-out waits for its kernel, every kernel for its copy-in
        ASB_CUDA(cudaStreamSynchronize(s_in));
        return ASB_OK;
    }
    ASB_CUDA(cudaMemcpy(c.dc, corners, (size_t)n * 8 * sizeof(float), cudaMemcpyHostToDevice));
    int rc = asb_ippe_square_device(c.dc, n, camera_matrix, dist_coeffs, marker_size, c.dout, out2 ? c.dout2 : nullptr,
                                    errs ? c.derr : nullptr, nullptr);
    if (rc != ASB_OK) return rc;
    ASB_CUDA(cudaMemcpy(out, c.dout, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost));
    if (out2) ASB_CUDA(cudaMemcpy(out2, c.dout2, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost));
    if (errs) ASB_CUDA(cudaMemcpy(errs, c.derr, (size_t)n * 2 * sizeof(float), cudaMemcpyDeviceToHost));
    return ASB_OK;
}

int asb_ippe_square(const float* corners, int64_t n, const double camera_matrix[9],
                    const double dist_coeffs[5], double marker_size, double* out, int device) {
    return ippe_host(corners, n, camera_matrix, dist_coeffs, marker_size, out, nullptr, nullptr, device);
}

// cv2.solvePnPGeneric(..., SOLVEPNP_IPPE_SQUARE): BOTH solutions, the better one first (SURVEY.md 8f row 3:
// what an orientation-ambiguity gate such as BaseFilter.correct_poses, BASE:274-325, chooses between).
int asb_ippe_square_both(const float* corners, int64_t n, const double camera_matrix[9], const double dist_coeffs[5],
                         double marker_size, double* out, double* out2, float* errs, int device) {
    if (n > 0 && !out2) return fail(ASB_ERR_BAD_ARG, "out2 is NULL");
    return ippe_host(corners, n, camera_matrix, dist_coeffs, marker_size, out, out2, errs, device);
}


// ---- allocation-free variant (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops) ----
// The device staging that asb_ippe_square keeps in a grow-only cudaMalloc cache lives here in a
// file-scope array instead: ASB_IPPE_STATIC_MAX corner sets x (32 B of corners + 48 B of poses) =
// 1.25 MiB, held by every GPU that loads the library until the process exits.  g_ippe_static is a
// host-side stand-in in host code, NOT a device address: the address the kernel gets comes from
// cudaGetSymbolAddress, taken once per device with that device current.
// Every call shares the array and it keeps the last call's contents.  Nothing has to start at zero:
// the copy overwrites corner rows [0,n) and the kernel writes every pose row [0,n) it returns; rows
// >= n are stale and never read.  Calls that overlap on the GPU (two host threads on different
// streams) would race on it: the entry point serialises callers with a mutex and runs on the legacy
// default stream, so calls through it do not overlap.
static __device__ __align__(256) unsigned char g_ippe_static[(size_t)ASB_IPPE_STATIC_MAX * 80];

static int ippe_static_scratch(int device, unsigned char** out) {
    static unsigned char* ptr[64] = {};
    if (device < 0 || device >= 64) return fail(ASB_ERR_BAD_ARG, "bad device %d", device);
    ASB_CUDA(cudaSetDevice(device));
    if (!ptr[device]) {
        void* p = nullptr;
        ASB_CUDA(cudaGetSymbolAddress(&p, g_ippe_static));
        if (!p) return fail(ASB_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = ptr[device];
    return ASB_OK;
}

int asb_ippe_square_static(const float* corners, int64_t n, const double camera_matrix[9],
                           const double dist_coeffs[5], double marker_size, double* out, int device) {
    if (n < 0 || !camera_matrix || (n > 0 && (!corners || !out))) return fail(ASB_ERR_BAD_ARG, "bad ippe arguments");
    if (n > ASB_IPPE_STATIC_MAX)
        return fail(ASB_ERR_CAPACITY, "asb_ippe_square_static: %lld corner sets exceed the static scratch of %d "
                    "(ASB_IPPE_STATIC_MAX); use asb_ippe_square", (long long)n, ASB_IPPE_STATIC_MAX);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(ASB_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (n == 0) return ASB_OK;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    unsigned char* base = nullptr;
    int rc = ippe_static_scratch(device, &base);
    if (rc != ASB_OK) return rc;
    float* dc = reinterpret_cast<float*>(base);
    double* dout = reinterpret_cast<double*>(base + (size_t)ASB_IPPE_STATIC_MAX * 32);
    ASB_CUDA(cudaMemcpy(dc, corners, (size_t)n * 8 * sizeof(float), cudaMemcpyHostToDevice));
    rc = asb_ippe_square_device(dc, n, camera_matrix, dist_coeffs, marker_size, dout, nullptr, nullptr, nullptr);
    if (rc != ASB_OK) return rc;
    ASB_CUDA(cudaMemcpy(out, dout, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost));
    return ASB_OK;
}

}  // extern "C"
