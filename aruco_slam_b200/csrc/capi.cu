// Library-level entry points of the C ABI and the measurement helpers
// (asb_fp64_peak, asb_copy_bandwidth) used by bench.py for the roofline denominators that
// MEASURED_PEAKS.json does not carry (FP64 pipe peak).
#include <mma.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace asb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
const char* get_error() { return g_err; }

// ---- FP64 FMA pipe: 16 independent chains per thread ------------------------------------
__global__ void __launch_bounds__(256) k_dfma_peak(double* out, int iters, double seed) {
    double a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed + i + threadIdx.x;
    const double x = 1.0000001, y = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fma(a[i], x, y);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- FP64 tensor path: mma.sync.m8n8k4.f64 (DMMA), 8 independent accumulator tiles -----
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) k_dmma_peak(double* out, int iters, double seed) {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = seed + i;
    const double a = 1e-3 + threadIdx.x * 1e-6, b = 1e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma884(c[i][0], c[i][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// both pipes in one kernel: even warps DFMA, odd warps DMMA
__global__ void __launch_bounds__(256) k_mixed_peak(double* out, int iters, double seed) {
    const int warp = threadIdx.x >> 5;
    double s = 0;
    if (warp & 1) {
        double c[8][2];
#pragma unroll
        for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = seed + i;
        const double a = 1e-3 + threadIdx.x * 1e-6, b = 1e-3;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma884(c[i][0], c[i][1], a, b);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    } else {
        double a[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = seed + i + threadIdx.x;
        const double x = 1.0000001, y = 1e-9;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fma(a[i], x, y);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) s += a[i];
    }
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// dependent-issue latencies (clocks per op) of one warp: DFMA, rsqrt(double), LDS.64, BAR.SYNC (256 thr)
__global__ void k_latency(double* out, double seed) {
    __shared__ double sm[64];
    const int n = 512;
    double a = seed, r[4];
    long long t0, t1;
    if (threadIdx.x < 64) sm[threadIdx.x] = (double)((threadIdx.x + 1) & 63);
    __syncthreads();
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) a = fma(a, 1.0000001, 1e-9);
    t1 = clock64();
    r[0] = (double)(t1 - t0) / n;
    double b = 1.5 + a * 1e-30;
    t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < 64; ++i) b = rsqrt(b) + 1.0;
    t1 = clock64();
    r[1] = (double)(t1 - t0) / 64;
    int idx = (int)(b * 1e-30);
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) idx = (int)sm[idx];
    t1 = clock64();
    r[2] = (double)(t1 - t0) / n;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < 64; ++i) __syncthreads();
    t1 = clock64();
    r[3] = (double)(t1 - t0) / 64;
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        out[0] = r[0]; out[1] = r[1]; out[2] = r[2]; out[3] = r[3];
        out[4] = a + b + idx;
    }
}

__global__ void k_copy(const double2* __restrict__ src, double2* __restrict__ dst, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        __stcs(dst + i, __ldcs(src + i));
}

// write-only / read-only / 1 read : 2 writes streams, 256-bit accesses
__global__ void k_stream(const double* __restrict__ src, double* __restrict__ d1, double* __restrict__ d2, long long n4,
                         int mode, double* sink) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    double acc = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        double a = 1.0, b = 2.0, c = 3.0, d = 4.0;
        if (mode != 0) asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(src + 4 * i));
        if (mode == 1) acc += a + b + c + d;
        if (mode == 0 || mode == 2)
            asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(d1 + 4 * i), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
        if (mode == 2)
            asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(d2 + 4 * i), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
    }
    if (acc == 123.456) *sink = acc;
}

// The sweep's memory pattern without any math: lower-triangle TS x TS tiles of an n x n
// matrix (leading dimension ld): read tile (I,J), write it back, write it to (J,I) (not
// transposed inside: same bytes, same addresses pattern as the mirror).  One warp per row
// slice: lanes cover 32 x 32 B = 1 KB or less of a row.
template <int TS>
__global__ void k_tilepattern(double* P, int ld, int ntile_side) {
    const int ntiles = ntile_side * (ntile_side + 1) / 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const int t_begin = (int)((long long)ntiles * blockIdx.x / gridDim.x);
    const int t_end = (int)((long long)ntiles * (blockIdx.x + 1) / gridDim.x);
    for (int t = t_begin; t < t_end; ++t) {
        int I = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
        while ((I + 1) * (I + 2) / 2 <= t) ++I;
        while (I * (I + 1) / 2 > t) --I;
        const int J = t - I * (I + 1) / 2;
        for (int e = warp * 32 + lane; e < TS * (TS / 4); e += nw * 32) {  // 32-byte pieces, row-major
            const int r = e / (TS / 4), c = 4 * (e % (TS / 4));
            double a, b, cc, d;
            asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(cc), "=d"(d)
                         : "l"(P + (long long)(I * TS + r) * ld + J * TS + c));
            asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(P + (long long)(I * TS + r) * ld + J * TS + c),
                         "d"(a), "d"(b), "d"(cc), "d"(d) : "memory");
            if (I != J)
                asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(P + (long long)(J * TS + r) * ld + I * TS + c),
                             "d"(a), "d"(b), "d"(cc), "d"(d) : "memory");
        }
    }
}

}  // namespace asb

using namespace asb;

extern "C" {

int asb_abi_version(void) { return ASB_ABI_VERSION; }

const char* asb_last_error(void) { return get_error(); }

int asb_device_count(int* count) {
    if (!count) return fail(ASB_ERR_BAD_ARG, "count is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        return fail(ASB_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    }
    *count = n;
    return ASB_OK;
}

int asb_default_noise(int kind, asb_noise* out) {
    if (!out || (kind != ASB_KIND_EKF && kind != ASB_KIND_ROT)) return fail(ASB_ERR_BAD_ARG, "bad kind/out");
    Noise n = default_noise(kind);
    out->init_cam = n.init_cam; out->init_lm = n.init_lm; out->r = n.r;
    out->q_cam = n.q_cam; out->q_err = n.q_err; out->q_lm = n.q_lm;
    return ASB_OK;
}

int asb_fp64_peak(int device, int which, double* tflops) {
    if (!tflops || which < 0 || which > 2) return fail(ASB_ERR_BAD_ARG, "bad which/tflops");
    ASB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    const int grid = prop.multiProcessorCount * 8, block = 256, iters = 20000;
    double* out = nullptr;
    ASB_CUDA(cudaMalloc(&out, (size_t)grid * block * sizeof(double)));
    cudaEvent_t e0, e1;
    ASB_CUDA(cudaEventCreate(&e0));
    ASB_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        ASB_CUDA(cudaEventRecord(e0));
        if (which == 0) k_dfma_peak<<<grid, block>>>(out, iters, 1.0);
        else if (which == 1) k_dmma_peak<<<grid, block>>>(out, iters, 1.0);
        else k_mixed_peak<<<grid, block>>>(out, iters, 1.0);
        ASB_CUDA(cudaEventRecord(e1));
        ASB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ASB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double threads = (double)grid * block, warps = threads / 32;
        double flop;
        if (which == 0) flop = threads * iters * 16.0 * 2.0;
        else if (which == 1) flop = warps * iters * 8.0 * (8 * 8 * 4 * 2.0);
        else flop = 0.5 * threads * iters * 16.0 * 2.0 + 0.5 * warps * iters * 8.0 * (8 * 8 * 4 * 2.0);
        best = std::max(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    *tflops = best;
    return ASB_OK;
}

int asb_latency_probe(int device, double clocks[4]) {
    if (!clocks) return fail(ASB_ERR_BAD_ARG, "clocks is NULL");
    ASB_CUDA(cudaSetDevice(device));
    double* out = nullptr;
    ASB_CUDA(cudaMalloc(&out, 8 * sizeof(double)));
    k_latency<<<1, 256>>>(out, 1.0);
    k_latency<<<1, 256>>>(out, 1.0);
    double h[8];
    ASB_CUDA(cudaMemcpy(h, out, sizeof h, cudaMemcpyDeviceToHost));
    cudaFree(out);
    for (int i = 0; i < 4; ++i) clocks[i] = h[i];
    return ASB_OK;
}

int asb_stream_bandwidth(int device, int64_t bytes, int mode, double* gbs) {
    if (!gbs || bytes < (1 << 20) || mode < 0 || mode > 2) return fail(ASB_ERR_BAD_ARG, "bad args");
    ASB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    double *a = nullptr, *b = nullptr, *c = nullptr;
    const long long n4 = bytes / 32;
    ASB_CUDA(cudaMalloc(&a, n4 * 32));
    ASB_CUDA(cudaMalloc(&b, n4 * 32));
    ASB_CUDA(cudaMalloc(&c, n4 * 32));
    ASB_CUDA(cudaMemset(a, 0, n4 * 32));
    cudaEvent_t e0, e1;
    ASB_CUDA(cudaEventCreate(&e0));
    ASB_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        ASB_CUDA(cudaEventRecord(e0));
        k_stream<<<prop.multiProcessorCount * 8, 512>>>(a, b, c, n4, mode, b);
        ASB_CUDA(cudaEventRecord(e1));
        ASB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ASB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double moved = (mode == 2 ? 3.0 : 1.0) * n4 * 32.0;
        best = std::max(best, moved / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(a); cudaFree(b); cudaFree(c);
    *gbs = best;
    return ASB_OK;
}

int asb_tilepattern_bandwidth(int device, int n, int tile, int ctas_per_sm, double* us) {
    if (!us || n < 256 || (tile != 64 && tile != 128 && tile != 256)) return fail(ASB_ERR_BAD_ARG, "bad args");
    ASB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    const int side = (n + tile - 1) / tile, ld = (side * tile + 127) / 128 * 128;
    double* P = nullptr;
    ASB_CUDA(cudaMalloc(&P, (size_t)ld * ld * sizeof(double)));
    ASB_CUDA(cudaMemset(P, 0, (size_t)ld * ld * sizeof(double)));
    cudaEvent_t e0, e1;
    ASB_CUDA(cudaEventCreate(&e0));
    ASB_CUDA(cudaEventCreate(&e1));
    double best = 1e30;
    for (int rep = 0; rep < 6; ++rep) {
        ASB_CUDA(cudaEventRecord(e0));
        const int grid = prop.multiProcessorCount * ctas_per_sm;
        if (tile == 64) k_tilepattern<64><<<grid, 512>>>(P, ld, side);
        else if (tile == 128) k_tilepattern<128><<<grid, 512>>>(P, ld, side);
        else k_tilepattern<256><<<grid, 512>>>(P, ld, side);
        ASB_CUDA(cudaEventRecord(e1));
        ASB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ASB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        best = std::min(best, 1e3 * (double)ms);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(P);
    *us = best;
    return ASB_OK;
}

int asb_copy_bandwidth(int device, int64_t bytes, double* gbs) {
    if (!gbs || bytes < (1 << 20)) return fail(ASB_ERR_BAD_ARG, "bad bytes/gbs");
    ASB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ASB_CUDA(cudaGetDeviceProperties(&prop, device));
    double2 *a = nullptr, *b = nullptr;
    const long long n = bytes / sizeof(double2);
    ASB_CUDA(cudaMalloc(&a, n * sizeof(double2)));
    ASB_CUDA(cudaMalloc(&b, n * sizeof(double2)));
    ASB_CUDA(cudaMemset(a, 0, n * sizeof(double2)));
    cudaEvent_t e0, e1;
    ASB_CUDA(cudaEventCreate(&e0));
    ASB_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        ASB_CUDA(cudaEventRecord(e0));
        k_copy<<<prop.multiProcessorCount * 16, 512>>>(a, b, n);
        ASB_CUDA(cudaEventRecord(e1));
        ASB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ASB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        best = std::max(best, 2.0 * n * sizeof(double2) / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(a);
    cudaFree(b);
    *gbs = best;
    return ASB_OK;
}

}  // extern "C"
