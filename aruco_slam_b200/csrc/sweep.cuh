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
    int reverse;           // walk each CTA's tile chunk backwards (odd frames): the tiles a CTA wrote last
                           // in the previous sweep are the ones it reads first, while they are still in L2
    unsigned long long* prof;   // CTA 0 stores globaltimer at start / end into prof[0..1] (may be null)
    unsigned long long* cta_prof;   // every CTA stores its start / end / %smid into cta_prof[3 b .. 3 b + 2] (may be null)
    int pol;               // cache policies (ld256 / st256 modes): bits 0-2 P loads, 3-5 direct stores,
                           // 6-8 mirror stores; 0 = take the dbg bits
    // Tile schedule.  tail == null: CTA b walks the contiguous range [ntiles b / grid, ntiles (b+1) / grid).
    // tail != null: CTA b walks n_static contiguous tiles [b n_static, (b+1) n_static) and then takes
    // single tiles >= grid * n_static from the shared counter *tail until they run out, so that an SM
    // that is ahead (or a CTA whose range was a tile shorter) absorbs the remainder instead of
    // idling: with contiguous ranges alone the slowest CTA ends 8 % after the median one.
    // Row-sharded engine (filter_shard.cu): this GPU owns the full rows [row0, row0 + 64 nI) of P, stored as
    // a (64 nI) x ld block.  rect != 0: the tile list is the nI x nJ rectangle of 64 x 64 tiles (t = I nJ + J),
    // every tile is stored directly (no mirror), VI comes from the LOCAL panel block I, the diagonal
    // (Q, predict) is where row0 + local row == column.  SweepDual only.
    int rect, row0, nJ;
    // lower_only: P_out keeps only its lower triangle (i >= j) up to date: no mirror stores, i.e. 8 N^2
    // instead of 12 N^2 bytes of DRAM traffic per sweep.  The readers of the pipelined engine
    // (k_f1, k_f2, asb_get_cov) take P[i][j], i < j, from P[j][i].
    int lower_only;
    // reserve_sm >= 0 (per-frame observe(), filter_single.cu): CTAs that land on that SM retire at once, so the SM
    // stays free for k_f1 of the NEXT frame, which then has an FP64 pipe to itself instead of sharing one with a
    // sweep CTA.  The static ranges are then not tied to blockIdx: a live CTA claims range after range from
    // *range_ctr (n_ranges of them) until none is left, then joins the dynamic tail - correct for any number of
    // live CTAs.
    int reserve_sm, n_ranges;
    unsigned int* range_ctr;   // zero on entry
    unsigned int* range_reset; // the NEXT launch's range counter: zeroed by CTA 0
    unsigned int* tail;        // counter of this launch (zero on entry)
    unsigned int* tail_reset;  // counter of the NEXT launch: zeroed by CTA 0
    int n_static;
};

// State update + multiplicative quaternion reset (EKF:134-152, ROT:143-177) by one CTA.
__device__ __forceinline__ void apply_state(double* state, const double* dx, int apply_kind, int as_written, int nlm,
                                            int tid, int nthreads) {
    if (tid == 0) {
        double cam[CAM], d[CAM];
#pragma unroll
        for (int c = 0; c < CAM; ++c) { cam[c] = state[c]; d[c] = __ldcg(dx + c); }
        camera_reset(cam, d, as_written);
#pragma unroll
        for (int c = 0; c < CAM; ++c) state[c] = cam[c];
    }
    if (apply_kind == 0) {
        for (int e = tid; e < 3 * nlm; e += nthreads) state[CAM + e] += __ldcg(dx + CAM + e);
    } else {
        for (int t = tid; t < nlm; t += nthreads) {
            double lm[10], d[10];
#pragma unroll
            for (int c = 0; c < 10; ++c) { lm[c] = state[CAM + 10 * t + c]; d[c] = __ldcg(dx + CAM + 10 * t + c); }
            landmark_update<1>(lm, d);
#pragma unroll
            for (int c = 0; c < 10; ++c) state[CAM + 10 * t + c] = lm[c];
        }
    }
}

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
// Programmatic dependent launch (kernels launched with cudaLaunchAttributeProgrammaticStreamSerialization):
// pdl_trigger lets the next kernel of the stream be scheduled once every CTA of this one has
// called it; pdl_wait blocks until the previous kernel of the stream has completed and its
// writes are visible.  Both are no-ops for a kernel launched without the attribute.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

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
// (address = (4 kk + t) * V_LDB + g + 8 r) hit 32 distinct banks.
//
// Fragment ownership is chosen for the GLOBAL side: lane (g = lane / 4, t4 = lane % 4) of a
// warp whose tile starts at (r0, c0) owns
//     rows     r0 + 4 g + rm                                     (DMMA tile rm = 0..3)
//     columns  c0 + 16 (rn / 2) + 4 t4 + 2 (rn % 2) + e          (DMMA tile rn, e = 0, 1)
// i.e. 4 consecutive rows and groups of 4 consecutive columns, so that every access to P is
// a 256-bit LDG/STG (SASS LDG.E.ENL2.256): a warp instruction covers whole 128-byte lines
// (8 rows x 128 B direct, 4 rows x 256 B mirrored) instead of 64-byte pieces.  The price is
// that the V panel is kept in two column orders, one for each role, so that the fragment
// loads stay g + 8 r:  position of state column c inside its block
//     as V_I (rows of the tile):     32 (c/32) + (c%32)/4 + 8 (c%4)
//     as V_J (columns of the tile):  16 (c/16) + 8 ((c%4)/2) + 2 ((c%16)/4) + c%2
// k_front writes both.
constexpr int V_BLK = 64, V_LDB = 68;
__host__ __device__ inline int v_pos_row(int c) { return (c & ~31) + ((c & 31) >> 2) + 8 * (c & 3); }
__host__ __device__ inline int v_pos_col(int c) { return (c & ~15) + 8 * ((c & 3) >> 1) + 2 * ((c & 15) >> 2) + (c & 1); }

// cache policies: 0 default, 1 evict-first (.cs), 2 L2-only (.cg), 3 L2 evict-last (SASS .ELL2), 4 L2 evict-first (.EFL2)
__device__ __forceinline__ void ld256(const double* p, double& a, double& b, double& c, double& d, int mode = 0) {
    if (mode == 1) asm volatile("ld.global.cs.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
    else if (mode == 2) asm volatile("ld.global.cg.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
    else if (mode == 3) asm volatile("ld.global.L2::evict_last.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
    else if (mode == 4) asm volatile("ld.global.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
    else asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}
__device__ __forceinline__ void st256(double* p, double a, double b, double c, double d, int mode = 0) {
    if (mode == 1) asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
    else if (mode == 2) asm volatile("st.global.cg.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
    else if (mode == 3) asm volatile("st.global.L2::evict_last.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
    else if (mode == 4) asm volatile("st.global.L2::evict_first.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
    else asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
// P sub-tile of one warp <-> registers, fragment ownership as above.  pr = P + (r0 + 4 g) * ld
// + c0 + 4 t4 (direct), pm = P + (c0 + 4 t4) * ld + r0 + 4 g (mirror).
template <int RM, int RN>
__device__ __forceinline__ void tile_load(const double* pr, int ld, double (&v)[RM][RN][2], int mode = 0) {
#pragma unroll
    for (int rm = 0; rm < RM; ++rm)
#pragma unroll
        for (int q = 0; q < RN / 2; ++q)
            ld256(pr + (long long)rm * ld + 16 * q, v[rm][2 * q][0], v[rm][2 * q][1], v[rm][2 * q + 1][0], v[rm][2 * q + 1][1], mode);
}
template <int RM, int RN>
__device__ __forceinline__ void tile_store(double* pr, int ld, const double (&v)[RM][RN][2], int mode = 0) {
#pragma unroll
    for (int rm = 0; rm < RM; ++rm)
#pragma unroll
        for (int q = 0; q < RN / 2; ++q)
            st256(pr + (long long)rm * ld + 16 * q, v[rm][2 * q][0], v[rm][2 * q][1], v[rm][2 * q + 1][0], v[rm][2 * q + 1][1], mode);
}
template <int RM, int RN>
__device__ __forceinline__ void tile_store_mirror(double* pm, int ld, const double (&v)[RM][RN][2], int mode = 0) {
    static_assert(RM == 4, "a lane owns 4 consecutive rows");
#pragma unroll
    for (int rn = 0; rn < RN; ++rn)
#pragma unroll
        for (int e = 0; e < 2; ++e)
            st256(pm + (long long)(16 * (rn >> 1) + 2 * (rn & 1) + e) * ld, v[0][rn][e], v[1][rn][e], v[2][rn][e], v[3][rn][e], mode);
}

// Tile shape: WM x WN warps, each owning (8 RM) x (8 RN) elements as RM x RN DMMA tiles.
// LATE: the P sub-tile is loaded AFTER the k-loop instead of being prefetched before it (32 registers
// fewer during the MMAs); NBUF: VJ buffers (1: the next panel is fetched after this tile's k-loop).
// Both trade per-CTA latency hiding for occupancy (a third CTA per SM keeps the FP64 tensor pipe fed
// while the other two are between k-loops).
template <int WM_, int WN_, int RM_, int RN_, int MINB_, bool LATE_ = false, int NBUF_ = 2>
struct SweepCfg {
    static constexpr int WM = WM_, WN = WN_, RM = RM_, RN = RN_, MINB = MINB_, NBUF = NBUF_;
    static constexpr bool LATE = LATE_;
    static constexpr int TI = WM * RM * 8, TJ = WN * RN * 8, THREADS = WM * WN * 32;
    static constexpr int RATIO = TI / TJ;             // column tiles per row tile on the diagonal
    static_assert(TJ == V_BLK && TI % V_BLK == 0, "tiles are made of whole V blocks");
    __host__ __device__ static size_t smem_bytes(int m4) {
        return (size_t)m4 * V_LDB * (RATIO + NBUF) * sizeof(double);
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
// latency hides behind the tile's own math.  Pin is read (lower triangle), P is written (both
// triangles): the same buffer for an in-place update, or the two halves of the pipelined
// engine's ping-pong pair (filter_single.cu), where the next frame's k_f1 still reads Pin.
// RESERVE: the per-frame instantiation (reserve_sm mode, see SweepExtra); the stream-mode / sharded instantiation
// compiles without it so that its tuned register allocation and control flow stay as they were.
template <class C, bool RESERVE = false>
__global__ void __launch_bounds__(C::THREADS, C::MINB)
k_sweep(const double* Pin, double* P, int ld, const double* __restrict__ Vr, const double* __restrict__ Vc, int m4,
        int ntiles, int dbg, SweepExtra ex) {
    // dbg (asb_bench_sweep only; 0 in the product path): 1 skip MMAs, 2 skip mirror stores,
    // 4 skip direct stores, 8 skip P loads; bits 4-5 load cache policy, bits 6-7 store cache
    // policy (0 default, 1 .cs evict-first, 2 .cg)
    constexpr int RM = C::RM, RN = C::RN, TI = C::TI, TJ = C::TJ;
    extern __shared__ __align__(128) unsigned char smraw[];
    const int blk = m4 * V_LDB;                      // doubles per 64-column block of V
    double* VI = reinterpret_cast<double*>(smraw);   // RATIO blocks
    double* VJ0 = VI + (size_t)C::RATIO * blk;       // 2 x 1 block
    __shared__ __align__(8) uint64_t bar[3];         // [0] VI, [1] VJ buf 0, [2] VJ buf 1
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    const int wi = warp / C::WN, wj = warp % C::WN;
    if (ex.prof && blockIdx.x == 0 && tid == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        ex.prof[0] = t;
    }
    if (ex.cta_prof && tid == 0) {
        unsigned long long t;
        unsigned int sm;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        ex.cta_prof[3 * blockIdx.x] = t;
        ex.cta_prof[3 * blockIdx.x + 2] = sm;
    }
    // Stream-mode ordering (filter_single.cu): k_f1(t+1) -> k_sweep(t) -> k_f2(t+1), each launched
    // as a programmatic dependent of the one before.  The sweep does not need k_f1's results, so it
    // only waits for it at the very end (its own completion then implies k_f1's).  The trigger for
    // k_f2 is at the end as well: released earlier, its CTAs pile up on the first SMs to drain
    // instead of spreading over the GPU (measured: k_f2 11 -> 15 us).
    if (blockIdx.x == 0 && ex.apply_kind >= 0)   // state update (k_apply folded in)
        apply_state(ex.state, ex.dx, ex.apply_kind, ex.as_written, ex.nlm, tid, C::THREADS);
    const bool dyn = ex.tail != nullptr;
    const bool reserve = RESERVE && dyn && ex.range_ctr != nullptr && ex.reserve_sm >= 0;   // a zero-initialised SweepExtra is 'off'
    if (dyn && blockIdx.x == 0 && tid == 0) {
        *ex.tail_reset = 0u;
        if (reserve) *ex.range_reset = 0u;
    }
    if (reserve) {
        unsigned int sm;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        if ((int)sm == ex.reserve_sm) { pdl_trigger(); pdl_wait(); return; }   // the whole CTA: %smid is CTA-uniform
    }
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        mbar_init(&bar[2], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __shared__ int s_next[2], s_nij[2], s_range;
    if (reserve && tid == 0) s_range = (int)atomicAdd(ex.range_ctr, 1u);
    __syncthreads();
    // virtual block index: blockIdx, or the claimed range in reserve mode (>= n_ranges: no static range)
    int vb = reserve ? s_range : (int)blockIdx.x;
    const int nvb = reserve ? ex.n_ranges : (int)gridDim.x;
    int t_begin = dyn ? (vb < nvb ? vb * ex.n_static : 0) : (int)((long long)ntiles * blockIdx.x / gridDim.x);
    int t_end = dyn ? (vb < nvb ? t_begin + ex.n_static : 0) : (int)((long long)ntiles * (blockIdx.x + 1) / gridDim.x);
    const int dyn_base = nvb * ex.n_static;
    const bool rev = ex.reverse != 0;
    if (t_begin >= t_end && !dyn) { pdl_trigger(); pdl_wait(); return; }
    const int pol_ld = ex.pol ? (ex.pol & 7) : ((dbg >> 4) & 3);
    const int pol_st = ex.pol ? ((ex.pol >> 3) & 7) : ((dbg >> 6) & 3);
    const int pol_mir = ex.pol ? ((ex.pol >> 6) & 7) : ((dbg >> 6) & 3);

    const uint32_t blk_bytes = (uint32_t)(blk * sizeof(double));
    auto load_vi = [&](int i0) {  // one thread
        mbar_expect_tx(&bar[0], C::RATIO * blk_bytes);
        bulk_g2s(VI, Vr + (size_t)(i0 / V_BLK) * blk, C::RATIO * blk_bytes, &bar[0]);
    };
    auto load_vj = [&](int buf, int j0) {  // one thread
        mbar_expect_tx(&bar[1 + buf], blk_bytes);
        bulk_g2s(VJ0 + (size_t)buf * blk, Vc + (size_t)(j0 / V_BLK) * blk, blk_bytes, &bar[1 + buf]);
    };
    // One thread draws the next tile of the dynamic tail: s_next[slot] = tile or -1, s_nij[slot] = (I << 16) | J,
    // and starts the bulk copy of its VJ panel into buffer vbuf.  t = RATIO I (I + 1) / 2 + J.
    auto draw = [&](int slot, int vbuf) {
        int nx = dyn_base + (int)atomicAdd(ex.tail, 1u);
        if (nx >= ntiles) nx = -1;
        s_next[slot] = nx;
        if (nx >= 0) {
            int i, j;
            if (ex.rect) {
                i = nx / ex.nJ;
                j = nx - i * ex.nJ;
            } else {
                i = (int)((sqrtf(8.0f * (float)nx / (float)C::RATIO + 1.0f) - 1.0f) * 0.5f);   // estimate, fixed up
                while (C::RATIO * (i + 1) * (i + 2) / 2 <= nx) ++i;
                while (C::RATIO * i * (i + 1) / 2 > nx) --i;
                j = nx - C::RATIO * i * (i + 1) / 2;
            }
            s_nij[slot] = (i << 16) | j;
            if (C::NBUF == 2) load_vj(vbuf, j * TJ);   // NBUF == 1: the single buffer is still in use, see do_tile
        }
    };

    uint32_t ph_i = 0, ph_j = 0;  // ph_j: bit b = phase of VJ buffer b
    bool vi_pending = true;

    // One tile: P sub-tile -> registers, out = P - V_I^T V_J, direct + mirror stores.  draw_slot >= 0: this
    // is where the next dynamic tile is drawn (after the P loads are in flight: the counter's round
    // trip hides behind them).  Inlined into both loops below.
    auto do_tile = [&](const int I, const int J, const int buf_in, const int draw_slot, const int next_j0) {
        const int buf = C::NBUF == 2 ? buf_in : 0;
        const int i0 = I * TI, j0 = J * TJ;
            // P sub-tile -> registers (fragment ownership of sweep.cuh's header: 256-bit accesses)
            const int r0 = i0 + 8 * RM * wi + 4 * g, c0 = j0 + 8 * RN * wj + 4 * t4;
            double pf[RM][RN][2];
            if (C::LATE) {
            } else if (dbg & 8) {
#pragma unroll
                for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                    for (int rn = 0; rn < RN; ++rn) pf[rm][rn][0] = pf[rm][rn][1] = 0.0;
            } else {
                tile_load<RM, RN>(Pin + (long long)r0 * ld + c0, ld, pf, pol_ld);
            }
        if (draw_slot >= 0 && tid == 0) draw(draw_slot, buf ^ 1);

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

            if (C::NBUF == 1) {
                __syncthreads();   // every warp is done with the VJ buffer: the next tile's panel flies during the epilogue
                if (tid == 0) {
                    int nj = next_j0;
                    if (draw_slot >= 0) nj = s_next[draw_slot] >= 0 ? (s_nij[draw_slot] & 0xffff) * TJ : -1;
                    if (nj >= 0) load_vj(0, nj);
                }
            }
            if (C::LATE) {
                if (dbg & 8) {
#pragma unroll
                    for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                        for (int rn = 0; rn < RN; ++rn) pf[rm][rn][0] = pf[rm][rn][1] = 0.0;
                } else {
                    tile_load<RM, RN>(Pin + (long long)r0 * ld + c0, ld, pf, pol_ld);
                }
            }
            const bool interior = i0 >= j0 + TJ;  // strictly below the diagonal
#pragma unroll
            for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                for (int rn = 0; rn < RN; ++rn) {
                    pf[rm][rn][0] -= acc[rm][rn][0];
                    pf[rm][rn][1] -= acc[rm][rn][1];
                }
            if (ex.rect) {   // row-sharded engine: local row block, every tile stored directly
                if (ex.add_q && ex.row0 + i0 < j0 + TJ && ex.row0 + i0 + TI > j0) {   // tile crosses the diagonal
#pragma unroll
                    for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                        for (int rn = 0; rn < RN; ++rn)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int gi = ex.row0 + r0 + rm, j = c0 + 16 * (rn >> 1) + 2 * (rn & 1) + e;
                                if (gi == j && gi < ex.n) pf[rm][rn][e] += process_noise(ex.nz, gi);
                            }
                }
                tile_store<RM, RN>(P + (long long)r0 * ld + c0, ld, pf, pol_st);
            } else if (interior) {
                if (!(dbg & 4)) tile_store<RM, RN>(P + (long long)r0 * ld + c0, ld, pf, pol_st);
                if (!(dbg & 2) && !ex.lower_only) tile_store_mirror<RM, RN>(P + (long long)c0 * ld + r0, ld, pf, pol_mir);
            } else {
#pragma unroll
                for (int rm = 0; rm < RM; ++rm)
#pragma unroll
                    for (int rn = 0; rn < RN; ++rn)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int i = r0 + rm, j = c0 + 16 * (rn >> 1) + 2 * (rn & 1) + e;
                            if (i >= j) {
                                double o = pf[rm][rn][e];
                                if (i == j && ex.add_q && i < ex.n) o += process_noise(ex.nz, i);
                                P[(long long)i * ld + j] = o;
                                if (i > j && !ex.lower_only) P[(long long)j * ld + i] = o;
                            }
                        }
            }
    };

    int buf = 0;
    int Ivi = -1;   // row block whose VI panel is (being) loaded
    bool drew = false;   // the first tail tile has been drawn (by the last tile of a static range)
    while (t_begin < t_end) {
        // ---- contiguous range: tile indices follow arithmetically, everything stays warp-uniform ----
        const int t_first = rev ? t_end - 1 : t_begin;
        int I = (int)((sqrt(8.0 * t_first / C::RATIO + 1.0) - 1.0) * 0.5);
        while (C::RATIO * (I + 1) * (I + 2) / 2 <= t_first) ++I;
        while (C::RATIO * I * (I + 1) / 2 > t_first) --I;
        int J = t_first - C::RATIO * I * (I + 1) / 2;
        if (ex.rect) {
            I = t_first / ex.nJ;
            J = t_first - I * ex.nJ;
        }
        if (tid == 0) {
            load_vi(I * TI);
            load_vj(buf, J * TJ);
        }
        vi_pending = true;
        for (int t = t_begin; t < t_end; ++t, buf ^= 1) {
            int In = I, Jn = rev ? J - 1 : J + 1;
            if (rev) {
                if (Jn < 0) { In = I - 1; Jn = C::RATIO * I - 1; }
            } else if (Jn == (ex.rect ? ex.nJ : C::RATIO * (I + 1))) { In = I + 1; Jn = 0; }
            const bool has_next = t + 1 < t_end;
            if (C::NBUF == 2 && tid == 0 && has_next) load_vj(buf ^ 1, Jn * TJ);  // buffer freed by the last __syncthreads
            do_tile(I, J, buf, (!has_next && dyn && !reserve) ? 0 : -1, has_next ? Jn * TJ : -1);
            if (!has_next && dyn && !reserve) drew = true;
            __syncthreads();  // every warp is done with VJ[buf] (and VI) before they are overwritten
            Ivi = I;
            if (has_next && In != I) {
                if (tid == 0) load_vi(In * TI);
                vi_pending = true;
                Ivi = In;
            }
            I = In;
            J = Jn;
        }
        // reserve mode: another static range may be unclaimed (fewer live CTAs than ranges)
        t_begin = t_end = 0;
        if (reserve) {
            if (tid == 0) s_range = (int)atomicAdd(ex.range_ctr, 1u);
            __syncthreads();
            vb = s_range;
            if (vb < nvb) { t_begin = vb * ex.n_static; t_end = t_begin + ex.n_static; }
            __syncthreads();   // s_range is read by everyone before a later claim overwrites it
        }
    }
    if (dyn && !drew) {
        if (tid == 0) {
            draw(0, buf);
            if (C::NBUF == 1 && s_next[0] >= 0) load_vj(0, (s_nij[0] & 0xffff) * TJ);
        }
        __syncthreads();
    }
    if (dyn) {
        // ---- dynamic tail: single tiles from the shared counter ----
        for (int slot = 0;; slot ^= 1, buf ^= 1) {
            if (s_next[slot] < 0) break;
            const int I = s_nij[slot] >> 16, J = s_nij[slot] & 0xffff;
            if (I != Ivi) {
                if (tid == 0) load_vi(I * TI);
                vi_pending = true;
                Ivi = I;
            }
            do_tile(I, J, buf, slot ^ 1, -1);
            __syncthreads();
        }
    }
    if (ex.prof && blockIdx.x == 0 && tid == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        ex.prof[1] = t;
    }
    if (ex.cta_prof && tid == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        ex.cta_prof[3 * blockIdx.x + 1] = t;
    }
    pdl_trigger();
    pdl_wait();
}


// =========================================================================================
}  // namespace asb
