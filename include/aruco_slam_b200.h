/*
 * aruco_slam_b200 — C ABI of the B200-native MEKF/EKF + IPPE hot path.
 *
 * This header is the drop-in boundary: plain pointers and sizes, no torch / numpy
 * types.  The Python classes in aruco_slam_b200/filters.py bind it with ctypes and
 * mirror the reference's BaseFilter API; INTEGRATION.md shows the stub a reference
 * maintainer would add.  Every entry point cites the reference interface it replaces
 * (paths relative to the reference tree: filters/extended_kalman_filter.py = EKF,
 * filters/ekf_with_rotations.py = ROT, filters/base_filter.py = BASE).
 *
 * Conventions
 *   - all floating point is IEEE binary64 unless a parameter says float;
 *   - state layout  [c(3), q(4: w x y z), e(3), landmark_0(d) ... landmark_{n-1}(d)],
 *     d = 3 (EKF) or 10 (ROT: xyz, quat wxyz, err xyz); N = 10 + d*n;
 *   - covariance exchanged with the host is dense row-major N x N;
 *   - an observation row is [t(3), rvec(3)] exactly as BASE:168-171 stacks it;
 *   - every function returns ASB_OK (0) or a negative asb_status; asb_last_error()
 *     gives the message of the last failure on the calling thread;
 *   - a handle owns one CUDA stream; calls on one handle are not thread-safe (neither
 *     is the reference); functions that return host data synchronise that stream;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails
 *     with ASB_ERR_CUDA.
 */
#ifndef ARUCO_SLAM_B200_H
#define ARUCO_SLAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASB_ABI_VERSION 1

#if defined(__GNUC__)
#define ASB_API __attribute__((visibility("default")))
#else
#define ASB_API
#endif

typedef enum asb_status {
    ASB_OK = 0,
    ASB_ERR_BAD_HANDLE = -1,
    ASB_ERR_BAD_ARG = -2,
    ASB_ERR_UNKNOWN_ID = -3,   /* update() saw an id that is not in the map: KeyError at EKF:180 */
    ASB_ERR_CAPACITY = -4,     /* more landmarks / observations than the handle was created for */
    ASB_ERR_NOT_SPD = -5,      /* innovation covariance S not positive definite / non-finite */
    ASB_ERR_CUDA = -6,
    ASB_ERR_NO_STREAM = -7     /* stream_run before stream_upload */
} asb_status;

typedef enum asb_kind {
    ASB_KIND_EKF = 0,          /* class EKF            (EKF:37)  d = 3,  3 rows / marker */
    ASB_KIND_ROT = 1           /* class EKF_Rotations  (ROT:43)  d = 10, 7 rows / marker */
} asb_kind;

typedef enum asb_quat_mode {
    ASB_QUAT_AS_WRITTEN = 0,   /* EKF:145-146 as written: scalar-first arrays parsed scalar-last */
    ASB_QUAT_SCALAR_FIRST = 1  /* those two calls with scalar_first=True (ROT:150-151 convention) */
} asb_quat_mode;               /* ignored for ASB_KIND_ROT, which is always scalar-first */

/* Module-level constants of the reference (EKF:21-27, ROT:23-31). NULL => defaults of the kind. */
typedef struct asb_noise {
    double init_cam;   /* INITIAL_CAMERA_UNCERTAINTY   0.1 */
    double init_lm;    /* INITIAL_LANDMARK_UNCERTAINTY 0.7 */
    double r;          /* R_UNCERTAINTY                0.9 */
    double q_cam;      /* Q_UNCERTAINTY_CAM            0.3 (EKF) / 0.2 (ROT) */
    double q_err;      /* Q_ERROR_UNCERTAINTY_CAM      0.5 */
    double q_lm;       /* Q_UNCERTAINTY_LM[_XYZ]       0.01 */
} asb_noise;

typedef struct asb_filter asb_filter;   /* one filter, covariance resident in HBM */
typedef struct asb_batch asb_batch;     /* many independent filters, one CTA each   */

/* ---- library ----------------------------------------------------------------- */
ASB_API int asb_abi_version(void);
ASB_API const char* asb_last_error(void);
ASB_API int asb_device_count(int* count);
/* Fill a noise struct with the reference constants of `kind`. */
ASB_API int asb_default_noise(int kind, asb_noise* out);

/* ---- single filter: EKF.__init__ (EKF:40-56) / EKF_Rotations.__init__ (ROT:46-64) ---- */
ASB_API int asb_filter_create(int kind, int quat_mode, const double pose0[10], const asb_noise* cfg,
                      int max_landmarks, int device, asb_filter** out);
ASB_API int asb_filter_destroy(asb_filter* f);
/* Run on a caller-owned CUDA stream (cudaStream_t cast to void*); NULL => the handle's own. */
ASB_API int asb_filter_set_stream(asb_filter* f, void* cuda_stream);
ASB_API int asb_filter_synchronize(asb_filter* f);

/* observe (EKF:58-82, ROT:66-90): add unseen ids in frame order, predict, update.
 * ids (k) int32, poses (k,6) row-major, HOST pointers. Asynchronous on the stream. */
ASB_API int asb_observe(asb_filter* f, const int32_t* ids, const double* poses, int k);
/* process_frame's hot path in one call (BASE:196-204): estimate_pose_of_markers (BASE:92-171) on
 * the k detected corner sets (k,4,2 float32, HOST) followed by observe(ids, poses); the poses stay
 * on the device in between.  poses_out (k,6) or NULL: the detected poses for the caller (BASE:212);
 * the call then waits for the IPPE kernel only, the update stays asynchronous. */
ASB_API int asb_observe_corners(asb_filter* f, const int32_t* ids, const float* corners, int k,
                        const double camera_matrix[9], const double dist_coeffs[5], double marker_size,
                        double* poses_out);
/* predict (EKF:95-105, ROT:103-113): P += diag(Q); the state is unchanged. */
ASB_API int asb_predict(asb_filter* f);
/* update (EKF:107-156, ROT:115-181). Unknown id => ASB_ERR_UNKNOWN_ID, nothing changed. */
ASB_API int asb_update(asb_filter* f, const int32_t* ids, const double* poses, int k);
/* add_marker (EKF:239-290, ROT:275-335). `uncertainity` = d diagonal values or NULL
 * (=> init_lm).  Re-adding a known id re-binds the id to a new index, as the reference's
 * dict assignment does. */
ASB_API int asb_add_marker(asb_filter* f, int32_t id, const double pose[6], const double* uncertainity);

/* Device-resident observation stream (bench "value" path: inputs already in HBM).
 * offsets (T+1) int64 prefix sums; ids/poses packed over all frames.  upload associates
 * ids on the host (first-seen order, EKF:75-78) and copies once; run enqueues frames
 * [t0,t1) with no host<->device traffic. */
ASB_API int asb_stream_upload(asb_filter* f, const int32_t* ids, const double* poses,
                      const int64_t* offsets, int n_frames);
ASB_API int asb_stream_run(asb_filter* f, int t0, int t1);

/* state / uncertainty / landmarks attributes (EKF:46-51), get_poses (EKF:84-89),
 * get_lm_uncertainties (EKF:91-93), get_lm_estimates (EKF:355-357). */
ASB_API int asb_get_dims(asb_filter* f, int* n_landmarks, int* n_state);
ASB_API int asb_get_state(asb_filter* f, double* out, int n_state);
ASB_API int asb_set_state(asb_filter* f, const double* in, int n_state);
ASB_API int asb_get_cov(asb_filter* f, double* out, int n_state);            /* N*N row-major */
ASB_API int asb_set_cov(asb_filter* f, const double* in, int n_state);
ASB_API int asb_get_cov_diag(asb_filter* f, double* out, int n_state);
ASB_API int asb_get_landmark_ids(asb_filter* f, int32_t* ids_in_index_order, int capacity);
/* Replace the id->index map (ids in index order); state/cov are then set with the
 * setters above.  This is the checkpoint / re-synchronisation path (SURVEY §8c ii). */
ASB_API int asb_set_landmarks(asb_filter* f, const int32_t* ids_in_index_order, int n);
/* Raw device pointers for zero-copy consumers (torch.from_blob / cupy): state (N),
 * covariance (row-major with leading dimension *ld doubles; only the LOWER triangle, i >= j, is
 * kept up to date by the update kernels - asb_get_cov returns the full symmetric matrix). */
ASB_API int asb_device_views(asb_filter* f, void** d_state, void** d_cov, int* ld);

/* Opt-in moving-average motion model.  The reference's README (README.md:55-56) describes
 * X_k|k-1 = X_k-1 + (X_k-1 - X_k-n) / n on the camera position with n = MOVING_AVG_WINDOW = 10 (EKF:19), but the
 * constant is never read and predict() leaves the state alone (EKF:95-105) - so window = 0 (the default)
 * IS the reference's behaviour and there is nothing to check parity of window > 0 against.  With window > 0 every
 * observe() first moves the camera position by that step (n = the positions held so far, at most `window`), before
 * augmentation and update.  Per-frame observe() only (asb_stream_run refuses); not in the batched engines. */
ASB_API int asb_set_motion_model(asb_filter* f, int window);

/* Engine choice of asb_observe / asb_observe_corners (no counterpart in the reference: EKF.observe, EKF:58-82, is one
 * code path).  A map of at most n_state_max state dimensions whose covariance fits one SM's shared memory is updated
 * by ONE kernel launch per frame (the resident batched kernel on a single sequence: the reference's own use case,
 * main/run_slam.py with a few dozen markers); larger maps run the pipelined k_f1 / k_f2 / k_sweep engine over the
 * whole GPU.  Default 192 (per-frame latency of the two engines crosses near N = 190 on a B200,
 * tools/small_map_exp.py); 0 = always the pipelined engine.  Both are checked against the same goldens. */
ASB_API int asb_set_small_map_limit(asb_filter* f, int n_state_max);

/* correct_poses (BASE:274-325, the orientation-ambiguity gate the reference defines and never calls; SURVEY 8f
 * row 3), EKF_Rotations only: observation o is kept (choice[o] = 1) when the angle between R(q_cam) *
 * from_euler('xyz', pose[3:6]) and the landmark's orientation in the state is below threshold_deg
 * (QUAT_THRESHOLD = 50, BASE:32); unknown ids are dropped (0).  poses2_or_null: the second IPPE solution per
 * observation (asb_ippe_square_both); it is tried when the first fails (choice 2) - an extension, NULL gives
 * exactly the reference's function.  Waits for the updates queued on the handle (the state must be final). */
ASB_API int asb_correct_poses(asb_filter* f, const int32_t* ids, const double* poses, const double* poses2_or_null,
                      int k, double threshold_deg, int32_t* choice_out);

/* Intermediates of the last update(), for parity tests (parse_poses EKF:158-200):
 * nu = z - h (m), blocks = per observation [r x 10 camera block | r x d landmark block]
 * row-major (k * r * (10+d)), dx = K nu (N).  Any pointer may be NULL. */
ASB_API int asb_get_last_update(asb_filter* f, double* nu, double* blocks, double* dx, int* m);

/* ---- batched independent filters (BASELINE configs C2 / C5) --------------------- */
ASB_API int asb_batch_create(int kind, int quat_mode, const double pose0[10], const asb_noise* cfg,
                     int n_seq, int max_landmarks, int max_obs_per_frame, int device,
                     asb_batch** out);
ASB_API int asb_batch_destroy(asb_batch* b);
/* Back to the constructor's state (pose0, P = init_cam I10, empty maps), asynchronous. */
ASB_API int asb_batch_reset(asb_batch* b);
/* Engine behind the handle (no reference counterpart; both give the reference's results).
 * 1 = resident: one CTA per sequence, the whole filter in shared memory for all frames (EKF up to 64 markers);
 * 2 = wide: frame-synchronous kernels over all sequences, covariance tiled in HBM (EKF_Rotations at 32 / 64
 *     markers, whose covariance does not fit an SM);  0 = automatic (resident when it fits).
 * set_engine resets the filters to their constructor state. */
ASB_API int asb_batch_set_engine(asb_batch* b, int engine);
ASB_API int asb_batch_get_engine(asb_batch* b, int* engine);
ASB_API int asb_batch_set_stream(asb_batch* b, void* cuda_stream);
ASB_API int asb_batch_synchronize(asb_batch* b);
/* offsets (n_seq*n_frames + 1) int64: frame t of sequence s = rows
 * [offsets[s*n_frames+t], offsets[s*n_frames+t+1]) of ids / poses (HOST pointers). */
ASB_API int asb_batch_upload(asb_batch* b, const int32_t* ids, const double* poses,
                     const int64_t* offsets, int n_frames);
/* Same with n_streams < n_seq distinct observation streams (n_seq a multiple of n_streams): sequence s
 * reads stream s % n_streams, offsets has n_streams*n_frames + 1 entries.  For large synthetic batches whose
 * host copy would not fit (BASELINE configs[4]: 32768 sequences x 64 markers x 1000 frames = 25 GB of poses);
 * every sequence still runs its own filter. */
ASB_API int asb_batch_upload_tiled(asb_batch* b, const int32_t* ids, const double* poses,
                           const int64_t* offsets, int n_frames, int n_streams);
/* observe() for frames [t0,t1) of every sequence; id association on the device. */
ASB_API int asb_batch_run(asb_batch* b, int t0, int t1);
ASB_API int asb_batch_get_dims(asb_batch* b, int seq, int* n_landmarks, int* n_state);
ASB_API int asb_batch_get_state(asb_batch* b, int seq, double* out, int n_state);
ASB_API int asb_batch_get_cov(asb_batch* b, int seq, double* out, int n_state);
ASB_API int asb_batch_get_landmark_ids(asb_batch* b, int seq, int32_t* ids_in_index_order, int capacity);
/* Checkpoint / re-synchronisation of one sequence (the batch counterpart of asb_set_landmarks / _set_state / _set_cov;
 * SURVEY 8c ii): landmarks first (sets the dimension), then state (n_state) and covariance (dense n x n, lower
 * triangle taken).  Synchronous. */
ASB_API int asb_batch_set_landmarks(asb_batch* b, int seq, const int32_t* ids_in_index_order, int n);
ASB_API int asb_batch_set_state(asb_batch* b, int seq, const double* in, int n_state);
ASB_API int asb_batch_set_cov(asb_batch* b, int seq, const double* in, int n_state);
/* camera pose (10) of every sequence, (n_seq,10) row-major, + per-sequence status. */
ASB_API int asb_batch_get_cameras(asb_batch* b, double* out, int32_t* status_or_null);
ASB_API int asb_batch_device_views(asb_batch* b, void** d_states, void** d_covs, int* state_stride,
                           int* cov_ld);

/* Clocks CTA 0 spent per phase since the last call (association, augment+predict, model, G, S,
 * Cholesky, solve, -, P update, apply), measurement aid. */
ASB_API int asb_batch_profile(asb_batch* b, double clocks[10]);

/* ---- row-sharded filter: ONE EKF / EKF_Rotations whose covariance rows are spread over the GPUs
 *      of a box, one process (rank) per GPU (BASELINE config C3 at N GPUs).
 * Same algorithm and same results as asb_filter (observe = EKF:58-82 / ROT:66-90); the state and
 * the id map are replicated, rank g owns rows [row0,row1) of `uncertainty` (EKF:48, EKF:155-156).
 * Every rank must make the same calls with the same arguments, in the same order: each update
 * runs two NCCL collectives (all-reduce of P[T,T], all-gather of the K H P panel) on the stream.
 * NCCL is opened at run time (libnccl.so.2, the copy torch has loaded when called from Python). */
typedef struct asb_shard asb_shard;
/* The row partition (callable without a GPU): rows [*row0,*row1) of rank `rank`, and the padded
 * dimension *ld (= world * rows per rank >= 10 + d*max_landmarks, rows per rank a multiple of 64). */
ASB_API int asb_shard_partition(int kind, int max_landmarks, int world, int rank, int* row0, int* row1, int* ld);
/* ncclGetUniqueId: call on rank 0, broadcast the 128 bytes to every rank (torch.distributed, MPI ...). */
ASB_API int asb_shard_unique_id(void* out128, int bytes);
ASB_API int asb_shard_create(int kind, int quat_mode, const double pose0[10], const asb_noise* cfg,
                     int max_landmarks, int device, int rank, int world, const void* unique_id128,
                     asb_shard** out);
ASB_API int asb_shard_destroy(asb_shard* s);
ASB_API int asb_shard_set_stream(asb_shard* s, void* cuda_stream);
ASB_API int asb_shard_synchronize(asb_shard* s);
ASB_API int asb_shard_observe(asb_shard* s, const int32_t* ids, const double* poses, int k);
ASB_API int asb_shard_stream_upload(asb_shard* s, const int32_t* ids, const double* poses,
                            const int64_t* offsets, int n_frames);
ASB_API int asb_shard_stream_run(asb_shard* s, int t0, int t1);
ASB_API int asb_shard_get_dims(asb_shard* s, int* n_landmarks, int* n_state);
ASB_API int asb_shard_get_state(asb_shard* s, double* out, int n_state);
/* This rank's live rows of `uncertainty`: (*n_rows, N) row-major, rows row0 .. min(row1, N). */
ASB_API int asb_shard_get_cov_rows(asb_shard* s, double* out, long long capacity, int* n_rows);
ASB_API int asb_shard_get_landmark_ids(asb_shard* s, int32_t* ids_in_index_order, int capacity);

/* ---- square-marker pose: estimate_pose_of_markers (BASE:92-171), i.e.
 *      cv2.solvePnP(marker_points, c, K, dist, flags=SOLVEPNP_IPPE_SQUARE) per marker ----
 * corners (n,4,2) float32 pixel corners in the order BASE:112-120 expects; camera_matrix
 * 3x3 row-major; dist_coeffs k1 k2 p1 p2 k3; out (n,6) = [tvec, rvec] rows (BASE:171). */
ASB_API int asb_ippe_square(const float* corners, int64_t n, const double camera_matrix[9],
                    const double dist_coeffs[5], double marker_size, double* out, int device);
/* cv2.solvePnPGeneric(..., SOLVEPNP_IPPE_SQUARE): both IPPE solutions per corner set, the one with the
 * smaller reprojection error first (out), the other in out2 (n,6); errs_or_null (n,2) float32 = the two
 * summed squared reprojection errors in normalised coordinates the ordering was decided on. */
ASB_API int asb_ippe_square_both(const float* corners, int64_t n, const double camera_matrix[9],
                         const double dist_coeffs[5], double marker_size, double* out, double* out2,
                         float* errs_or_null, int device);
/* Same on DEVICE buffers, asynchronous on `cuda_stream`.  out2_or_null receives the
 * second IPPE solution (n,6) and err_or_null (n,2) the two float reprojection errors. */
ASB_API int asb_ippe_square_device(const float* d_corners, int64_t n, const double camera_matrix[9],
                           const double dist_coeffs[5], double marker_size, double* d_out,
                           double* d_out2_or_null, float* d_err_or_null, void* cuda_stream);

/* Allocation-free variant of asb_ippe_square.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture,
 * tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * asb_ippe_square (HOST buffers), but the device staging is a file-scope static __device__ array of
 * ASB_IPPE_STATIC_MAX corner sets x 80 B = 1.25 MiB (held by every GPU that loads the library until
 * exit) instead of the grow-only cudaMalloc cache; the call makes no cudaMalloc / cudaFree / allocating
 * library call.  n > ASB_IPPE_STATIC_MAX => ASB_ERR_CAPACITY.  All calls share the one array (it keeps
 * the previous call's contents; rows [0,n) are overwritten, nothing needs zeroing); callers are
 * serialised by a mutex on the legacy default stream - work that overlaps on the GPU would race. */
#define ASB_IPPE_STATIC_MAX 16384
ASB_API int asb_ippe_square_static(const float* corners, int64_t n, const double camera_matrix[9],
                           const double dist_coeffs[5], double marker_size, double* out, int device);

/* ---- measurement helpers (bench / profiles) -------------------------------------- */
/* FP64 pipe micro-benchmarks: achieved TFLOP/s of dependent-free DFMA chains and of
 * mma.sync m8n8k4 f64 (DMMA), timed with CUDA events. which: 0 = DFMA, 1 = DMMA, 2 = both. */
ASB_API int asb_fp64_peak(int device, int which, double* tflops);
/* The covariance sweep kernel alone on this filter's current P (state dims as they are),
 * `reps` back-to-back launches timed with CUDA events.  config 0 = 128x64 tiles 1 CTA/SM,
 * 1 = 64x64 tiles 2 CTAs/SM.  dbg != 0 disables parts of the kernel (experiments; the
 * filter's P is garbage afterwards).  V is whatever the last update left in the panel. */
ASB_API int asb_bench_sweep(asb_filter* f, int m4, int config, int dbg, int reps, double* us_per_launch);
/* Durations (us, globaltimer) of the phases of the last k_front launch as seen by CTA 0:
 * index set + copy issue, model + copy wait, S build, Cholesky, G slice, V = W G, panel store +
 * growth; then us[7..11] = clocks of one Cholesky step: diagonal block, panel, barrier,
 * trailing update, barrier. */
ASB_API int asb_front_profile(asb_filter* f, double us[12]);
/* Pipelined engine: globaltimer stamps (us, relative to the earliest) of the last two updates, the
 * older one first; per update 16 slots: [0..5] k_f1 phases, [8..11] k_f2 phases, [12] / [13] start / end
 * of CTA 0 of k_sweep; -1 = not written. */
ASB_API int asb_timeline(asb_filter* f, double us[32]);
/* Per-CTA (start us, end us, %smid) of the last k_sweep, n CTAs; needs ASB_CTA_PROF=1 at creation. */
ASB_API int asb_cta_profile(asb_filter* f, double* us, int n);
/* Dependent-issue latencies in clocks: DFMA, rsqrt(double)+add, LDS.64 (with cvt), BAR.SYNC of 256 threads. */
ASB_API int asb_latency_probe(int device, double clocks[4]);
/* HBM stream bandwidth (GB/s of bytes moved): mode 0 write-only, 1 read-only, 2 one read + two writes
 * (the sweep's mix), 256-bit accesses. */
ASB_API int asb_stream_bandwidth(int device, int64_t bytes, int mode, double* gbs);
/* The sweep's memory pattern with no math (read lower-triangle tile, write it, write its mirror
 * location) on an n x n FP64 matrix with tile = 64 / 128 / 256: best time in us. */
ASB_API int asb_tilepattern_bandwidth(int device, int n, int tile, int ctas_per_sm, double* us);
/* Streaming copy bandwidth (GB/s, read+write bytes) over `bytes` of HBM. */
ASB_API int asb_copy_bandwidth(int device, int64_t bytes, double* gbs);

#ifdef __cplusplus
}
#endif
#endif /* ARUCO_SLAM_B200_H */
