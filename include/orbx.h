/* orbx.h -- C ABI of the B200-native ORB front-end (liborbx.so).
 *
 * This is the drop-in boundary for the reference's feature front-end.  Every entry point
 * is `extern "C"`, takes plain pointers and sizes, never throws and never aborts; the C++
 * classes in include/Features/ and include/SlamTypes/ (same names and signatures as the
 * reference) are thin host wrappers over these calls.  There is no CPU fallback: every
 * compute entry runs hand-written sm_100a kernels and returns ORBX_ERR_CUDA if no device
 * is usable.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   orbx_create / orbx_get_level_tables   ORBextractor::ORBextractor + accessors
 *                                         Features/ORBextractor.hpp:68-108, .cpp:492-595
 *   orbx_extract / orbx_extract_batch     ORBextractor::operator()      Features/ORBextractor.cpp:1531-1653
 *   orbx_get_pyramid_level                public member mvImagePyramid  Features/ORBextractor.hpp:111
 *   orbx_set_camera / orbx_frame_glue     Frame::Frame ctor glue        SlamTypes/Frame.cpp:40-161
 *   orbx_features_in_area                 Frame::GetFeaturesInArea      SlamTypes/Frame.cpp:163-206
 *   orbx_search_for_initialization        ORBmatcher::SearchForInitialization  Features/ORBmatcher.cpp:11-150
 *   orbx_descriptor_distance              DBoW2::FORB::distance         Thirdparty/DBoW2/src/FORB.cpp:77-100
 *   orbx_check_homography / _fundamental  Initializer::CheckHomography / CheckFundamental (+ the RANSAC selection)
 *   orbx_stream_score_models              Initialization/Initializer.cpp:268-438, :205-209, :257-261
 *   orbx_stream_*                         the demo's per-pair sequence  demo/demo_initialization.cpp:65-108
 *                                         kept resident on the device (throughput path)
 *
 * Threading: an orbx_ctx is not thread-safe; distinct contexts are independent (one per
 * host thread / GPU), each owning its stream and all of its device memory.
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes (all entry points returning int) */
#define ORBX_OK 0
#define ORBX_ERR_ARG (-1)        /* null pointer, bad size, image shape differs from the context's */
#define ORBX_ERR_EMPTY_IMAGE (-2)/* reference: `if (_image.empty()) return -1;`  ORBextractor.cpp:1536 */
#define ORBX_ERR_CTOR (-3)       /* reference would exit(1): scaleFactor == 1 with nlevels > 1, .cpp:502-505 */
#define ORBX_ERR_CUDA (-4)       /* CUDA runtime error; see orbx_last_error() */
#define ORBX_ERR_CAPACITY (-5)   /* caller's output buffer or an internal table is too small */
#define ORBX_ERR_ASPECT (-6)     /* level aspect ratio rounds to 0 roots (reference indexes an empty vector, .cpp:706) */
#define ORBX_ERR_STATE (-7)      /* call order (e.g. glue before orbx_set_camera) */

#define ORBX_MAX_LEVELS 16
#define ORBX_GRID_COLS 64 /* FRAME_GRID_COLS, SlamTypes/Frame.hpp:16 */
#define ORBX_GRID_ROWS 48 /* FRAME_GRID_ROWS, SlamTypes/Frame.hpp:15 */
#define ORBX_GRID_CELLS (ORBX_GRID_COLS * ORBX_GRID_ROWS)

/* Byte-compatible with cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle, response, octave, class_id. */
typedef struct orbx_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orbx_keypoint;

/* A Frame as the matcher sees it (host memory, borrowed).  grid_offsets has
 * ORBX_GRID_CELLS + 1 entries in cell order ix * ORBX_GRID_ROWS + iy; grid_indices lists the
 * keypoint indices of each cell in ascending order (= mGrid[ix][iy], Frame.hpp:98). */
typedef struct orbx_frame {
  int32_t n;
  const orbx_keypoint* keys_un; /* mvKeysUn */
  const uint8_t* desc;          /* mDescriptors, n x 32, continuous */
  const uint32_t* grid_offsets;
  const uint32_t* grid_indices;
} orbx_frame;

typedef struct orbx_ctx orbx_ctx;

/* ---- lifetime ------------------------------------------------------------------------ */
/* One context = one extractor configuration on one image shape on one GPU.  max_batch is the
 * number of frames processed per launch sequence (device memory is sized for it once).
 * Returns NULL on failure; *status (may be NULL) receives the reason. */
orbx_ctx* orbx_create(int device, int width, int height, int nfeatures, float scale_factor, int nlevels,
                      int ini_th_fast, int min_th_fast, int max_batch, int* status);
void orbx_destroy(orbx_ctx* ctx);
const char* orbx_last_error(const orbx_ctx* ctx); /* never NULL */
/* cos / sin of the keypoint angle in computeOrbDescriptor (Features/ORBextractor.cpp:173-174: the float overloads, glibc's
 * cosf / sinf).  2 [default]: glibc's algorithm restated on the device, bit-equal to the host libm on every float of
 * [0, 2 pi]; 0: (float)cos((double)a); 1: CUDA's cosf(a).  See DESIGN.md section 5. */
int orbx_set_trig_mode(orbx_ctx* ctx, int mode);
/* Test hook: sin / cos of n host floats computed exactly as trig mode 2 does on the device (host buffers in and out). */
int orbx_sincosf_probe(orbx_ctx* ctx, const float* y, long long n, float* sin_out, float* cos_out);
/* Keypoint selection policy of the context: 0 = ComputeKeyPointsOctTree (default; what the reference's operator() calls,
 * Features/ORBextractor.cpp:1552), 1 = ComputeKeyPointsOld (:1207-1496, the alternative the reference keeps behind a
 * commented-out call, :1555): its own cell grid, the minThFAST retry for cells with at most 3 keypoints, per-cell quotas
 * and cv::KeyPointsFilter::retainBest (libstdc++'s std::nth_element order).  Everything else of operator() is shared.
 * Returns ORBX_ERR_ARG for shapes on which that function divides by zero (a level quota too small for one cell column),
 * whose cells are wider than 480 tested columns, or whose cell grid runs past the tested region of a level (very wide or
 * very tall images: the reference then keeps keypoints closer than 19 px to the border, whose descriptors it reads from
 * outside the blurred level, and asserts in cv::Mat::rowRange when a cell leaves the image). */
int orbx_set_keypoint_policy(orbx_ctx* ctx, int policy);
/* Schedule options of a context (they never change results).  The environment is read ONCE, in orbx_create
 * (ORBX_SINGLE_STREAM, ORBX_RUN_PARTS, ORBX_LANES, ORBX_RESIZE_ROWS, ORBX_MATCH_SERIAL, ORBX_FAST_TILE, ORBX_PYRAMID_CHAIN); this call
 * changes an option of a live context: name = "single_stream" | "run_parts" | "lanes" | "resize_rows" | "match_serial" |
 * "match_path" (0 auto, 1 k_match_cta, 2 k_match, 3 brute force) | "frame_graph" (1: single-frame calls use a CUDA graph) |
 * "qt_tight" (test hook: the quadtree squeezes erased nodes out of its table at the smallest size its bound allows) |
 * "pyramid_chain" (1, default: batches of at most 2 frames build their pyramids in one launch with a grid barrier between levels).
 * No launch path reads the environment.  Returns ORBX_ERR_ARG for an unknown name. */
int orbx_set_option(orbx_ctx* ctx, const char* name, int value);

/* ---- extractor tables (host arithmetic of the reference constructor) -------------------- */
/* Any output pointer may be NULL.  Arrays have nlevels entries; umax has 16.  Returns nlevels. */
int orbx_get_level_tables(const orbx_ctx* ctx, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int* n_per_level, int* umax16);
int orbx_get_level_size(const orbx_ctx* ctx, int level, int* w, int* h);
/* Same tables without a device (used by the C++ wrapper's accessors before the first image). */
int orbx_compute_level_tables(int nfeatures, float scale_factor, int nlevels, float* scale, float* inv_scale,
                              float* sigma2, float* inv_sigma2, int* n_per_level, int* umax16);

/* ---- extraction ---------------------------------------------------------------------------- */
/* img: 8-bit single channel, ctx's width x height, row stride `step` bytes (host memory).
 * kps/desc: caller-owned, room for `cap` keypoints (cap >= nfeatures is always enough).
 * lapping: vLappingArea {x0, x1} (NULL = {0, 0}).  *n_out = number of keypoints,
 * *mono_index = the reference's return value (may be NULL). */
int orbx_extract(orbx_ctx* ctx, const uint8_t* img, size_t step, const int lapping[2], orbx_keypoint* kps,
                 uint8_t* desc, int cap, int* n_out, int* mono_index);
/* Frame::Frame in one call (SlamTypes/Frame.cpp:57-77): extraction + UndistortKeyPoints + grid assignment of one frame
 * through ONE CUDA graph launch and ONE synchronisation (the single-frame latency path; orbx_extract of one image runs
 * through the same kind of graph).  Same results as orbx_extract followed by orbx_frame_glue.  kps_un / grid_offsets
 * (ORBX_GRID_CELLS + 1) / grid_indices (cap) may be NULL.  Needs orbx_set_camera. */
int orbx_frame_construct(orbx_ctx* ctx, const uint8_t* img, size_t step, const int lapping[2], orbx_keypoint* kps,
                         uint8_t* desc, int cap, int* n_out, int* mono_index, orbx_keypoint* kps_un,
                         uint32_t* grid_offsets, uint32_t* grid_indices);
/* n_img <= max_batch images; outputs are n_img consecutive blocks of `cap` keypoints. */
int orbx_extract_batch(orbx_ctx* ctx, const uint8_t* const* imgs, size_t step, int n_img, const int lapping[2],
                       orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index);

/* Inspection of the last batch (parity tests; mvImagePyramid).  slot = index in the batch.
 * padded != 0: (w+38) x (h+38) including the 19-px BORDER_REFLECT_101 frame of
 * ComputePyramid (.cpp:1689,1708); else w x h. */
int orbx_get_pyramid_level(orbx_ctx* ctx, int slot, int level, int padded, uint8_t* dst, size_t dst_step);
int orbx_get_blurred_level(orbx_ctx* ctx, int slot, int level, uint8_t* dst, size_t dst_step);
/* FAST candidates of a level in the reference's order (cell row-major, then y, then x), as
 * (x, y, response) float triples relative to (16, 16).  Returns the count (may exceed cap). */
int orbx_get_candidates(orbx_ctx* ctx, int slot, int level, float* xyr, int cap);
/* keypoints of one level after distribution + orientation (+ descriptors), level coordinates */
int orbx_get_level_keypoints(orbx_ctx* ctx, int slot, int level, orbx_keypoint* kps, uint8_t* desc, int cap);

/* Copy-only probe of the host link: reps x (in_bytes pinned host -> device, out_bytes device -> pinned host, both
 * directions at once, no kernels); *ms_per_rep = device time of one repetition.  bench.py runs it on every rank at the same
 * time and reports the e2e step against it (e2e.host_link_frac). */
void* orbx_host_alloc_ex(size_t bytes, int write_combined); /* page-locked; write-combined for upload-only buffers */
int orbx_hostlink_probe(orbx_ctx* ctx, const void* host_in, size_t in_bytes, void* host_out, size_t out_bytes, int reps,
                        float* ms_per_rep);

/* ---- radius queries on device-resident frames ------------------------------ */
/* Frame::GetFeaturesInArea (SlamTypes/Frame.cpp:163-206) for n_queries windows on the frames orbx_stream_run left on the
 * device: query i = queries5[5 i .. 5 i + 4] = (x, y, r, minLevel, maxLevel) on frame slots[i].  out[i * cap ..] receives
 * the keypoint indices in the reference's visiting order, counts[i] the result size (may exceed cap).  One launch, a
 * warp per query; nothing is uploaded but the queries. */
int orbx_stream_features_in_area(orbx_ctx* ctx, int n_queries, const int* slots, const float* queries5, int* out, int cap,
                                 int* counts);

/* ---- brute-force matching: the distance matrix of one pair sharded by query rows ---------- */
/* SearchForInitialization(F1, F2) with a window that makes most of F2 a candidate of every query is a
 * (level-0 keys of F1) x (level-0 keys of F2) Hamming matrix.  orbx_search_for_initialization computes it on
 * all SMs of one GPU; these two calls split it over GPUs by query rows (north_star: "Hamming matrix sharded by
 * query rows", no collective -- F2's descriptors are uploaded to every rank):
 *   orbx_match_rows     rows [q0, q1) of the matrix, reduced to each query's sorted list of its 7 best
 *                       candidates: rows_out[(q1 - q0)][8] words (7 keys = distance << 20 | visiting position,
 *                       1 info word).  *n_queries = number of level-0 keys of F1.  Returns the rows written.
 *   orbx_match_resolve  the reference's ordered loop (ORBmatcher.cpp:36-141) from the complete table
 *                       rows[n_queries][8] gathered from the ranks; results are identical to
 *                       orbx_search_for_initialization(F1, F2, ...). */
int orbx_match_rows(orbx_ctx* ctx, const orbx_frame* F1, const orbx_frame* F2, int window_size, int q0, int q1,
                    uint32_t* rows_out, int* n_queries);
int orbx_match_resolve(orbx_ctx* ctx, const orbx_frame* F1, const orbx_frame* F2, float nnratio, int check_orientation,
                       int window_size, const uint32_t* rows, int n_queries, int* matches12, int* nmatches,
                       int counters[3]);
/* Raw all-pairs 256-bit Hamming (FORB::distance over a matrix): for each of n_a descriptors the nearest of n_b
 * (first minimum in index order), its distance and the second smallest distance (-1 if n_b == 1).  Output
 * pointers may be NULL.  kernel_ms (may be NULL): device time of one pass, averaged over `repeats` passes. */
int orbx_hamming_best2(orbx_ctx* ctx, const uint8_t* a, int n_a, const uint8_t* b, int n_b, int* best_idx, int* best_dist,
                       int* second_dist, int repeats, float* kernel_ms);
/* The same result for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid allocation).  No counterpart in the reference.
 * d_a, d_b and the three outputs (each may be NULL) are DEVICE pointers on the context's device; `stream` is a
 * cudaStream_t (NULL = the context's stream); synchronize != 0 waits for the stream before returning, otherwise the two
 * kernels are only enqueued.  The call allocates nothing: the per-query table (8 words per query) lives in a fixed
 * device array inside the library, ORBX_NOALLOC_MAX_QUERIES rows = 512 KiB, held by every GPU that loads the library
 * until the process exits.  n_a beyond that limit returns ORBX_ERR_CAPACITY (text in orbx_last_error); split the
 * queries or use orbx_hamming_best2.  Every call on a GPU shares the array (it keeps the previous call's rows; each
 * call overwrites the rows it reads, nothing needs clearing): calls that overlap on the GPU race on it -- keep them on
 * one stream or order them with events.  The first call per device looks up the array's device address
 * (cudaGetSymbolAddress); make that call outside stream capture. */
#define ORBX_NOALLOC_MAX_QUERIES 16384
int orbx_hamming_best2_noalloc(orbx_ctx* ctx, const uint8_t* d_a, int n_a, const uint8_t* d_b, int n_b, int* d_best_idx,
                               int* d_best_dist, int* d_second_dist, void* stream, int synchronize);

/* ---- Frame glue ---------------------------------------------------------------------------- */
/* K4 = fx, fy, cx, cy; dist4 = k1, k2, p1, p2 (float32, as the reference's CV_32F matrices).
 * bounds = mnMinX, mnMaxX, mnMinY, mnMaxY or NULL to compute them as ComputeImageBounds does
 * (Frame.cpp:101-134).  bounds_out (may be NULL) receives the values in use. */
int orbx_set_camera(orbx_ctx* ctx, const float K4[4], const float dist4[4], const int bounds[4],
                    int bounds_out[4]);
/* UndistortKeyPoints + grid assignment (Frame.cpp:64-77) on the device. */
int orbx_frame_glue(orbx_ctx* ctx, const orbx_keypoint* kps, int n, orbx_keypoint* kps_un,
                    uint32_t* grid_offsets, uint32_t* grid_indices);
/* GetFeaturesInArea on a host frame (device kernel).  Returns the count (may exceed cap). */
int orbx_features_in_area(orbx_ctx* ctx, const orbx_frame* F, float x, float y, float r, int min_level,
                          int max_level, int* out, int cap);

/* ---- matcher ------------------------------------------------------------------------------- */
/* matches12: F1->n ints (-1 = none).  *nmatches = the reference's return value (with its
 * double-decrement quirk).  counters: rejected by distance, ratio, orientation (the three
 * values the reference prints, ORBmatcher.cpp:144-147); may be NULL. */
int orbx_search_for_initialization(orbx_ctx* ctx, const orbx_frame* F1, const orbx_frame* F2, float nnratio,
                                   int check_orientation, int window_size, int* matches12, int* nmatches,
                                   int counters[3]);
/* 256-bit Hamming distance of two 32-byte descriptors (host utility, no device needed). */
int orbx_descriptor_distance(const uint8_t a[32], const uint8_t b[32]);
/* n distances a[i] vs b[i] on the device (32 bytes each). */
int orbx_descriptor_distance_batch(orbx_ctx* ctx, const uint8_t* a, const uint8_t* b, int n, int* out);

/* ---- Initializer scoring (the step after the matcher) --------------------------- */
/* Initializer::CheckHomography (Initialization/Initializer.cpp:268-353) for n_hyp hypotheses on one pair.
 * kps1_un / kps2_un = mvKeysUn of the reference and the current frame, matches12 = vnMatches12 (n1 ints); the match
 * list mvMatches12 is the ascending i with matches12[i] >= 0 (Initializer.cpp:27-34), *n_list its length N.
 * H21, H12 (= its inverse, as the caller computed it): n_hyp x 9 floats, row-major.  scores: n_hyp floats, bit-equal
 * to the reference's.  inliers (may be NULL): n_hyp x N bytes, vbMatchesInliers per hypothesis.  *best (may be NULL):
 * the selection of the RANSAC loop FindHomography (:205-209) over the set -- the first hypothesis whose score is
 * strictly larger than every earlier one and than 0, or -1. */
int orbx_check_homography(orbx_ctx* ctx, const orbx_keypoint* kps1_un, int n1, const orbx_keypoint* kps2_un, int n2,
                          const int* matches12, const float* H21, const float* H12, int n_hyp, float sigma,
                          float* scores, uint8_t* inliers, int* n_list, int* best);
/* Initializer::CheckFundamental (Initializer.cpp:355-438); selection as FindFundamental (:257-261). */
int orbx_check_fundamental(orbx_ctx* ctx, const orbx_keypoint* kps1_un, int n1, const orbx_keypoint* kps2_un, int n2,
                           const int* matches12, const float* F21, int n_hyp, float sigma, float* scores,
                           uint8_t* inliers, int* n_list, int* best);
/* Both models on the first n_pairs pairs still resident from the last orbx_stream_run / orbx_stream_process (their
 * undistorted keys and vnMatches12 never leave the device).  H21, H12, F21: n_pairs x n_hyp x 9 floats (H21 / H12 or
 * F21 may be NULL to skip a model).  scores_*: n_pairs x n_hyp; best_*: n_pairs; inliers_*: blocks of `cap` bytes per
 * pair holding the flags of the selected hypothesis by list position (0 beyond N or when best = -1); n_list: n_pairs.
 * Any output may be NULL. */
int orbx_stream_score_models(orbx_ctx* ctx, int n_pairs, const float* H21, const float* H12, const float* F21,
                             int n_hyp, float sigma, float* scores_h, float* scores_f, int* best_h, int* best_f,
                             uint8_t* inliers_h, uint8_t* inliers_f, int cap, int* n_list);

/* ---- device-resident stream (throughput path) ---------------------------------------------- */
/* The frames of a batch stay in HBM from the image to the match list: extract -> glue ->
 * SearchForInitialization on the pairs (2k, 2k+1).  orbx_set_camera must have been called. */
/* copy n_img continuous width x height images (host, ideally pinned) into the batch slots */
int orbx_stream_upload(orbx_ctx* ctx, const uint8_t* imgs, int n_img);
/* n_img interleaved 3-channel frames (W*H*3 bytes each); level 0 = Converter::toGray (Utils/Converter.cpp:5-19),
 * i.e. cv::cvtColor BGR2GRAY (rgb = 0, demo/demo_initialization.cpp:67-68) or RGB2GRAY (rgb != 0), on the device */
int orbx_stream_upload_color(orbx_ctx* ctx, const uint8_t* imgs, int n_img, int rgb);
/* enqueue the whole pipeline on the context's stream for the first n_img slots; asynchronous */
int orbx_stream_run(orbx_ctx* ctx, int n_img, float nnratio, int check_orientation, int window_size);
/* wait, then copy results to the host.  Any pointer may be NULL.  kps/kps_un/desc: blocks of
 * `cap` per frame; matches12: blocks of `cap` per pair; nmatches/counters per pair.
 * Rows are copied whole, so `cap` must hold a frame's keypoint capacity (max(nfeatures, sum of the level quotas)):
 * a smaller `cap` with any of the four array pointers set returns ORBX_ERR_CAPACITY instead of truncating
 * (also for orbx_stream_process / orbx_stream_process_async). */
int orbx_stream_download(orbx_ctx* ctx, int n_img, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc,
                         int cap, int* n_out, int* matches12, int* nmatches, int* counters3);
/* compact result of the e2e path: per pair nmatches + matches12 only */
/* Host frames in, everything out, in one call: n_img frames (W*H bytes each, contiguous) are cut into
 * chunks of `chunk` frames (even; <= 0 picks 128) whose H2D copy, kernels and D2H copy overlap on three
 * streams.  Buffers as in orbx_stream_download; page-locked memory (orbx_host_alloc) keeps the copies
 * asynchronous.  Replaces the per-pair sequence of demo/demo_initialization.cpp:65-108 for a stream. */
int orbx_stream_process(orbx_ctx* ctx, const uint8_t* imgs, int n_img, float nnratio, int check_orientation,
                        int window_size, int chunk, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc, int cap,
                        int* n_out, int* matches12, int* nmatches, int* counters3);
/* Streaming form of orbx_stream_process: returns a ticket (0 or 1; negative = error) as soon as the step is enqueued.
 * chunk <= 0 cuts the step into three equal chunks, one per compute lane.  Up to two steps are in flight, so the upload of step s+1 overlaps the kernels and the download of step s; the caller
 * keeps one set of input / output buffers per ticket untouched until orbx_stream_wait(ticket) has returned (it blocks
 * until that step's results are on the host and returns its status).  Other orbx_stream_* calls are ordered after the
 * steps in flight; orbx_stream_sync waits for all of them. */
int orbx_stream_process_async(orbx_ctx* ctx, const uint8_t* imgs, int n_img, float nnratio, int check_orientation,
                              int window_size, int chunk, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc, int cap,
                              int* n_out, int* matches12, int* nmatches, int* counters3);
int orbx_stream_wait(orbx_ctx* ctx, int ticket);
int orbx_stream_sync(orbx_ctx* ctx);
/* CUDA-event timing on the context's stream: start/stop bracket enqueued work; elapsed in ms.
 * which = 0: whole bracket; 1..: per-stage timers recorded inside orbx_stream_run when
 * enabled with orbx_set_stage_timing(ctx, 1). */
int orbx_timer_start(orbx_ctx* ctx);
int orbx_timer_stop(orbx_ctx* ctx, float* ms);
int orbx_set_stage_timing(orbx_ctx* ctx, int enable);
/* 8 slots: pyramid, fast, quadtree, blur, describe+pack, scoring kernels of the last orbx_stream_score_models, glue, match */
int orbx_get_stage_times(orbx_ctx* ctx, float* ms8);
/* number of kernel launches issued by the last orbx_stream_run / orbx_extract_batch */
int orbx_last_launch_count(const orbx_ctx* ctx);
/* page-locked host buffers for orbx_stream_upload / orbx_stream_download (NULL on failure) */
void* orbx_host_alloc(size_t bytes);
void orbx_host_free(void* p);
/* raw cudaStream_t of the context (for callers that record their own events) */
void* orbx_get_stream(orbx_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
