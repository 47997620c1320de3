/* include/orbx.h -- C-ABI of liborbx.so: the B200-native (sm_100a) ORB front-end.
 *
 * This is the drop-in boundary for the reference's feature hot path.  Plain C types, caller-allocated
 * outputs, int status codes; no OpenCV / torch types cross it.  The C++ adapter classes in
 * orbslam2commentedbyxcm_b200/host/ (ORB_SLAM2::ORBextractor / ORBmatcher with the reference's public surface) are
 * thin marshalling layers over these entry points; INTEGRATION.md shows the reference-side binding.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference tree).
 * There is no CPU fallback: every compute entry point fails with ORBX_ERR_NO_DEVICE / ORBX_ERR_CUDA when
 * no sm_100 device is usable.
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_ERR_BAD_ARG (-1)   /* null pointer, non-positive size, nlevels > ORBX_MAX_LEVELS, ... */
#define ORBX_ERR_TOO_SMALL (-2) /* image too small for the pyramid / 30-px cell grid at some level */
#define ORBX_ERR_CUDA (-3)      /* a CUDA call failed; orbx_last_error() has the text */
#define ORBX_ERR_CAPACITY (-4)  /* a caller-provided output buffer is too small */
#define ORBX_ERR_NO_DEVICE (-5) /* no CUDA device */

#define ORBX_MAX_LEVELS 16
#define ORBX_EDGE 19       /* EDGE_THRESHOLD, src/ORBextractor.cc:46 */
#define ORBX_DESC_BYTES 32 /* 256-bit rBRIEF */

typedef struct orbx_handle orbx_handle;

/* ORBextractor constructor arguments, include/ORBextractor.h:93, src/ORBextractor.cc:438-443 */
typedef struct orbx_config {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int ini_th_fast;
    int min_th_fast;
} orbx_config;

/* cv::KeyPoint as the reference fills it (src/ORBextractor.cc:1146-1155, 1613-1622); class_id is always -1 */
typedef struct orbx_keypoint {
    float x, y;     /* pt, level-0 pixel frame (level coords * mvScaleFactor[octave]) */
    float size;     /* (int)(31 * mvScaleFactor[octave]) */
    float angle;    /* degrees, [0, 360], cv::fastAtan2 semantics */
    float response; /* FAST score */
    int octave;
} orbx_keypoint;

/* Text of the last error on this thread (never NULL). */
const char* orbx_last_error(void);

/* Number of CUDA devices visible, or a negative ORBX_ERR_* code. */
int orbx_device_count(void);

/* ---- extractor ------------------------------------------------------------------------------------ */

/* Replaces ORBextractor::ORBextractor (src/ORBextractor.cc:438-550): computes the scale tables and per-level
 * quotas, creates the device context (stream, buffers are sized lazily per image shape) on `device`. */
int orbx_create(const orbx_config* cfg, int device, orbx_handle** out);
void orbx_destroy(orbx_handle* h);

/* Replaces GetLevels / GetScaleFactor(s) / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares (include/ORBextractor.h:119-159) and exposes mnFeaturesPerLevel.
 * Any pointer may be NULL; arrays must hold nlevels entries. */
int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                    int* features_per_level);

/* Upper bound on keypoints per frame for this configuration and image shape (rows x cols): the quad-tree can
 * return up to quota+3 (or 4*nIni) nodes per level.  Size kps/desc buffers with it. */
int orbx_max_keypoints(orbx_handle* h, int rows, int cols);

/* Replaces ORBextractor::operator()(image, mask, keypoints, descriptors) (src/ORBextractor.cc:1513-1629) for
 * n frames of identical shape.  HOST buffers in, HOST buffers out; H2D, all kernels, D2H and one stream
 * synchronise happen inside.
 *   imgs          n frames of CV_8UC1, frame i at imgs + i*frame_stride, rows `step` bytes apart
 *   kps, desc     n * capacity entries / n * capacity * 32 bytes; frame i at offset i*capacity
 *   counts        n entries: keypoints found per frame (0 for an "empty image" is not produced here:
 *                 rows/cols <= 0 or imgs == NULL is ORBX_ERR_BAD_ARG; the adapter maps it to a silent return)
 * Keypoints are ordered level 0..nlevels-1 and, inside a level, in the reference's quad-tree list order. */
int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs, int n, int rows, int cols, size_t step,
                       size_t frame_stride, orbx_keypoint* kps, uint8_t* desc, int capacity, int* counts);

/* Page-locked host memory for images and results.  When kps / desc / counts of orbx_extract_batch live in such memory
 * and capacity == orbx_max_keypoints(), results are DMA'd straight into them (no staging copy on the host); note that
 * then all `capacity` entries of a frame are transferred and only the first counts[i] are meaningful. */
void* orbx_alloc_pinned(size_t bytes);
void orbx_free_pinned(void* p);

/* Single frame convenience (n = 1). */
int orbx_extract(orbx_handle* h, const uint8_t* img, int rows, int cols, size_t step, orbx_keypoint* kps,
                 uint8_t* desc, int capacity, int* count);

/* Same computation with DEVICE-resident inputs and outputs, asynchronous on the handle's stream (no copy, no
 * host synchronisation): d_imgs are tight or pitched device frames, d_kps/d_desc/d_counts device buffers laid
 * out as in orbx_extract_batch.  Use orbx_sync() (or your own event on orbx_stream()) before reading results.
 * capacity below orbx_max_keypoints() is accepted; a frame with more keypoints than capacity is truncated to it AND
 * reported: the next orbx_sync() returns ORBX_ERR_CAPACITY. */
int orbx_extract_batch_device(orbx_handle* h, const uint8_t* d_imgs, int n, int rows, int cols, size_t step,
                              size_t frame_stride, orbx_keypoint* d_kps, uint8_t* d_desc, int capacity,
                              int* d_counts);
int orbx_sync(orbx_handle* h);
/* cudaStream_t of the handle, as void*. */
void* orbx_stream(orbx_handle* h);

/* Replaces the public member mvImagePyramid (include/ORBextractor.h:162) of the LAST extracted batch:
 * copies level `level` of frame `frame` INCLUDING its 19-px REFLECT_101 frame into dst
 * ((rows_l + 38) rows of (cols_l + 38) bytes, dst_step bytes apart), so that an ROI at (19,19) of dst behaves
 * like the reference's Mat (adjustROI by <= 19 stays valid, src/Frame.cc:782-831).
 * Residency: `frame` indexes the last batch.  After orbx_extract_batch_device and after host batches of up to 48
 * frames every frame of the batch is resident.  A larger host batch is pipelined through two arena slots and only
 * the frames of its last two chunks (at most the last 64 frames) stay resident; asking for an earlier frame -- here,
 * in orbm_stereo_match (which reads the pyramids through the handle) or in the orbx_debug_* getters -- returns
 * ORBX_ERR_BAD_ARG ("no longer resident"), never another frame's data. */
int orbx_get_level_size(orbx_handle* h, int level, int* rows_l, int* cols_l);
int orbx_get_pyramid_level(orbx_handle* h, int frame, int level, uint8_t* dst, size_t dst_step);
/* Every padded level of one frame in ONE call (ORBextractor::mvImagePyramid, include/ORBextractor.h:162; consumer
 * Frame::ComputeStereoMatches, src/Frame.cc:782-831): level l is written at dst + level_offsets[l] with row step
 * level_steps[l] >= cols_l + 38, (rows_l + 38) rows.  The copies are enqueued back to back and waited for once; give it
 * page-locked memory (orbx_alloc_pinned) and the copies run at link speed instead of through the driver's staging buffer. */
int orbx_get_pyramid_all(orbx_handle* h, int frame, uint8_t* dst, const size_t* level_offsets, const size_t* level_steps);

/* Per-stage device time of the last batch in milliseconds, measured with CUDA events when profiling is on
 * (profiling serialises the stages; leave it off for throughput).  Stage order: ORBX_STAGE_*. */
#define ORBX_STAGE_H2D 0
#define ORBX_STAGE_PYRAMID 1
#define ORBX_STAGE_FAST 2
#define ORBX_STAGE_OCTREE 3 /* quad-tree distribution (IC_Angle is part of ORBX_STAGE_DESC: k_describe) */
#define ORBX_STAGE_BLUR 4
#define ORBX_STAGE_DESC 5
#define ORBX_STAGE_D2H 6
#define ORBX_NUM_STAGES 7
int orbx_set_profiling(orbx_handle* h, int on);
int orbx_stage_times(orbx_handle* h, float* ms /* ORBX_NUM_STAGES */);
/* Number of kernel launches issued by the last extract call (for bench.py's gpu_launches). */
int orbx_last_launch_count(orbx_handle* h);
/* Use a captured CUDA graph per (shape, batch) for the kernel sequence (default on). */
int orbx_set_use_graph(orbx_handle* h, int on);

/* Stage introspection for parity tests (last batch): FAST candidates of one level as (x, y, response) triples
 * in level pixel coordinates, unordered; the blurred interior of one level (rows_l x cols_l, tight). */
int orbx_debug_get_candidates(orbx_handle* h, int frame, int level, int* xyr, int cap, int* n);
int orbx_debug_get_blurred(orbx_handle* h, int frame, int level, uint8_t* dst);

/* Domain sweeps of the two transcendental device functions behind IC_Angle and the steered descriptor (parity tests):
 * the device cosf/sinf restatement of glibc (src/ORBextractor.cc:109,123-125: angle = deg * factorPI) over the float bit
 * patterns first_bits .. first_bits + count - 1 of `deg`, and the device cv::fastAtan2 (src/ORBextractor.cc:105) over
 * generated integer moment pairs (mode 0: pair i = (i / 10001 - 5000, i % 10001 - 5000), the dense grid [-5000, 5000]^2;
 * mode k >= 1: two signed 22-bit values from splitmix64(i + k * 0x9E3779B97F4A7C15)).  The *_checksum forms fold every
 * `chunk` consecutive values into sum_i bits(value_i) * (2 i + 1) mod 2^64 (i = index inside the chunk), one word per
 * chunk, so all 1 135 869 953 angles can be compared with the CPU oracle without moving them; the plain forms return
 * the values (n <= 2^24 is sensible). */
int orbx_debug_sincosf_checksum(int device, uint32_t first_bits, uint64_t count, uint64_t chunk, uint64_t* sum_sin,
                                uint64_t* sum_cos);
int orbx_debug_sincosf(int device, uint32_t first_bits, int n, float* out_sin, float* out_cos);
int orbx_debug_fast_atan2_checksum(int device, int mode, uint64_t first, uint64_t count, uint64_t chunk, uint64_t* sum);
int orbx_debug_fast_atan2(int device, int mode, uint64_t first, int n, float* out);

/* ---- matcher -------------------------------------------------------------------------------------- */

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1983-2003): 256-bit Hamming distance of two 32-byte rows.
 * Host-side (a single pair never justifies a launch). */
int orbm_descriptor_distance(const uint8_t* a, const uint8_t* b);

typedef struct orbm_context orbm_context;
int orbm_create(int device, orbm_context** out);
void orbm_destroy(orbm_context* m);
void* orbm_stream(orbm_context* m);
int orbm_sync(orbm_context* m);

/* Brute-force best / second-best Hamming search (north_star item 6, config 5): for every query row the smallest
 * distance over db rows [0, ndb), the lowest index attaining it (reported + idx_offset) and the second
 * smallest distance of the multiset -- the result of the reference's sequential
 * `if(d<best){second=best;best=d;idx=j} else if(d<second) second=d` scan (src/ORBmatcher.cc:134-151) over the
 * whole database.  Empty db: best = second = 256, idx = -1.  mem_kind 0: all pointers are HOST (copies inside,
 * synchronous); 1: all pointers are DEVICE (asynchronous on orbm_stream()). */
int orbm_bruteforce(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, int idx_offset,
                    int32_t* best, int32_t* best_idx, int32_t* second, int mem_kind);

/* orbm_bruteforce for callers that are not allowed to allocate device memory.  (No counterpart in the reference: it
 * exists for the integration, not for parity -- the result is orbm_bruteforce's, bit for bit.)
 *
 * Who needs it: a SLAM front end that was handed a fixed device-memory budget and runs next to another owner of the GPU
 * (a planner, an inference engine, a caching allocator that has already reserved the rest of the memory), where a
 * cudaMalloc in the middle of tracking can fail or stall; and hard-real-time loops, because cudaMalloc / cudaFree
 * serialise against all work on the device and have no latency bound.  orbm_bruteforce grows a scratch buffer with
 * cudaMalloc whenever a larger search arrives; this entry point never does.  It needs no orbm_context, creates no
 * stream or event, and calls no cudaMalloc, cudaFree or any other allocating routine.
 *
 * The scratch (one best / index / second triple per database chunk and query) is a fixed array inside liborbx.so:
 * ORBM_NOALLOC_SCRATCH_BYTES = 12 MiB of device memory on EVERY GPU that loads the library, held from load to process
 * exit, whether or not this function is called.  That bounds the search: nq <= ORBM_NOALLOC_MAX_NQ and
 * ndb <= ORBM_NOALLOC_MAX_NDB, anything larger is refused with ORBX_ERR_CAPACITY (orbx_last_error() says why) and
 * nothing is launched.
 *
 *   device       CUDA device ordinal (< 64); made current by the call
 *   stream       a cudaStream_t of that device, or NULL for the default stream; both kernels are enqueued on it and the
 *                call returns without waiting -- synchronise the stream before reading the outputs
 *   q, db, best, best_idx, second   DEVICE pointers, laid out as for orbm_bruteforce with mem_kind = 1
 *
 * The first call on each device looks up the scratch array's device address and the SM count (a few microseconds of
 * host work, no allocation) and caches them; make that first call during set-up, before any stream capture begins.
 * All calls on one device share the one array.  Nothing in it has to be cleared between calls -- each call overwrites
 * every slot it later reads -- but two calls whose kernels can overlap on the GPU (different streams, or different host
 * threads on different streams) race on it and return garbage: keep calls on one device in one stream, or order them
 * with events. */
#define ORBM_NOALLOC_MAX_NQ 4096                /* queries per call (two frames' worth of features) */
#define ORBM_NOALLOC_MAX_NDB (1 << 24)          /* database rows per call (512 MiB of descriptors) */
#define ORBM_NOALLOC_SCRATCH_SLOTS (1 << 20)    /* (chunk, query) triples the static array holds */
#define ORBM_NOALLOC_SCRATCH_BYTES (3 * 4 * ORBM_NOALLOC_SCRATCH_SLOTS)
int orbm_bruteforce_noalloc(int device, void* stream, const uint8_t* q, int nq, const uint8_t* db, int ndb,
                            int idx_offset, int32_t* best, int32_t* best_idx, int32_t* second);

/* Deterministic merge of `nparts` partial results (e.g. one per database shard, parts ordered by ascending
 * index range): equals the sequential scan over the concatenated database.  DEVICE pointers,
 * best/best_idx/second laid out [part][nq].  Used after the all-gather in the sharded search. */
int orbm_merge_partials(orbm_context* m, const int32_t* best, const int32_t* best_idx, const int32_t* second,
                        int nparts, int nq, int32_t* out_best, int32_t* out_idx, int32_t* out_second);

/* ---- Brute force on the tensor cores (sm_100a: tcgen05.mma kind::i8, TMEM accumulators, TMA operand tiles).  The same
 * result as orbm_bruteforce, bit for bit: with descriptors expanded to +-1 bytes the dot product is 256 - 2 x Hamming
 * distance (src/ORBmatcher.cc:1983-2003), an exact integer, so query tile x database tile distances are an int8 GEMM and
 * the best / second selection runs as its epilogue.  orbm_db_create expands a database (mem_kind 0 = host pointer, 1 =
 * device pointer; 256 bytes per row of device memory) once; searches against it are orbm_bruteforce_db (mem_kind as in
 * orbm_bruteforce) and, for a shard of a distributed map, orbm_bruteforce_sharded_db (DEVICE pointers). */
typedef struct orbm_db orbm_db;
int orbm_db_create(orbm_context* m, const uint8_t* db, int ndb, int mem_kind, orbm_db** out);
void orbm_db_destroy(orbm_db* db);
int orbm_bruteforce_db(orbm_context* m, const orbm_db* db, const uint8_t* q, int nq, int idx_offset, int32_t* best,
                       int32_t* best_idx, int32_t* second, int mem_kind);

/* ---- Sharded brute-force search with a fused merge + exchange over peer memory (north_star: "local-map matching shards
 * ... by query set across 2, 4 and 8 GPUs"; replaces the all-gather + orbm_merge_partials pair).
 * One process (or context) per GPU.  Each rank creates an exchange buffer, publishes its CUDA IPC handle through any
 * side channel (torch.distributed / MPI / a pipe), and connects to the peers' handles once.  After that a search is
 * asynchronous device work only: the rank's shard is searched, one kernel merges the chunk partials, stores the
 * per-query (best, index, second) triple into every peer's buffer with NVLink P2P stores, waits for the peers' triples
 * in its own buffer and merges them (same deterministic rule as orbm_merge_partials, so the result equals the
 * unsharded sequential scan on every rank).  All ranks must call orbm_bruteforce_sharded the same number of times with
 * the same queries; nq may differ from call to call (any value up to nq_max): the slot freshness tag wraps every 16384
 * calls, and each call rewrites a rotating window of unused slots so that an old tag can never pass for the current one.
 * Pointers of orbm_bruteforce_sharded are DEVICE pointers; call orbm_sync() before reading. */
#define ORBM_IPC_HANDLE_BYTES 64
#define ORBM_MAX_WORLD 16
int orbm_exchange_create(orbm_context* m, int world, int rank, int nq_max, uint8_t* handle_out /* 64 B or NULL */,
                         void** local_ptr_out /* or NULL */);
/* handles: world x ORBM_IPC_HANDLE_BYTES, rank order (the entry of this rank is ignored) */
int orbm_exchange_connect(orbm_context* m, const uint8_t* handles);
/* same-process variant (several contexts / GPUs in one process): raw device pointers from orbm_exchange_create */
int orbm_exchange_connect_local(orbm_context* m, void* const* peer_ptrs);
void orbm_exchange_destroy(orbm_context* m);
int orbm_bruteforce_sharded(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db_shard, int ndb_shard, int idx_offset,
                            int32_t* best, int32_t* best_idx, int32_t* second);
/* The same in two steps, for callers that overlap other work with the exchange (and for single-GPU protocol tests):
 * phases = 1 searches the shard and publishes the triples to the peers, phases = 2 waits for the peers' triples of the
 * latest published search and writes the merged result, phases = 3 is orbm_bruteforce_sharded. */
int orbm_bruteforce_sharded_phases(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db_shard, int ndb_shard,
                                   int idx_offset, int32_t* best, int32_t* best_idx, int32_t* second, int phases);
/* orbm_bruteforce_sharded with the shard prepared by orbm_db_create (tensor-core search of the shard, same exchange). */
int orbm_bruteforce_sharded_db(orbm_context* m, const orbm_db* db_shard, const uint8_t* q, int nq, int idx_offset, int32_t* best,
                               int32_t* best_idx, int32_t* second);
/* 1 if a sharded search gave up waiting for a peer (the results of that call are invalid) */
int orbm_exchange_timed_out(orbm_context* m);
/* test hook: copies the first nwords 64-bit slot words (tag << 50 | best << 41 | second << 32 | index; layout
 * [parity][source rank][nq_max]) of this rank's exchange buffer to the host, and the number of searches published so far */
int orbm_debug_exchange_snapshot(orbm_context* m, uint64_t* words, size_t nwords, uint32_t* epoch);

/* Frame-side inputs of the windowed searches: Frame::mvKeysUn, mvuRight, mDescriptors, mvpMapPoints occupancy,
 * mGrid (src/Frame.cc:351-370) flattened as CSR with cell (ix, iy) at ix*48+iy, and the grid constants. */
typedef struct orbm_frame_view {
    int n;
    const float* x;            /* mvKeysUn[i].pt.x */
    const float* y;
    const int32_t* octave;
    const float* u_right;      /* mvuRight, may be NULL (monocular: all -1) */
    const uint8_t* desc;       /* n x 32 */
    const uint8_t* occupied;   /* 1 iff mvpMapPoints[i] && mvpMapPoints[i]->Observations() > 0 on entry */
    const int32_t* cell_start; /* 64*48 + 1 */
    const int32_t* cell_items;
    float min_x, min_y, grid_w_inv, grid_h_inv; /* mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv */
    const float* scale_factors; /* mvScaleFactors */
    int nlevels;
} orbm_frame_view;

/* MapPoint-side inputs: the tracking scratch fields written by Frame::isInFrustum (include/MapPoint.h:149-156). */
typedef struct orbm_mappoint_view {
    int n;
    const float* proj_x;     /* mTrackProjX */
    const float* proj_y;     /* mTrackProjY */
    const float* proj_xr;    /* mTrackProjXR */
    const int32_t* level;    /* mnTrackScaleLevel */
    const float* view_cos;   /* mTrackViewCos */
    const uint8_t* valid;    /* mbTrackInView && !isBad() */
    const uint8_t* obs_pos;  /* Observations() > 0 (a claim by such a point blocks later points) */
    const uint8_t* desc;     /* GetDescriptor(), n x 32 */
} orbm_mappoint_view;

/* Replaces ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, float th) (src/ORBmatcher.cc:61-173,
 * RadiusByViewingCos :176-183, Frame::GetFeaturesInArea src/Frame.cc:488-548) including its order-dependent
 * claims.  HOST pointers.  out_match[i] = index of the map point the call would store in F.mvpMapPoints[i], or -1
 * when it leaves the slot untouched.  *nmatches = the function's return value. */
int orbm_search_by_projection(orbm_context* m, const orbm_frame_view* f, const orbm_mappoint_view* p, float th,
                              float nnratio, int32_t* out_match, int* nmatches);

/* Source features of a windowed search AFTER the caller's projection step (the adapter keeps the reference's own
 * cv::Mat arithmetic for Rcw*x3Dw+tcw, fx*xc*invzc+cx, PredictScale, ...). */
typedef struct orbm_window_query_view {
    int n;
    const float* x;              /* projected u */
    const float* y;              /* projected v */
    const float* radius;         /* window half-size handed to Frame::GetFeaturesInArea */
    const int32_t* min_level;    /* minLevel / maxLevel arguments of GetFeaturesInArea (-1 = open) */
    const int32_t* max_level;
    const float* xr;             /* predicted right coordinate u - mbf*invz for the stereo gate; NULL = no gate */
    const float* angle;          /* source keypoint angle (rotation histogram); NULL when check_orientation == 0 */
    const uint8_t* valid;        /* 0 = skipped by the reference before the search (no map point, outlier, behind camera, ...) */
    const uint8_t* claims_block; /* 1 = a keypoint claimed by this query is skipped by later queries */
    const uint8_t* desc;         /* n x 32 */
} orbm_window_query_view;

/* Windowed search with the reference's order-dependent claims, shared by
 *   ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)          src/ORBmatcher.cc:1620-1789
 *   ORBmatcher::SearchByProjection(Frame& Cur, KeyFrame*, sAlreadyFound, th, ORBdist)  src/ORBmatcher.cc:1792-1924
 *   ORBmatcher::SearchByProjection(KeyFrame*, cv::Mat Scw, vpPoints, vpMatched, th)   src/ORBmatcher.cc:398-520
 *     (loop closing; KeyFrame::GetFeaturesInArea is the same window walk: occupied[i] = (vpMatched[i] != NULL),
 *      claims_block = 1, max_dist = TH_LOW, ratio_mode 0, check_orientation 0, levels [pred-1, pred], xr = NULL)
 * (max_dist = TH_HIGH resp. ORBdist, ratio_mode 0, check_orientation = mbCheckOrientation; for the KeyFrame overload
 * pass occupied[i] = (mvpMapPoints[i] != NULL), claims_block = 1 and xr = NULL) and, with ratio_mode 1, by
 * orbm_search_by_projection.  f_angle = Cur.mvKeysUn[i].angle.  HOST pointers.
 * out_match[i]: -1 = slot untouched, -2 = set to NULL by the rotation-histogram filter, >= 0 = query index stored. */
int orbm_search_window(orbm_context* m, const orbm_frame_view* f, const float* f_angle, const orbm_window_query_view* q,
                       int max_dist, int ratio_mode, float nnratio, int check_orientation, int32_t* out_match,
                       int* nmatches);

/* Replaces the coarse Hamming search of Frame::ComputeStereoMatches (src/Frame.cc:684-765).  HOST pointers.
 * best_r[i] = right keypoint index for left keypoint i or -1 (distance >= (TH_HIGH+TH_LOW)/2 or no candidate);
 * best_dist[i] = the best distance found (TH_HIGH when none). */
int orbm_stereo_rows(orbm_context* m, int nl, const float* lx, const float* ly, const int32_t* loct,
                     const uint8_t* ldesc, int nr, const float* rx, const float* ry, const int32_t* roct,
                     const uint8_t* rdesc, const float* scale_factors, int nlevels, int n_rows, float min_d,
                     float max_d, int32_t* best_r, int32_t* best_dist);

/* Replaces the WHOLE of Frame::ComputeStereoMatches (src/Frame.cc:673-885): coarse row search, SAD refinement on the
 * pyramid levels (11x11 window, 11 shifts, sub-pixel parabola fit, :767-866) and removal of matches whose SAD is >=
 * 2.1 x median (:868-884).  In this reference that block sits INSIDE the per-keypoint loop, so a match is tested against
 * the running median after its own insertion and after every later one; that is the default (bit-identical to the
 * reference's Frame constructor, tests/test_oracle_matcher.py).  orbm_set_stereo_outlier_mode(m, 1) selects the upstream
 * ORB-SLAM2 placement instead (one test against the final median after the loop).  The level images are read straight
 * from the device arenas of the two extractor handles (frame indices refer to their last extracted batch), so
 * mvImagePyramid never has to be downloaded.  Keypoint arrays are HOST pointers.  mb = baseline in metres (mbf / fx):
 * minZ = mb, maxD = mbf / minZ.  u_right / depth: nl entries, -1 where unmatched (mvuRight, mvDepth). */
int orbm_stereo_match(orbm_context* m, orbx_handle* ex_left, int frame_left, orbx_handle* ex_right, int frame_right,
                      int nl, const float* lx, const float* ly, const int32_t* loct, const uint8_t* ldesc, int nr,
                      const float* rx, const float* ry, const int32_t* roct, const uint8_t* rdesc, float mbf, float mb,
                      float* u_right, float* depth, int* n_matched);
int orbm_set_stereo_outlier_mode(orbm_context* m, int mode /* 0 = as this reference (default), 1 = upstream */);

/* DBoW2::FeatureVector (std::map<NodeId, vector<unsigned>>) flattened as CSR with ascending node ids. */
typedef struct orbm_featvec_view {
    int n_nodes;
    const int32_t* node_id;
    const int32_t* node_start; /* n_nodes + 1 */
    const int32_t* node_feat;
} orbm_featvec_view;

/* Replaces ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (src/ORBmatcher.cc:228-392 incl. the
 * rotation-histogram filter, ComputeThreeMaxima :1935-1977).  HOST pointers.  kf_has_mp[i] = KF feature i has a
 * non-bad MapPoint.  out_match[f] = KF feature index matched to frame feature f, or -1. */
int orbm_search_by_bow(orbm_context* m, int nkf, const uint8_t* kf_desc, const float* kf_angle,
                       const uint8_t* kf_has_mp, const orbm_featvec_view* kf_fv, int nf, const uint8_t* f_desc,
                       const float* f_angle, const orbm_featvec_view* f_fv, float nnratio, int check_orientation,
                       int32_t* out_match, int* nmatches);

/* Replaces ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12)
 * (src/ORBmatcher.cc:696-850; loop closing).  has_mp1/2[i] = feature i of that key frame has a non-bad MapPoint.
 * out_match12[i1] = matched feature index of KF2 or -1 (the caller stores vpMapPoints2[that index]).  HOST pointers. */
int orbm_search_by_bow_kf(orbm_context* m, int n1, const uint8_t* desc1, const float* angle1, const uint8_t* has_mp1,
                          const orbm_featvec_view* fv1, int n2, const uint8_t* desc2, const float* angle2,
                          const uint8_t* has_mp2, const orbm_featvec_view* fv2, float nnratio, int check_orientation,
                          int32_t* out_match12, int* nmatches);

/* Claim-free windowed search ("next" row): the candidate loop of
 *   ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th)          src/ORBmatcher.cc:1067-1222  gate = 1
 *   ORBmatcher::Fuse(KeyFrame*, cv::Mat Scw, vpPoints, th, vpReplace)  src/ORBmatcher.cc:1225-1357  gate = 0
 * after the caller's projection (x, y = u, v; xr = u - bf*invz; radius = th * mvScaleFactors[nPredictedLevel];
 * min_level / max_level = nPredictedLevel-1 / nPredictedLevel; valid = 0 for the points the reference skips before the
 * search).  f describes the KeyFrame (mvKeysUn, mvuRight, mDescriptors, mGrid as in orbm_frame_view; `occupied`,
 * q->angle and q->claims_block are not read).  Candidates are KeyFrame::GetFeaturesInArea(u, v, radius)
 * (src/KeyFrame.cc:623-658); gate 1 adds Fuse's reprojection test e2 * inv_level_sigma2[octave] > 7.8 (stereo keypoint,
 * e2 includes (xr - mvuRight)^2) resp. > 5.99 (monocular).  max_dist = TH_LOW (50) for both Fuse overloads.
 * best_idx[q] = KeyFrame feature (first minimum in visiting order) or -1 (no candidate / distance > max_dist);
 * best_dist (may be NULL) = the smallest distance seen (256 when none).  The map mutations (Replace / AddObservation /
 * AddMapPoint / vpReplacePoint, :1198-1217 and :1333-1350) stay with the caller.  HOST pointers. */
int orbm_search_best(orbm_context* m, const orbm_frame_view* f, const orbm_window_query_view* q, int gate,
                     const float* inv_level_sigma2, int max_dist, int32_t* best_idx, int32_t* best_dist);

/* Replaces the two projection passes and the agreement check of ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1361-1602).
 * q1_in_2: one entry per KF1 feature = its map point transformed into camera 2 and projected (valid = has a good map
 * point, !vbAlreadyMatched1, in image, inside the scale-invariance distances), searched in kf2; q2_in_1 likewise, one
 * entry per KF2 feature, searched in kf1; radius = th * mvScaleFactors[nPredictedLevel], levels [pred-1, pred],
 * acceptance bestDist <= TH_HIGH.  out_match12[i1] = idx2 when both directions agree, else -1 (the caller stores
 * vpMapPoints2[idx2] into vpMatches12[i1]); *nfound = the return value.  HOST pointers. */
int orbm_search_by_sim3(orbm_context* m, const orbm_frame_view* kf1, const orbm_window_query_view* q2_in_1,
                        const orbm_frame_view* kf2, const orbm_window_query_view* q1_in_2, int32_t* out_match12,
                        int* nfound);

/* Replaces ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
 * (src/ORBmatcher.cc:539-683) including vMatchedDistance / vnMatches21 (a later, strictly better match steals an F2
 * feature; the stolen match still counts in the rotation histogram).  F1 side: octave1 = mvKeysUn[i].octave (only level
 * 0 searches), angle1, desc1 = mDescriptors, prev_x / prev_y = vbPrevMatched.  f2 = F2 as orbm_frame_view (`occupied`
 * and `u_right` not read), f2_angle = F2.mvKeysUn[i].angle.  out_match12[i1] = vnMatches12; the caller updates
 * vbPrevMatched[i1] = F2.mvKeysUn[out_match12[i1]].pt for the matched ones (:677-680).  HOST pointers. */
int orbm_search_for_initialization(orbm_context* m, int n1, const int32_t* octave1, const float* angle1,
                                   const uint8_t* desc1, const float* prev_x, const float* prev_y,
                                   const orbm_frame_view* f2, const float* f2_angle, int window_size, float nnratio,
                                   int check_orientation, int32_t* out_match12, int* nmatches);

/* One key frame of SearchForTriangulation: mvKeysUn (pt, octave, angle), mvuRight (>= 0 = stereo), mDescriptors,
 * has_mp[i] = (GetMapPoint(i) != NULL), mFeatVec. */
typedef struct orbm_kf_view {
    int n;
    const float* x;
    const float* y;
    const int32_t* octave;
    const float* angle;
    const float* u_right;
    const uint8_t* desc;
    const uint8_t* has_mp;
    const orbm_featvec_view* fv;
} orbm_kf_view;

/* Replaces ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)
 * (src/ORBmatcher.cc:850-1056) with CheckDistEpipolarLine (:186-213).  F12 = 3x3 row-major; (ex, ey) = projection of
 * KF1's camera centre into KF2 (:858-865, computed by the caller); scale_factors2 / level_sigma2_2 = pKF2->mvScaleFactors
 * / mvLevelSigma2.  out_match12[idx1] = idx2 or -1: vMatchedPairs is the list of (i, out_match12[i]) with
 * out_match12[i] >= 0 in ascending i.  HOST pointers. */
int orbm_search_for_triangulation(orbm_context* m, const orbm_kf_view* kf1, const orbm_kf_view* kf2, const float* F12,
                                  float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                  int nlevels2, int only_stereo, int check_orientation, int32_t* out_match12,
                                  int* nmatches);

/* ---- DBoW2 vocabulary descent ("next" row) ------------------------------------ */

/* A vocabulary tree on the device.  Flat form of TemplatedVocabulary::m_nodes (Thirdparty/DBoW2/DBoW2/
 * TemplatedVocabulary.h:1338-1420 builds it from ORBvoc.txt): node 0 is the root, the children of node i are
 * child_ids[child_start[i] .. child_start[i+1]) in the order of m_nodes[i].children, node_desc is n_nodes x 32 bytes,
 * word_id[i] >= 0 and weight[i] are meaningful for leaves.  L = tree depth (m_L). */
typedef struct orbm_vocab orbm_vocab;
int orbm_vocab_create(orbm_context* m, int n_nodes, int L, const int32_t* child_start, const int32_t* child_ids,
                      const uint8_t* node_desc, const int32_t* word_id, const double* weight, orbm_vocab** out);
void orbm_vocab_destroy(orbm_vocab* v);

/* Replaces the per-feature TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup) (:1218-1259) that
 * Frame::ComputeBoW / KeyFrame::ComputeBoW run for every descriptor: word id, weight and the node id `levelsup` levels
 * above the leaves (0 = root when L - levelsup <= 0).  HOST pointers.  The caller folds the triples into
 * DBoW2::BowVector / FeatureVector in feature order exactly as :1127-1192 does (addWeight / addFeature, L1 normalise). */
int orbm_bow_transform(orbm_context* m, const orbm_vocab* v, const uint8_t* desc, int n, int levelsup, int32_t* word_id,
                       int32_t* node_id, double* weight);

/* Replaces the arithmetic of MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:295-360) for a batch of map
 * points ("next" row): descriptor sets in CSR form (rows set_start[s] .. set_start[s+1] of desc);
 * best_idx[s] = index inside set s of the descriptor with the least median Hamming distance to the set (first minimum;
 * -1 for an empty set).  HOST pointers. */
int orbm_distinctive_descriptors(orbm_context* m, int n_sets, const int32_t* set_start, const uint8_t* desc,
                                 int32_t* best_idx);

/* Frame preparation between extraction and matching ("next" row).  mem_kind 0 = HOST pointers
 * (synchronous), 1 = DEVICE pointers (asynchronous on the context's stream), as in orbm_bruteforce.
 *
 * orbm_undistort_keypoints replaces Frame::UndistortKeyPoints (src/Frame.cc:586-628): copies the coordinates when
 * dist[0] == 0 (:589-592), else cv::undistortPoints(pts, K, dist, R = I, P = K) -- 5 iterations in double, result
 * narrowed to float; `dist` = (k1, k2, p1, p2[, k3[, k4, k5, k6[, s1..s4]]]) is always a HOST array (mDistCoef).
 *
 * orbm_assign_grid replaces Frame::AssignFeaturesToGrid + PosInGrid (:351-370, :558-567): the 64 x 48 mGrid as CSR
 * (cell (ix, iy) at ix * 48 + iy; cell_start has 64*48+1 entries, cell_items n), indices ascending inside a cell --
 * the layout orbm_frame_view takes. */
int orbm_undistort_keypoints(orbm_context* m, const float* x, const float* y, int n, float fx, float fy, float cx, float cy,
                             const float* dist, int ndist, float* xu, float* yu, int mem_kind);
int orbm_assign_grid(orbm_context* m, const float* x, const float* y, int n, float min_x, float min_y, float grid_w_inv,
                     float grid_h_inv, int32_t* cell_start, int32_t* cell_items, int mem_kind);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
