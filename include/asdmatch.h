/*
 * asdmatch.h -- C ABI of libasdmatch_cuda.so: B200-native (sm_100a) ASD descriptor matching.
 *
 * Drop-in boundary for the one data-parallel hot path of mataiyuan/ASD-SLAM: the descriptor loops of
 * ORB_SLAM2::ORBmatcher (reference src/vslam/include/ORBmatcher.h:48-90, src/vslam/src/ORBmatcher.cc).
 * The reference has no FFI for this path: ORBmatcher is a C++ class inside
 * libvslam.so.  The C++ adaptor asd-slam_b200/host/ORBmatcher.{h,cc} keeps that class's method names
 * and argument meaning and marshals Frame / KeyFrame / MapPoint state into the flat calls below;
 * INTEGRATION.md shows the binding a maintainer adds to the reference.
 *
 * Conventions
 *  - plain pointers and sizes only; no torch / OpenCV types; MapPoint* never crosses the boundary
 *    (the adaptor maps indices back to pointers);
 *  - every function returns ASDM_OK (0) or a negative error code and never throws;
 *    asdm_last_error() gives the message of the last failure on that context;
 *  - host buffers are caller-owned; device buffers are library-owned (asdm_set, asdm_job) and come from the
 *    device's stream-ordered memory pool, so creating / releasing sets and jobs never synchronises the device;
 *  - a context owns its CUDA streams; calls on one context are serialised by an internal mutex. Threads that must not
 *    wait for each other take a context each (asdm_fork) and share the descriptor sets;
 *  - there is NO CPU fallback: without a CUDA device asdm_create fails with ASDM_ERR_CUDA.
 *
 * Result contract: index lists are bit-identical to the reference's CPU matcher on the same inputs;
 * returned distances are the reference's own fp32 values (sequential index-ascending accumulation of
 * (a-b)*(a-b), no FMA: ORBmatcher.cc:1637-1644).
 */
#ifndef ASDMATCH_H
#define ASDMATCH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASDM_DESC_DIM 128     /* ASD descriptors: N x 128 CV_32FC1 (ORBextractor.cc:1209) */
#define ASDM_HISTO_LENGTH 30  /* ORBmatcher.cc:39 */
#define ASDM_TH_LOW 0.5f      /* ORBmatcher.cc:38 */
#define ASDM_TH_HIGH 1.5f     /* ORBmatcher.cc:37 */

enum {
  ASDM_OK = 0,
  ASDM_ERR_INVALID = -1,     /* bad argument */
  ASDM_ERR_CUDA = -2,        /* CUDA runtime / driver failure (including: no device) */
  ASDM_ERR_OOM = -3,
  ASDM_ERR_INTERNAL = -4
};

typedef struct asdm_ctx asdm_ctx;
typedef struct asdm_set asdm_set; /* device-resident descriptors + keypoints + grid of one Frame / KeyFrame */
typedef struct asdm_job asdm_job; /* a prepared batched search whose inputs are resident in HBM */

/* cv::KeyPoint fields the matcher reads (mvKeysUn[i].pt.x, .pt.y, .angle, .octave) */
typedef struct {
  float x, y, angle;
  int32_t octave;
} asdm_kp;

/* Frame/KeyFrame grid geometry: Frame.h:37-38 (64x48), Frame.cc:34, KeyFrame.h:217-220.
 * KeyFrame's integer bounds are passed converted to float (exact). */
typedef struct {
  float min_x, min_y, max_x, max_y;
  float inv_w, inv_h;
  int32_t cols, rows;
} asdm_grid;

/* DBoW2::FeatureVector (src/dbow2/include/FeatureVector.h:21-22) as CSR: ascending node ids,
 * idx[off[k]..off[k+1]) = feature indices of node k in insertion order. A feature appears at most once
 * (TemplatedVocabulary.h:1125-1192 adds each feature to one node, or to none when its word is stopped). */
typedef struct {
  int32_t n_nodes;
  const uint32_t* node_ids;
  const int32_t* off;
  const int32_t* idx;
} asdm_featvec;

/* ---------------------------------------------------------------------------------------- context */
int asdm_create(int device, asdm_ctx** out);
/* A second context on the same GPU for another thread of the SLAM system (upstream ORB-SLAM2 runs tracking, local mapping
 * and loop closing concurrently): its own streams, staging arena, scratch and lock -- a thread never waits for another
 * thread's search -- and the right to search the sets the other contexts of the family created (a Frame's descriptors are
 * uploaded once, by the thread that builds it). The library orders the streams (an event per set behind whatever last
 * wrote it); the caller keeps a set alive while a call on it is in progress, as with any shared object, and destroys the
 * forked contexts before the parent. */
int asdm_fork(asdm_ctx* parent, asdm_ctx** out);
int asdm_destroy(asdm_ctx* ctx);
const char* asdm_last_error(const asdm_ctx* ctx);
void* asdm_stream(asdm_ctx* ctx); /* the context's cudaStream_t, for event timing by the caller */
int asdm_synchronize(asdm_ctx* ctx);
/* run on a caller-owned cudaStream_t instead (e.g. the stream an NCCL all-gather of the results is queued on);
 * NULL switches back to a stream owned by the context */
int asdm_set_stream(asdm_ctx* ctx, void* stream);
/* page-locked host memory for inputs / results, so that the copies inside the calls are true async DMA */
int asdm_host_alloc(void** out, size_t bytes);
int asdm_host_free(void* p);
/* options: "force_exact" (1 = skip the tensor-core shortlist and compute every gated pair in exact fp32),
 *          "cand_cap" (initial per-query candidate slots; grows automatically on overflow) */
int asdm_set_option(asdm_ctx* ctx, const char* key, int value);
/* counters of the last call/run on this context: "gpu_launches", "cand_cap", "retries", "tc_used" */
int asdm_get_stat(asdm_ctx* ctx, const char* key, int64_t* value);

/* ---------------------------------------------------------- descriptor sets (Frame.h:147, KeyFrame.h:199) */
/* Uploads once: fp32 rows (512 B, padded), the bf16 tensor-core copy and row norms, the keypoint SoA and
 * the 64x48 grid bucket (PosInGrid/AssignFeaturesToGrid, Frame.cc:123-138, 276-286). kps/grid may be NULL
 * for a pure descriptor pool (map-point descriptors). Refcounted: Frame copies and KeyFrame clones
 * (Frame.cc:47, KeyFrame.cc:153) retain instead of re-uploading. */
int asdm_set_create(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
                    const asdm_grid* grid, asdm_set** out);
/* `desc` may also point to memory of the context's device (descriptors produced on the GPU, or keyframes replicated
 * to every GPU over NVLink by an all-gather): the rows are then copied device-to-device, ordered on the context
 * stream, so the producer must have been queued on that stream (asdm_set_stream). kps stay host pointers. */
/* Same, but returns without waiting for the descriptor upload: `desc` must stay valid and unmodified until the next
 * call on this context that synchronises (asdm_synchronize, any search / fetch that returns results). kps / grid are
 * staged and may be reused at once. With page-locked `desc` (asdm_host_alloc) the copy is a true async DMA, so the
 * uploads of a batch of keyframes queue back to back and overlap the host-side preparation of the job that uses them. */
int asdm_set_create_async(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
                          const asdm_grid* grid, asdm_set** out);
/* A batch of sets in one call (keyframes that arrive together: a map being loaded, a cold loop-closure sweep). Same
 * result as asdm_set_create_async + asdm_set_featvec per set, but the small inputs of each set (keypoints, FeatureVector)
 * travel in one copy, one kernel launch prepares the whole batch, and the call costs a few microseconds of host time
 * per set -- so the host keeps up with the PCIe copy engine (2 MB of descriptors per 43 us). featvec may be NULL. */
typedef struct {
  const float* desc;
  int32_t n;
  int32_t row_stride_bytes; /* 0 = 512 */
  const asdm_kp* kps;
  const asdm_grid* grid;
  const asdm_featvec* featvec;
} asdm_set_desc;
int asdm_set_create_batch(asdm_ctx* ctx, const asdm_set_desc* descs, int n_sets, asdm_set** out);
int asdm_set_retain(asdm_set* set);
int asdm_set_release(asdm_set* set);
int asdm_set_size(const asdm_set* set);
/* attaches Frame::mFeatVec / KeyFrame::mFeatVec (after ComputeBoW, Frame.cc:289-296) */
int asdm_set_featvec(asdm_set* set, const asdm_featvec* fv);
/* reads rows back (map save / checkpoint: System.cc:296-439 writes descriptors from host Mats) */
int asdm_set_read_desc(asdm_set* set, int row0, int n_rows, float* out);

/* Map save / load. The descriptor block of one frame inside the reference's binary map file `chamo.map`
 * (vm::save_visual_map, src/visual_map/src/visual_map_seri.cc:134-146; vm::loader_visual_map, :272-285) is
 *     int32 desc_width (= 128)   int32 desc_count (= N)   N x 128 fp32, descriptor after descriptor
 * (System::saveToVisualMap fills the column-major Eigen D x N matrix from mDescriptors, System.cc:335-343): byte for
 * byte the resident rows, so saving is one device-to-host copy behind an 8-byte header and loading is one upload.
 * asdm_set_save_blocks writes the blocks of n_sets sets back to back into `out` (all copies queued, one wait);
 * offsets[k] (n_sets + 1 entries, may be NULL) receives where block k starts. Returns ASDM_ERR_INVALID when cap is too
 * small (asdm_desc_block_bytes says how much a set of n descriptors needs). */
size_t asdm_desc_block_bytes(int n);
int asdm_set_save_blocks(asdm_ctx* ctx, asdm_set* const* sets, int n_sets, void* out, size_t cap, size_t* offsets);
/* Creates a set from one block (the float descriptors System::LoadORBMap should hand to KeyFrame::setData; the
 * reference converts them to CV_8UC1 there, System.cc:64-69, which is stale ORB code). *consumed = bytes of the block.
 * Fails with ASDM_ERR_INVALID on a truncated block or desc_width != 128. Asynchronous like asdm_set_create_async:
 * `block` must stay valid until the next synchronising call. */
int asdm_set_load_block(asdm_ctx* ctx, const void* block, size_t bytes, const asdm_kp* kps, const asdm_grid* grid,
                        asdm_set** out, size_t* consumed);

/* --------------------------------------------------------------------------- BoW-gated dense search */
/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)  (ORBmatcher.cc:156-297).
 * valid_kf[i]: KF feature i holds a map point that is not bad.
 * match_f[j] (n_f entries): KF feature whose map point is assigned to frame feature j, or -1.
 * dist_f may be NULL. *nmatches = return value of the reference. */
int asdm_search_by_bow_kf_f(asdm_ctx* ctx, asdm_set* kf, const uint8_t* valid_kf, asdm_set* f, float nnratio,
                            int check_ori, int32_t* match_f, float* dist_f, int32_t* nmatches);

/* ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)  (ORBmatcher.cc:533-666).
 * match12[i] (n1 entries): feature of KF2 matched to feature i of KF1, or -1. */
int asdm_search_by_bow_kf_kf(asdm_ctx* ctx, asdm_set* kf1, const uint8_t* valid1, asdm_set* kf2,
                             const uint8_t* valid2, float nnratio, int check_ori, int32_t* match12, float* dist12,
                             int32_t* nmatches);

/* ORBmatcher::SearchForTriangulation (ORBmatcher.cc:669-822; CheckDistEpipolarLine :136-153).
 * has_mp1/2[i]: feature already holds a map point. F12 row-major 3x3; (ex,ey) = epipole in image 2
 * (:675-682). scale_factors2 / level_sigma2_2: pKF2->mvScaleFactors / mvLevelSigma2 (n_levels entries).
 * Pairs = {(i, match12[i]) : match12[i] >= 0}, ascending i (:814-819). */
int asdm_search_for_triangulation(asdm_ctx* ctx, asdm_set* kf1, const uint8_t* has_mp1, asdm_set* kf2,
                                  const uint8_t* has_mp2, const float* F12, float ex, float ey,
                                  const float* scale_factors2, const float* level_sigma2_2, int n_levels,
                                  int check_ori, int32_t* match12, float* dist12, int32_t* nmatches);

/* ----------------------------------------------------------------------- projection-gated searches */
/* Common descriptor loop of the four ORBmatcher::SearchByProjection overloads (ORBmatcher.cc:44-122, 300-413,
 * 1318-1452, 1455-1582). The adaptor supplies, per query, what the reference derives from poses and map points
 * before it reaches GetFeaturesInArea: the window (u, v, r), the octave range handed to Frame::GetFeaturesInArea
 * (Frame.cc:219-274; KeyFrame::GetFeaturesInArea has none: min_level = max_level = -1), whether the query's map
 * point blocks the keypoint for later queries (Observations()>0) and the query angle for the histogram. */
typedef struct {
  uint8_t valid;  /* query survived the reference's pre-descriptor gates (mbTrackInView, !isBad, in image ...) */
  uint8_t blocks; /* once assigned, later queries skip that keypoint (:86-88, :1393-1395) */
  int16_t pad;
  float u, v, r;
  int32_t min_level, max_level;
  float angle;
} asdm_proj_query;

enum {
  ASDM_PROJ_RATIO_LEVEL = 0, /* :44-122  best & second with octaves; ratio test only when both share an octave */
  ASDM_PROJ_BEST_ONLY = 1    /* :300-413, :1318-1452, :1455-1582 */
};

typedef struct {
  int32_t mode;
  float th_accept;     /* accept if best <= th_accept: TH_HIGH (:110, :1408), ORBdist (:1538), TH_LOW (:404) */
  float best_init;     /* 256 (:75, :380, :1518) or 100 (:1387) */
  float nnratio;       /* mode RATIO_LEVEL only */
  int32_t check_ori;   /* rotation histogram (:1413-1449, :1543-1579) */
  int32_t count_twice; /* :116-117 increments nmatches twice */
  /* Fuse's reprojection gate (:919-926): when chi2_max > 0 a keypoint is skipped if
   * e2 * inv_level_sigma2[octave] > chi2_max, e2 = (u-x)^2 + (v-y)^2 in fp32, compared in double (5.99) */
  double chi2_max;
  int32_t n_levels;
  int32_t reserved;
  float inv_level_sigma2[32];
} asdm_proj_params;

/* target: set with keypoints + grid (the Frame / KeyFrame searched). Query descriptors: either desc_q (host,
 * nq x 128, e.g. MapPoint::GetDescriptor of each local map point) or rows q_rows[i] of the resident set q_set
 * (q_rows NULL = row i). blocked0[j] (may be NULL): keypoint j unavailable from the start.
 * assign[j] (target->n entries) = query finally assigned to keypoint j; -1 = never assigned (the caller keeps what
 * the keypoint held); -2 = assigned and then cleared by the rotation check (the reference stores NULL, :1441). *nmatches = reference's return. */
int asdm_search_by_projection(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq, const float* desc_q,
                              asdm_set* q_set, const int32_t* q_rows, const uint8_t* blocked0,
                              const asdm_proj_params* p, int32_t* assign, float* dist_t, int32_t* nmatches);

/* Same search seen from the queries: query_best[i] = keypoint chosen by query i (or -1), query_dist[i] its distance.
 * For the entry points without exclusivity, where several queries may choose one keypoint: SearchBySim3
 * (ORBmatcher.cc:1090-1314, both directions) and the match phase of Fuse (:825-960, :963-1086; set every
 * query's `blocks` to 0 and apply the graph mutations on the host in query order). */
int asdm_search_by_projection_q(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq, const float* desc_q,
                                asdm_set* q_set, const int32_t* q_rows, const uint8_t* blocked0,
                                const asdm_proj_params* p, int32_t* query_best, float* query_dist, int32_t* nmatches);

/* ORBmatcher::SearchForInitialization (ORBmatcher.cc:416-531). prev_matched: 2*n1 floats (x,y), read and rewritten
 * (:526-528). matches12[i] (n1 entries) = index in F2 or -1. */
int asdm_search_for_initialization(asdm_ctx* ctx, asdm_set* f1, asdm_set* f2, float* prev_matched, int window_size,
                                   float nnratio, int check_ori, int32_t* matches12, float* dist12, int32_t* nmatches);

/* Descriptor stage of Frame::ComputeStereoMatches (Frame.cc:360-444): left keypoints = mvKeys, right = mvKeysRight.
 * best_r[i] = right index when bestDist < thOrbDist, else -1. */
int asdm_stereo_row_match(asdm_ctx* ctx, asdm_set* left, asdm_set* right, const float* scale_factors, int n_levels,
                          int n_rows, float min_d, float max_d, int32_t* best_r, float* best_dist);

/* The same search for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE. It calls no
 * cudaMalloc / cudaFree / cudaHostAlloc and takes nothing from the context's pools: its device scratch (ticket + device
 * copy of the results) is one file-scope device array of the library, 128.25 KB, held by every GPU that loads the library
 * until exit. Limits: left->n <= ASDM_STATIC_MAX_LEFT (larger: ASDM_ERR_INVALID with a message in asdm_last_error; use
 * asdm_stereo_row_match), right->n <= 10240 as for asdm_stereo_row_match.
 *   out    : caller-owned, 16-byte aligned memory the GPU can write (device memory, or page-locked host memory such as
 *            asdm_host_alloc returns) of 2 * pad * 4 bytes, pad = left->n rounded up to a multiple of 4:
 *            int32 best_r[pad] followed by float best_dist[pad] (entries >= left->n are unspecified).
 *   stream : a cudaStream_t of the context's device the work is queued on, or NULL for the context's own stream.
 * ASYNCHRONOUS: the call returns when the work is queued; `out` is complete once the stream has been waited for
 * (cudaStreamSynchronize, or asdm_synchronize(ctx) when stream is NULL). The first call on a device resolves the
 * array's device address (cudaGetSymbolAddress): make it once outside a capture. All calls share the one array, which
 * keeps the last call's contents (the ticket is cleared on the stream before every launch): calls that overlap on the
 * GPU -- two streams, two threads, two contexts of one device -- race on it; keep them ordered. */
#define ASDM_STATIC_MAX_LEFT 16384
int asdm_stereo_row_match_static(asdm_ctx* ctx, asdm_set* left, asdm_set* right, const float* scale_factors, int n_levels,
                                 int n_rows, float min_d, float max_d, void* out, void* stream);

/* ------------------------------------------------------------------ batched keyframe-pair jobs */
/* Loop bodies of LoopClosing.cc:291-318 (kind BOW_KF_KF) and LocalMapping.cc:335-370 (kind TRIANGULATION)
 * over independent keyframe pairs (ia[p], ib[p]) drawn from n_kf keyframes. masks[k] = valid (BoW) or
 * has_mp (triangulation) of keyframe k. prepare() uploads the per-call inputs; run() launches the kernels
 * on the context stream without synchronising; fetch() copies results to the host (row stride
 * asdm_job_out_stride()) and synchronises. Pairs pair_begin..pair_end-1 are processed (sharding by rank). */
enum { ASDM_JOB_BOW_KF_KF = 1, ASDM_JOB_TRIANGULATION = 2 };

typedef struct {
  int32_t kind;
  int32_t n_kf;
  asdm_set* const* sets;        /* n_kf */
  const uint8_t* const* masks;  /* n_kf */
  int32_t n_pairs;
  const int32_t* ia;
  const int32_t* ib;
  float nnratio;                /* BoW */
  int32_t check_ori;
  const float* F12;             /* triangulation: 9*n_pairs */
  const float* epipoles;        /* triangulation: 2*n_pairs */
  const float* scale_factors;   /* triangulation: n_levels */
  const float* level_sigma2;    /* triangulation: n_levels */
  int32_t n_levels;
  /* Optional (NULL = whole keyframe): pair p only searches query rows [row_begin[p], row_end[p]) of sets[ia[p]], so a
   * pair can be split into row-block units across GPUs. Triangulation with check_ori == 0 only: that search has no
   * cross-query state (vbMatched2 is never set, ORBmatcher.cc:669-822), the BoW searches are greedy and the rotation
   * histogram is per pair. Rows outside the range come back as -1. Multiples of 128 avoid wasted tensor-core tiles. */
  const int32_t* row_begin;
  const int32_t* row_end;
} asdm_job_desc;

int asdm_job_prepare(asdm_ctx* ctx, const asdm_job_desc* desc, asdm_job** out);
int asdm_job_run(asdm_job* job);
int asdm_job_fetch(asdm_job* job, int32_t* match12, float* dist12, int32_t* nmatches);
/* Queues the device-to-host copies of the results behind the run and returns at once, so the copies of one job overlap
 * the kernels of the next (page-locked destinations). The buffers hold the results only after asdm_job_fetch on the same
 * pointers returned: it validates the run (candidate pools, speculative tensor-core route) and copies again only if the
 * job had to be rerun. */
int asdm_job_fetch_async(asdm_job* job, int32_t* match12, float* dist12, int32_t* nmatches);
/* waits for the last run and, if a candidate pool overflowed, grows it and runs again; after it returns the device
 * outputs are final (fetch() calls it) */
int asdm_job_settle(asdm_job* job);
int asdm_job_out_stride(const asdm_job* job);
/* device pointers of the results (match12: n_pairs*stride int32, dist12: same float, nmatches: n_pairs),
 * e.g. for an NCCL all-gather by the caller */
int asdm_job_device_outputs(asdm_job* job, void** match12, void** dist12, void** nmatches);
int asdm_job_destroy(asdm_job* job);

/* ---- results published into every rank's gathered arrays by the job itself (one process per GPU) ----
 * The exchange step of a sharded sweep -- every rank ends up with the match lists of all ranks -- as the epilogue of the
 * job's last kernel instead of a collective call behind it: each rank allocates its gathered record arrays with
 * asdm_shared_alloc, hands the 64-byte handle to the other processes of the node (any transport: torch.distributed
 * object all-gather, a pipe), opens theirs with asdm_shared_open (cudaIpc mapping, NVLink peer access) and tells its job
 * where row `row0 + pair` of every rank's arrays is. From then on asdm_job_run stores each finished row into all of
 * them; a barrier of the group on the same stream (e.g. a one-word NCCL all-reduce) is all that is left for the caller
 * before reading. Replaces the three ncclAllGather calls of the north-star design, not a reference function. */
#define ASDM_SHARED_HANDLE_BYTES 64
int asdm_shared_alloc(asdm_ctx* ctx, size_t bytes, void** dptr, uint8_t handle[ASDM_SHARED_HANDLE_BYTES]);
int asdm_shared_open(asdm_ctx* ctx, const uint8_t handle[ASDM_SHARED_HANDLE_BYTES], void** dptr);
int asdm_shared_close(asdm_ctx* ctx, void* dptr); /* a pointer from asdm_shared_open */
int asdm_shared_free(asdm_ctx* ctx, void* dptr);  /* a pointer from asdm_shared_alloc */
/* n_ranks <= 16 destinations (this rank's own arrays included; 0 switches publishing off). match / dist: [rows][stride]
 * int32 / fp32 with stride >= asdm_job_out_stride (the tail of a row is filled with -1), nmatches: [rows] int32. May be
 * called between runs (e.g. to alternate between two generations of arrays). */
#define ASDM_PUBLISH_MULTICAST 1 /* the destinations are NVSwitch multicast mappings (e.g. torch symmetric memory's
                                  * multicast_ptr + offset): one store reaches every rank bound to them, n_ranks = 1 */
#define ASDM_PUBLISH_NO_DIST 2   /* publish match indices + counts only -- what the reference's calls return (vpMatches12,
                                  * vMatchedPairs); the distance rows stay in the job's local outputs */
int asdm_job_set_publish(asdm_job* job, int n_ranks, void* const* match, void* const* dist, void* const* nmatches,
                         int row0, int stride, int flags);

/* ------------------------------------------------------------------ multi-GPU groups (north_star (3)) */
/* The batched keyframe-pair searches over several GPUs from plain C / C++: no libtorch, no NCCL, no Python on the path.
 * A group is `world` ranks, one GPU each. Two ways to get one:
 *   asdm_group_create  one process drives every GPU (the SLAM process: LocalMapping::CreateNewMapPoints over
 *                      LocalMapping.cc:335-370, LoopClosing::ComputeSim3 over LoopClosing.cc:291-318). `devices` may name a
 *                      GPU more than once (two ranks on one GPU: how the tests run on a single-GPU box).
 *   asdm_group_join    one process per GPU (the bench contract). The processes of the node exchange the handles of their
 *                      device windows through `allgather`, a caller-supplied all-gather of small host blobs (MPI, a pipe,
 *                      torch.distributed ...), called a few times inside asdm_group_join / asdm_group_host_alloc only --
 *                      never on the search path. All ranks must make the same group calls in the same order.
 * Every rank owns a WINDOW of device memory that every other rank can store into (peer access inside one process, cudaIpc
 * mappings between processes). Replicated keyframes and gathered results live at the same offset of every window, so no
 * handle is exchanged after the group exists. Exchange steps are fused into the producing kernels (peer stores over
 * NVLink) and ordered by a device-side barrier (one flag store per peer + a bounded spin: asdm_group_* never blocks the
 * host on another rank except in asdm_group_job_wait / asdm_group_host_barrier). */
typedef struct asdm_group asdm_group;
typedef struct asdm_gjob asdm_gjob;
typedef int (*asdm_allgather_fn)(void* user, const void* send, void* recv, size_t bytes_per_rank); /* 0 = ok; rank order */

int asdm_group_create(const int* devices, int n, size_t window_bytes /* 0 = 2 GiB */, asdm_group** out);
int asdm_group_join(asdm_ctx* ctx, int rank, int world, size_t window_bytes, asdm_allgather_fn allgather, void* user,
                    asdm_group** out);
int asdm_group_destroy(asdm_group* g);
const char* asdm_group_last_error(const asdm_group* g);
int asdm_group_world(const asdm_group* g);
int asdm_group_local_count(const asdm_group* g);       /* ranks hosted by this process */
int asdm_group_local_rank(const asdm_group* g, int l); /* global rank of local member l */
asdm_ctx* asdm_group_ctx(asdm_group* g, int l);        /* its context (streams, options, counters) */
/* Host memory every rank's GPU can store into and every process of the group can read: page-locked; one process: plain
 * pinned memory; several: one shared mapping (memfd) registered with CUDA in each. Collective. */
int asdm_group_host_alloc(asdm_group* g, size_t bytes, void** out);
int asdm_group_host_free(asdm_group* g, void* p);
int asdm_group_host_barrier(asdm_group* g); /* all processes of the group (no-op with one process) */

/* A batch of keyframes replicated on every GPU of the group: keyframe k crosses PCIe ONCE -- rank k % world
 * uploads it from descs[k] (only the process hosting that rank reads descs[k].desc / kps / featvec; n, and whether kps /
 * grid / featvec are given, must agree everywhere) -- and the owner's pack kernel stores the fp32 rows, the bf16 copy, the
 * norms, the keypoints and the FeatureVector arrays straight into every peer's window. out[k * local_count + l] = the
 * set of keyframe k on local member l: an ordinary asdm_set (a view; release each with asdm_set_release). Asynchronous:
 * descs[k].desc must stay valid until the next synchronising call. Searches queued later on the members' streams see
 * the whole batch (a device-side barrier closes the exchange). */
int asdm_group_sets_create(asdm_group* g, const asdm_set_desc* descs, int n_sets, asdm_set** out);

/* A batched search dealt over the group: unit u (a pair, or one of `blocks_per_pair` 128-row-aligned query row blocks of a
 * pair: triangulation without rotation check only, see asdm_job_desc) belongs to rank u % world. job.sets = the array
 * asdm_group_sets_create returned ([n_kf * local_count]). */
enum {
  ASDM_GATHER_DEVICE = 1, /* every rank's GPU also receives every unit's records (asdm_group_job_gathered) */
  ASDM_GATHER_NO_DIST = 2 /* ... index lists + counts only, what the reference's calls return (vpMatches12, vMatchedPairs): the
                           * distance rows of a unit stay on the rank that computed it (half the NVLink traffic) */
};
typedef struct {
  asdm_job_desc job;
  int32_t blocks_per_pair; /* >= 1 */
  int32_t gather;          /* 0, ASDM_GATHER_DEVICE, or ASDM_GATHER_DEVICE | ASDM_GATHER_NO_DIST */
} asdm_gjob_desc;
int asdm_group_job_prepare(asdm_group* g, const asdm_gjob_desc* desc, asdm_gjob** out);
/* The unit plan asdm_group_job_prepare derives, as plain host arithmetic (no device, no communication; every rank computes
 * the same table): pair p with query_rows[p] rows becomes min(blocks_per_pair, ceil(rows / 128)) units whose row ranges
 * [unit_row0, unit_row1) start on 128-row boundaries and tile [0, rows); unit u is searched by rank u % world. Output
 * arrays of `cap` entries (each may be NULL); *n_units = units planned. Replaces the per-pair loop bodies the reference
 * runs serially (LocalMapping.cc:335-370, LoopClosing.cc:291-318) as the scheduling step of the sharded form. */
int asdm_group_plan_units(const int32_t* query_rows, int n_pairs, int blocks_per_pair, int world, int32_t* unit_pair,
                          int32_t* unit_row0, int32_t* unit_row1, int32_t* unit_rank, int cap, int* n_units);
int asdm_group_job_units(const asdm_gjob* j);  /* n_pairs, or more when pairs are split */
int asdm_group_job_stride(const asdm_gjob* j); /* entries per result row */
/* Launches the units of every local member; returns without waiting. h_match / h_dist: [n_pairs][stride] in the caller's
 * pair order, h_unit_nm: [units] -- device-visible host memory (asdm_group_host_alloc / asdm_host_alloc) the kernels
 * store the finished rows into as each unit retires, every rank over its own PCIe link; any of them may be NULL. */
int asdm_group_job_run(asdm_gjob* j, int32_t* h_match, float* h_dist, int32_t* h_unit_nm);
/* Waits for the local units (growing candidate pools and re-running if one overflowed) and for the other processes of the
 * group: afterwards the host arrays handed to asdm_group_job_run are complete. */
int asdm_group_job_wait(asdm_gjob* j);
/* nmatches[p] = sum of the counts of pair p's units */
int asdm_group_job_counts(const asdm_gjob* j, const int32_t* h_unit_nm, int32_t* nmatches);
/* run + wait + copy into ordinary (pageable) arrays: the whole batch on every process, as one un-sharded asdm_job
 * over all pairs would report it. Any pointer may be NULL. */
int asdm_group_job_fetch(asdm_gjob* j, int32_t* match12, float* dist12, int32_t* nmatches);
/* ASDM_GATHER_DEVICE: device pointers, on local member l, of the records of ALL units as the last run left them
 * (match / dist: [units or pairs][stride] as above, unit_nm: [units]); complete after asdm_group_job_wait, or -- without
 * a host wait -- for work queued on the member's stream behind asdm_group_job_flush. The barrier that closes the exchange
 * of a run is queued on a stream of its own: the next run's kernels do not wait for the last rows of this one to cross
 * NVLink (three generations of arrays alternate, so the exchange of a run overlaps the next run). */
int asdm_group_job_gathered(asdm_gjob* j, int l, void** match, void** dist, void** unit_nm);
int asdm_group_job_flush(asdm_gjob* j);
int asdm_group_job_destroy(asdm_gjob* j);

/* ------------------------------------------------------------------ map-point descriptor maintenance */
/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:271-338) for a batch of map points, e.g. all points matched by
 * a new keyframe (LocalMapping.cc:245, 535, 628). Observations of point p are rows (sets[obs_set[e]], obs_row[e]),
 * e in [obs_off[p], obs_off[p+1]), listed in the reference's iteration order (its std::map<KeyFrame*, size_t>), bad
 * keyframes already dropped (:296). best[p] = position in the point's list of the descriptor to become mDescriptor
 * (least median distance to the rest, first wins, :321-334), or -1 when the point has no observation (the reference
 * returns without touching mDescriptor). best_median may be NULL. */
int asdm_distinctive_descriptors(asdm_ctx* ctx, asdm_set* const* sets, int n_sets, int n_points, const int32_t* obs_off,
                                 const int32_t* obs_set, const int32_t* obs_row, int32_t* best, float* best_median);

/* ------------------------------------------------------------------ vocabulary tree (DBoW2) on the device */
/* TemplatedVocabulary<FSift::TDescriptor, FSift> (src/dbow2/include/TemplatedVocabulary.h; the ORBVocabulary typedef,
 * src/vslam/include/ORBVocabulary.h:34) as flat arrays, in the state TemplatedVocabulary::load leaves it (:1455-1504):
 * node 0 is the root, children of a node are listed in push_back order (:1486), desc holds Node::descriptor
 * (FSift::fromString), weight Node::weight, word_id Node::word_id of the leaves (ignored for inner nodes). */
typedef struct asdm_vocab asdm_vocab;
typedef struct {
  int32_t k, L;             /* m_k, m_L */
  int32_t weighting;        /* DBoW2::WeightingType: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY (BowVector.h:36-42) */
  int32_t scoring;          /* DBoW2::ScoringType: 0 L1_NORM, 1 L2_NORM, 2 CHI_SQUARE, 3 KL, 4 BHATTACHARYYA, 5 DOT_PRODUCT */
  int32_t n_nodes;          /* including the root */
  const int32_t* child_off; /* n_nodes + 1 */
  const int32_t* child_idx; /* child_off[n_nodes] node ids */
  const float* desc;        /* n_nodes x 128 (row 0, the root, is not read) */
  const double* weight;     /* n_nodes */
  const int32_t* word_id;   /* n_nodes */
} asdm_vocab_desc;

int asdm_vocab_create(asdm_ctx* ctx, const asdm_vocab_desc* desc, asdm_vocab** out);
int asdm_vocab_release(asdm_vocab* voc);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1125-1192) over the rows of `set`,
 * as called by Frame::ComputeBoW (Frame.cc:289-296) and KeyFrame::ComputeBoW (KeyFrame.cc:203-219) with levelsup = 4.
 * The tree descent (:1216-1257, FSift::distance in double, FSift.cpp:86-101) runs on the device and the resulting
 * FeatureVector is attached to the set there (as asdm_set_featvec would), so the gating data of the BoW searches never
 * crosses PCIe. Returned to the host for the caller's mBowVec / mFeatVec:
 *   word_id[i]  word of feature i (n entries);  node_id[i]  its node levelsup levels above the leaves (0 = root);
 *   in_fv[i]    1 when the word is not stopped (weight > 0, :1155) -- only those features enter either vector;
 *   bow_id / bow_val (capacity n; *bow_n entries, ascending word id) = the BowVector after weighting and
 *   normalisation (:1137-1189, BowVector.cpp:61-84), bit-identical doubles.
 * Any output pointer may be NULL. An empty vocabulary cannot be created, so the early return of :1132 does not arise. */
int asdm_vocab_transform(asdm_ctx* ctx, asdm_vocab* voc, asdm_set* set, int levelsup, int32_t* word_id,
                         int32_t* node_id, uint8_t* in_fv, int32_t* bow_n, uint32_t* bow_id, double* bow_val);

/* ------------------------------------------------------------------------------------------------------------------
 * Keyframe database: BoW candidate detection -- replaces ORB_SLAM2::KeyFrameDatabase (reference
 * src/vslam/include/KeyFrameDatabase.h:40-62, src/vslam/src/KeyFrameDatabase.cc:34-320) with the L1 scoring of the
 * vocabulary (DBoW2::L1Scoring::score, src/dbow2/DBoW2/ScoringObject.cpp:24-61). The BowVectors of all keyframes stay
 * resident on the device; one kernel counts the common words and scores every keyframe against the query
 * (k_bowdb_query), the list logic of the reference (gates, covisibility-group accumulation, order) is finished on the
 * host inside the call. Keyframes are named by SLOT = the order of asdm_bowdb_add (= their order inside every inverted
 * list of the reference); slots are never reused.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct asdm_bowdb asdm_bowdb;
int asdm_bowdb_create(asdm_ctx* ctx, asdm_bowdb** out);                  /* KeyFrameDatabase(voc) :34-38 */
int asdm_bowdb_release(asdm_bowdb* db);
/* KeyFrameDatabase::add (:41-47): word ids ascending (std::map order) with their values; *slot receives the new slot */
int asdm_bowdb_add(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n_words, int32_t* slot);
int asdm_bowdb_erase(asdm_bowdb* db, int32_t slot);                       /* KeyFrameDatabase::erase :49-70 */
int asdm_bowdb_clear(asdm_bowdb* db);                                     /* KeyFrameDatabase::clear :72-76 */
int asdm_bowdb_size(asdm_bowdb* db);                                      /* slots handed out so far (erased included) */
/* The device pass alone: for every slot the number of words shared with the query, the smallest shared word
 * (0xffffffff if none) and the L1 score as the reference's float. Arrays of asdm_bowdb_size entries; erased slots: 0. */
int asdm_bowdb_query(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n_words, int32_t* common,
                     uint32_t* first_word, float* score);
typedef struct {
  const uint8_t* connected;   /* [size] pKF->GetConnectedKeyFrames() membership (loop detection only), may be NULL */
  const uint8_t* global_flag; /* [size] KeyFrame::GetGlobalMapFlag(), may be NULL when only_global_map == 0 */
  const int32_t* neigh_off;   /* [size + 1] CSR of GetBestCovisibilityKeyFrames(10) per slot, may be NULL (no neighbours) */
  const int32_t* neigh_slot;
  float min_score;            /* loop detection: minScore */
  int32_t only_global_map;
} asdm_bowdb_gates;
/* KeyFrameDatabase::DetectLoopCandidates(pKF, minScore, only_global_map) :80-216. out: candidate slots in the
 * reference's order; ASDM_ERR_INVALID if more than cap. */
int asdm_bowdb_detect_loop(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n_words,
                           const asdm_bowdb_gates* gates, int32_t* out, int cap, int32_t* n_out);
/* KeyFrameDatabase::DetectRelocalizationCandidates(F, only_global_map) :219-320. The keyframes' mRelocScore persists
 * between calls as in the reference (it starts at 0 here; the reference leaves it uninitialised, KeyFrame.cc:150). */
int asdm_bowdb_detect_reloc(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n_words,
                            const asdm_bowdb_gates* gates, int32_t* out, int cap, int32_t* n_out);
int asdm_bowdb_reloc_scores(asdm_bowdb* db, float* score);               /* [size] the persistent mRelocScore values */
/* The C++ drop-in (host/KeyFrameDatabase_cuda.cc) keeps the reference's bookkeeping fields on its KeyFrames
 * (KeyFrame.h:175-180: mnLoopWords / mLoopScore / mnRelocWords / mRelocScore): the same two calls that also hand back the
 * per-slot arrays of the device pass (as asdm_bowdb_query; any of the three may be NULL), and a setter for mRelocScore so
 * that the KeyFrame member stays the source of truth. */
int asdm_bowdb_detect_loop_full(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n_words,
                                const asdm_bowdb_gates* gates, int32_t* out, int cap, int32_t* n_out, int32_t* common,
                                uint32_t* first_word, float* score);
int asdm_bowdb_detect_reloc_full(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n_words,
                                 const asdm_bowdb_gates* gates, int32_t* out, int cap, int32_t* n_out, int32_t* common,
                                 uint32_t* first_word, float* score);
int asdm_bowdb_set_reloc_scores(asdm_bowdb* db, const float* score);     /* [size] */

#ifdef __cplusplus
}
#endif
#endif
