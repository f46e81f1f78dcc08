// internal.cuh -- shared host/device declarations of libasdmatch_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/asdmatch.h"

#define ASDM_MAX_LEVELS 32
#define ASDM_BEST_INIT_BOW 256.0f  // ORBmatcher.cc:204-206, 577-579

// A device allocation that only ever grows (scratch reused across calls on one context).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

// ------------------------------------------------------------------------------------------ host structs
struct asdm_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  // descriptor uploads ride their own stream so the DMA of keyframe k+1 overlaps the pack / grid kernels of keyframe k
  cudaStream_t upload_stream = nullptr;
  cudaEvent_t upload_done = nullptr;
  // ... and the per-set preparation (bf16 copy, norms, range flag) a third one, so that neither the copy engine nor the
  // search kernels on the main stream wait behind it: searches of earlier keyframes overlap the upload AND the packing
  // of later ones (StreamedPairSweep). The main stream waits for prep_done of every set at creation time, so any
  // search queued later is ordered after the sets it may touch.
  cudaStream_t prep_stream = nullptr;
  cudaEvent_t prep_done = nullptr;
  // results queued by asdm_job_fetch_async leave on a fourth stream, so the copies of one wave overlap the next wave
  cudaStream_t download_stream = nullptr;
  cudaEvent_t run_done = nullptr;
  bool downloads_pending = false;
  // Scratch blocks of jobs / calls (DevBuf) are recycled through this cache instead of going back to the driver pool:
  // they are only ever touched on the context stream, so stream order makes the reuse safe, and a job costs no
  // cudaMallocAsync / cudaFreeAsync calls in steady state. Keyed by size class.
  std::map<size_t, std::vector<void*>> block_cache;
  size_t block_cache_bytes = 0;
  std::recursive_mutex mu;
  std::string err;
  // Contexts forked from one another (asdm_fork: one per thread of the SLAM system, each with its own streams, staging
  // arena and scratch) form a family on one GPU and may search each other's sets; `family` names the root.
  asdm_ctx* family = nullptr;
  std::mutex slot_mu;  // the pinned slot slab is handed out / taken back from any thread that releases a set
  int opt_force_exact = 0;
  int opt_cand_cap = 8;
  int opt_list_cap_per_query = 4;  // tensor-core short list slots per query row (grows on overflow)
  int opt_tc_epi_warps = 8;        // epilogue warps of the tensor-core kernel: 4 or 8
  int opt_voc_direct = 0;          // vocabulary descent: 1 = children read straight from global memory (no staging)
  int opt_bowdb_filter_bits = 128; // keyframe database: Bloom filter bits per query word (rounded up to a power of two overall)
  int opt_bowdb_search = 0;        // keyframe database: 1 = binary-search kernel instead of the shared-memory hash table
  int opt_voc_exact = 0;           // vocabulary descent: 1 = always run the exact double chain (no fp32 fast pass)
  int64_t stat_voc_exact = 0;      // (feature, level) decisions of the last transform that needed the exact chain
  int opt_init_serial = 0;         // SearchForInitialization: 1 = the serial replay kernel instead of the Jacobi sweeps
  int64_t stat_init_sweeps = 0;    // 1 when the last SearchForInitialization was resolved by the sweeps
  int opt_tc_variant = 0;          // pipeline-shape experiments of the tensor-core kernel (tc_shortlist.cu)
  int64_t stat_launches = 0;       // kernels launched by the last call / job run
  int64_t stat_total_launches = 0;
  int64_t stat_cand_cap = 0;
  int64_t stat_retries = 0;
  int64_t stat_tc_used = 0;
  int64_t stat_tc_pairs = 0;  // (i,j) pairs emitted by the tensor-core shortlist in the last run
  int sm_count = 148;
  void* tmap_encode = nullptr;  // cuTensorMapEncodeTiled, resolved through the runtime
  // Pinned staging arena (bump allocator): small per-call inputs (keypoints, masks, FeatureVector flattening, work
  // lists) are written here and DMA'd from here, so no call has to wait for its own upload. Everything staged is
  // consumed in stream order; the arena rewinds whenever the library synchronises the stream.
  char* stage = nullptr;
  size_t stage_cap = 0, stage_off = 0;
  char* res_host = nullptr;  // page-locked result block of the single-pair searches (their kernels store the rows here)
  size_t res_cap = 0;
  // Pinned 64-byte slots with object lifetime (per-job overflow flags, per-set tensor-core range flag)
  char* slot_slab = nullptr;
  std::vector<int> slot_free;
  DevBuf scratch[12];
  int opt_time_kernels = 0;     // record CUDA events around the distance stage of every job run
  std::vector<cudaEvent_t> ev;  // pairs (begin, end), drained by asdm_get_stat("dist_stage_us")
  cudaEvent_t stage_ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // time_kernels == 2: begin | zero | distance | rerank | resolve of the last run
  int proj_ent_cap = 1 << 16;  // entries of the projection candidate pool (grows on overflow)
  int resolve_level_hint = 0;  // highest k_resolve_bow budget level a job of this context needed (or option "resolve_level")
  int opt_proj_legacy = 0;     // 1 = the three-kernel projection search (gather / dist / resolve) instead of k_proj_fused
  int32_t stat_proj_ns[6] = {0, 0, 0, 0, 0, 0};  // k_proj_fused of the last call: phase durations (ns) and sweeps
  int32_t stat_proj_host_ns[3] = {0, 0, 0};      // host side of the same call: staging | launch | wait
  void* proj_ctl = nullptr;    // ProjCtl of k_proj_fused: zero between calls (the kernel's last CTA resets it)
};

// small per-set arrays (range flag | keypoints | FeatureVector) of a batch of sets live in ONE device block, filled by
// one copy (asdm_set_create_batch); the sets of the batch share it
struct SideBlock {
  void* p = nullptr;
  std::atomic<int> ref{0};  // sets may be released from any thread of the family
};

struct asdm_set {
  asdm_ctx* ctx = nullptr;
  SideBlock* side = nullptr;
  std::atomic<int> ref{1};
  int n = 0;
  int n_pad = 0;                     // rows allocated (multiple of 256; padding rows are zero)
  float* desc = nullptr;             // [n_pad][128] fp32, 512 B rows
  __nv_bfloat16* desc_bf = nullptr;  // [n_pad][128] bf16 (tensor-core operand, K-major)
  float* norm2 = nullptr;            // [n_pad] sum of squares of the fp32 row
  int32_t* d_flags = nullptr;        // [0] = 1 when some row is non-finite or its norm is out of the TC range
  int tc_ok = -1;                    // -1 unknown (flags not read back yet), 0 no, 1 yes
  int32_t* h_flag = nullptr;         // pinned slot the flag is copied to right after k_pack_desc (resolved lazily)
  void* fv_mem = nullptr;            // allocation holding node_id / fv_pos
  bool has_kps = false, has_grid = false, has_fv = false;
  asdm_kp* kps = nullptr;  // [n] cv::KeyPoint fields as uploaded (16 B each: one LDG.128 per keypoint)
  asdm_grid grid{};
  int32_t* cell_start = nullptr;  // [cols*rows+1]
  int32_t* cell_items = nullptr;  // [n] insertion (index) order inside a cell
  int32_t* cell_tmp = nullptr;    // [n] scratch of k_build_grid
  int4* cell_kp = nullptr;        // [n] keypoints in cell order: {x bits, y bits, octave, index} (one LDG.128 per candidate)
  bool grid_built = false;        // the grid bucket is built by the first projection-gated search that needs it
  uint32_t* node_id = nullptr;    // [n] FeatureVector node of the feature
  int32_t* fv_pos = nullptr;      // [n] CSR position of the feature in the FeatureVector, -1 if absent
  std::vector<uint8_t> tmap_bf;   // CUtensorMap (128 B) over desc_bf, built lazily
  // Sets replicated over a multi-GPU group (group.cu) are VIEWS into the group's device window: rows, bf16 copy, norms,
  // keypoints and FeatureVector arrays belong to the slab of their batch (refcounted), only the grid bucket is the
  // set's own (allocated by the first projection-gated search).
  struct GroupSlab* slab = nullptr;
  void* grid_mem = nullptr;
  // Shared use across the contexts of a family: `ready` is recorded behind whatever last wrote the set's arrays (pack
  // kernel, FeatureVector upload, vocabulary transform); a context other than the owner orders its stream behind it
  // (use_set). The lazily built parts (grid bucket, tensor map, range flag) are guarded by `mu`.
  std::mutex mu;
  cudaEvent_t ready = nullptr;
  cudaEvent_t grid_ev = nullptr;   // behind k_build_grid, recorded on the stream of the context that built it
  asdm_ctx* grid_builder = nullptr;
};

struct asdm_group;
struct GroupSlab {
  asdm_group* g = nullptr;
  int member = 0;            // local member whose window holds this copy
  size_t off = 0, bytes = 0; // range of the window
  std::atomic<int> ref{0};   // sets alive (released from any thread)
  int32_t* h_flags = nullptr; // page-locked [n_sets]: tensor-core range flags of the batch, copied back after the exchange
};
void group_slab_release(GroupSlab* s);  // group.cu

// DBoW2 vocabulary tree resident on the device (kernels_vocab.cuh) plus the host copies of what the BowVector needs
struct asdm_vocab {
  asdm_ctx* ctx = nullptr;
  int k = 0, L = 0, weighting = 0, scoring = 0, n_nodes = 0;
  int G = 1;               // lanes per feature in k_voc_descend
  int kmax = 1;            // widest fan-out of the tree
  void* mem = nullptr;     // one arena: slot records | slot rows (kernels_vocab.cuh)
  int4* meta = nullptr;
  float* rows = nullptr;
  int root_k = 0;
  std::vector<double> h_weight;
  std::vector<int32_t> h_word;
};

// --------------------------------------------------------------------------------- device-visible structs
// Where a sharded job publishes its results: the rank-major gathered record arrays of every rank of the
// group -- this GPU's own and its NVLink peers' (cudaIpc mappings) -- so that the all-gather of the match lists is the
// epilogue of the resolve kernel: each CTA stores its pair's finished row into all of them.
#define ASDM_MAX_PEERS 16
struct PublishArgs {
  int n = 0;                        // 0 = the job keeps its results local
  int row0 = 0;                     // record slot of local unit k = row0 + k * row_step (counts always; rows unless row_map)
  int row_step = 1;                 // 1: this rank's units are contiguous (rank-major); world: round-robin deal, unit-major
  const int32_t* row_map = nullptr; // device array [n_pairs], or null: destination ROW of unit k when units are row blocks
                                    // of a pair (several units share a row and each publishes only its own columns)
  int col_range = 0;                // 1: publish columns [rowA0, rowA1) of the unit only (row-block units)
  int stride = 0;                   // entries per row of the gathered arrays
  int multicast = 0;                // destinations are NVLink multicast (NVLS) addresses: one store reaches every rank
  int no_dist = 0;                  // publish index lists + counts only (what the reference calls return); distances stay local
  unsigned dist_mask = 0xffffffffu; // bit p: destination p takes the distance rows too (when no_dist == 0)
  int32_t* match[ASDM_MAX_PEERS];   // [world * per_rank][stride]
  float* dist[ASDM_MAX_PEERS];
  int32_t* nm[ASDM_MAX_PEERS];      // [world * per_rank]
};

// One keyframe pair of a BoW-gated dense search (a8 / a9 / a11).
struct PairDesc {
  const float* descA;
  const float* descB;
  const __nv_bfloat16* bfA;
  const __nv_bfloat16* bfB;
  const float* normA;
  const float* normB;
  const uint32_t* nodeA;
  const uint32_t* nodeB;
  const int32_t* posA;  // query processing order (CSR position); <0 = not in the FeatureVector
  const int32_t* posB;  // target scan order
  const uint8_t* maskA;
  const uint8_t* maskB;
  const asdm_kp* kpA;
  const asdm_kp* kpB;
  int32_t nA, nB;
  int32_t rowA0, rowA1;  // query rows [rowA0, rowA1) of set A belong to this unit (whole set: 0, nA)
  int32_t setA, setB;    // index into the job's tensor-map table
  // tensor-core shortlist gates, precomputed per keyframe by k_tc_row_bounds (tc_shortlist.cu):
  const float* rbA;    // [n_pad] KAPPA*|a_i|^2 when query i is usable (mask, FeatureVector membership), else +inf
  const float* cminB;  // [n_pad/128] min over each 128-row chunk of the same bound on the target side
  float F12[9];
  float ex, ey;
};

struct BowParams {
  int32_t kind;        // KIND_*
  int32_t mask_sense;  // feature usable iff mask == mask_sense
  float t_keep;        // keep candidate iff d <= t_keep
  float th_accept;     // TH_LOW
  int32_t strict;      // a9: best < TH_LOW (ORBmatcher.cc:609); a8: <= (:231)
  float nnratio;
  int32_t check_ori;
  int32_t n_levels;
  float scale_factors[ASDM_MAX_LEVELS];
  float level_sigma2[ASDM_MAX_LEVELS];
};

#define KIND_BOW_KF_KF 1
#define KIND_TRIANGULATION 2
#define KIND_BOW_KF_F 3

struct Cand {  // 16 B: one LDG.128 brings everything the resolve kernels rank a candidate by
  int32_t j;
  float d;
  int32_t pos;  // posB[j]: the target's place in the reference's scan (tie order among equal distances)
  int32_t pad;  // posA[i] << 5 | rotation bin of the pair (cand_meta)
};

struct TcHit {  // one (pair, query, target) short-listed by the tensor-core kernel
  int32_t pair, i, j;
};

// ------------------------------------------------------------------------------------------ device helpers
// ORBmatcher::DescriptorDistance (ORBmatcher.cc:1637-1644): one fp32 accumulator, index ascending,
// separate sub / mul / add with round-to-nearest -- never contracted into FMA.
__device__ __forceinline__ float asdm_exact_dist(const float* __restrict__ a, const float* __restrict__ b) {
  const float4* a4 = reinterpret_cast<const float4*>(a);
  const float4* b4 = reinterpret_cast<const float4*>(b);
  float s = 0.f;
#pragma unroll 8
  for (int k = 0; k < ASDM_DESC_DIM / 4; k++) {
    const float4 x = __ldg(a4 + k), y = __ldg(b4 + k);
    float d;
    d = __fsub_rn(x.x, y.x); s = __fadd_rn(s, __fmul_rn(d, d));
    d = __fsub_rn(x.y, y.y); s = __fadd_rn(s, __fmul_rn(d, d));
    d = __fsub_rn(x.z, y.z); s = __fadd_rn(s, __fmul_rn(d, d));
    d = __fsub_rn(x.w, y.w); s = __fadd_rn(s, __fmul_rn(d, d));
  }
  return s;
}

// rotation-bin idiom, e.g. ORBmatcher.cc:245-250 (factor = 1.0f/HISTO_LENGTH, so only bins 0..12 occur)
__device__ __forceinline__ int asdm_rot_bin(float a1, float a2) {
  const float factor = 1.0f / ASDM_HISTO_LENGTH;
  float rot = __fsub_rn(a1, a2);
  if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
  int bin = (int)roundf(__fmul_rn(rot, factor));
  if (bin == ASDM_HISTO_LENGTH) bin = 0;
  if (bin < 0 || bin >= ASDM_HISTO_LENGTH) bin = 0;  // the reference asserts here
  return bin;
}

// ORBmatcher::ComputeThreeMaxima (ORBmatcher.cc:1584-1625) on bin sizes
__device__ inline void asdm_three_maxima(const int* sizes, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  ind1 = ind2 = ind3 = -1;
  for (int i = 0; i < L; i++) {
    const int s = sizes[i];
    if (s > max1) {
      max3 = max2; max2 = max1; max1 = s;
      ind3 = ind2; ind2 = ind1; ind1 = i;
    } else if (s > max2) {
      max3 = max2; max2 = s;
      ind3 = ind2; ind2 = i;
    } else if (s > max3) {
      max3 = s; ind3 = i;
    }
  }
  if ((float)max2 < __fmul_rn(0.1f, (float)max1)) {
    ind2 = -1; ind3 = -1;
  } else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) {
    ind3 = -1;
  }
}

// --------------------------------------------------------------------------------------- launch wrappers
// tc_shortlist.cu: tcgen05/TMA distance contraction that lists every (pair,i,j) whose bf16 distance estimate
// is within the keep threshold plus the proven error bound.
struct TcWork {   // one unit of the persistent tensor-core kernel
  int32_t pair;
  int32_t n_tile;      // 256-row tile of set B kept resident in shared memory
  int32_t m_tile0;     // first 128-row tile of set A streamed against it
  int32_t m_tiles;
};
// The hit list is segmented: epilogue warp e of CTA c appends to list[(c*EPI_WARPS+e)*seg_cap ..) without atomics and
// leaves its count in seg_count[c*EPI_WARPS+e] (a count above seg_cap means that segment overflowed).
#define ASDM_TC_MAX_EPI_WARPS 8
cudaError_t asdm_tc_shortlist_launch(asdm_ctx* ctx, const PairDesc* d_pairs, const void* d_tmaps, const TcWork* d_work,
                                     int n_work, float t_keep, TcHit* d_list, int seg_cap, int32_t* d_seg_count,
                                     int32_t* d_list_total);
// per-keyframe row bounds for the shortlist: one launch over every 128-row chunk of every keyframe of a job
struct TcSetDesc {
  const float* norm2;
  const uint8_t* mask;
  const int32_t* pos;
  int32_t n;
  int32_t chunk0;  // first chunk of this keyframe in rb / cmin
};
cudaError_t asdm_tc_row_bounds_launch(asdm_ctx* ctx, const TcSetDesc* d_sets, const int32_t* d_chunk_set, int n_chunks,
                                      int mask_sense, float* d_rb, float* d_cmin);
int asdm_tc_build_tmap(asdm_ctx* ctx, asdm_set* set);
