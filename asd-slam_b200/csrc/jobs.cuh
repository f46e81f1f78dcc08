// jobs.cuh -- the batched keyframe-pair job (jobs.cu), shared with the multi-GPU group (group.cu).
#pragma once
#include "host_util.cuh"

struct asdm_job {
  asdm_ctx* ctx = nullptr;
  int kind = 0;
  int n_pairs = 0;
  int n_stride = 0;    // max nA over pairs (query slots per pair)
  int m_stride = 0;    // max nB over pairs
  int out_stride = 0;  // entries per pair in match12 / dist12
  PublishArgs pub{};   // gathered arrays of every rank the resolve kernel also stores into (asdm_job_set_publish)
  BowParams P{};
  std::vector<asdm_set*> sets;  // retained
  std::vector<PairDesc> h_pairs;
  DevBuf d_pairs, d_masks, d_cnt, d_cand, d_owner, d_acc, d_flags, d_match, d_dist, d_nm;
  DevBuf d_work, d_tmaps, d_list, d_seg_count, d_rb, d_cmin, d_setdesc;
  // d_masks / d_pairs / d_work / d_tmaps / d_setdesc are views into d_in (every input of the job, brought over by ONE
  // pull); d_flags / d_seg_count / d_cnt are views into d_zero (everything a run has to clear, ONE fill)
  DevBuf d_in, d_zero;
  size_t zero_bytes = 0;
  int cap = 8;
  int use_tc = 0;
  int resolve_level = 0;  // k_resolve_bow record budget: 0 = 100 KB, 1 = 200 KB, 2 = global-memory kernel
  int n_work = 0;
  int list_cap = 0;  // total hit slots; split evenly into n_seg segments of seg_cap
  int n_seg = 0, seg_cap = 0;
  float t_keep_tc = 0.f;
  bool ran = false;
  void* prefetched[3] = {nullptr, nullptr, nullptr};  // destinations of asdm_job_fetch_async
  bool prefetch_valid = false;
  bool tc_pending = false;     // tensor-core route taken before every set's range flag was back (validated in settle)
  int32_t* h_flags = nullptr;  // pinned, 8 ints: [0] cand overflow, [1] list overflow, [2] list count
  bool h_flags_owned = false;  // true when the context's slot slab was exhausted and the job allocated its own
};

void job_free(asdm_job* j);
int job_alloc_scratch(asdm_job* j);
int job_prepare(asdm_ctx* ctx, const asdm_job_desc* d, int kind_internal, asdm_job** out);
int job_run(asdm_job* j);
int job_settle(asdm_job* j);
