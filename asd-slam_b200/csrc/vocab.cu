// vocab.cu -- DBoW2 vocabulary tree on the device.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"
#include "kernels_vocab.cuh"

// ============================================================================ vocabulary tree
extern "C" int asdm_vocab_create(asdm_ctx* ctx, const asdm_vocab_desc* d, asdm_vocab** out) {
  if (!ctx || !d || !out) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  *out = nullptr;
  if (d->n_nodes < 2 || !d->child_off || !d->child_idx || !d->desc || !d->weight || !d->word_id)
    return fail(ctx, ASDM_ERR_INVALID, "vocabulary is empty or incomplete");
  if (d->weighting < 0 || d->weighting > 3 || d->scoring < 0 || d->scoring > 5)
    return fail(ctx, ASDM_ERR_INVALID, "unknown weighting / scoring type");
  const int nn = d->n_nodes;
  if (d->child_off[0] != 0 || d->child_off[nn] != nn - 1)
    return fail(ctx, ASDM_ERR_INVALID, "children lists must name every node but the root exactly once");
  if (d->child_off[1] == 0) return fail(ctx, ASDM_ERR_INVALID, "the root has no children");
  std::vector<uint8_t> seen(nn, 0);
  int kmax = 0;
  for (int p = 0; p < nn; p++) {
    const int kp = d->child_off[p + 1] - d->child_off[p];
    if (kp < 0) return fail(ctx, ASDM_ERR_INVALID, "child offsets not ascending");
    kmax = std::max(kmax, kp);
    for (int e = d->child_off[p]; e < d->child_off[p + 1]; e++) {
      const int c = d->child_idx[e];
      if (c <= 0 || c >= nn || seen[c] || c == p) return fail(ctx, ASDM_ERR_INVALID, "children lists do not form a tree");
      seen[c] = 1;
    }
  }
  // a node reachable from itself cannot also be reachable from the root: every node has one parent (checked above),
  // so walking down from the root must visit all n_nodes
  {
    std::vector<int> stack{0};
    int visited = 0;
    while (!stack.empty()) {
      const int p = stack.back();
      stack.pop_back();
      visited++;
      for (int e = d->child_off[p]; e < d->child_off[p + 1]; e++) stack.push_back(d->child_idx[e]);
    }
    if (visited != nn) return fail(ctx, ASDM_ERR_INVALID, "children lists contain a cycle");
  }
  asdm_vocab* v = new (std::nothrow) asdm_vocab;
  if (!v) return fail(ctx, ASDM_ERR_OOM, "host allocation");
  v->ctx = ctx;
  v->k = d->k; v->L = d->L; v->weighting = d->weighting; v->scoring = d->scoring; v->n_nodes = nn;
  v->G = std::min(std::max(kmax, 1), 32);
  v->kmax = kmax;
  v->h_weight.assign(d->weight, d->weight + nn);
  v->h_word.assign(d->word_id, d->word_id + nn);
  // one descriptor row and one record per child slot, built on the host once (vocabulary load time), one upload
  const size_t ns = (size_t)nn - 1, o_meta = 0, o_rows = align_up(ns * sizeof(int4), 512), total = o_rows + ns * ASDM_DESC_DIM * 4;
  std::vector<char> hb(total, 0);
  int4* meta = (int4*)(hb.data() + o_meta);
  float* rows = (float*)(hb.data() + o_rows);
  for (size_t sl = 0; sl < ns; sl++) {
    const int c = d->child_idx[sl];
    meta[sl] = make_int4(c, d->child_off[c], d->child_off[c + 1] - d->child_off[c], d->weight[c] > 0.0 ? 1 : 0);
    memcpy(rows + sl * ASDM_DESC_DIM, d->desc + (size_t)c * ASDM_DESC_DIM, ASDM_DESC_DIM * 4);
  }
  cudaError_t e = dev_alloc(ctx, &v->mem, total);
  if (e == cudaSuccess) e = cudaMemcpyAsync(v->mem, hb.data(), total, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // hb is pageable and dies with this scope
  if (e != cudaSuccess) {
    dev_free(ctx, v->mem);
    delete v;
    return fail(ctx, e == cudaErrorMemoryAllocation ? ASDM_ERR_OOM : ASDM_ERR_CUDA, "vocabulary upload", e);
  }
  synced(ctx);
  v->meta = (int4*)((char*)v->mem + o_meta);
  v->rows = (float*)((char*)v->mem + o_rows);
  v->root_k = d->child_off[1];
  *out = v;
  return ASDM_OK;
}

extern "C" int asdm_vocab_release(asdm_vocab* v) {
  if (!v) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = v->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  cudaSetDevice(ctx->device);
  dev_free(ctx, v->mem);
  delete v;
  return ASDM_OK;
}

extern "C" int asdm_vocab_transform(asdm_ctx* ctx, asdm_vocab* voc, asdm_set* s, int levelsup, int32_t* word_id,
                                    int32_t* node_id, uint8_t* in_fv, int32_t* bow_n, uint32_t* bow_id, double* bow_val) {
  if (!ctx || !voc || !s || voc->ctx != ctx || s->ctx != ctx) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  if (bow_n) *bow_n = 0;
  const int n = s->n;
  if (!s->node_id) {
    CK(dev_alloc(ctx, &s->fv_mem, (size_t)s->n_pad * 8));
    s->node_id = (uint32_t*)s->fv_mem;
    s->fv_pos = (int32_t*)((char*)s->fv_mem + (size_t)s->n_pad * 4);
  }
  // padding rows: node 0, not in the FeatureVector (as asdm_set_featvec leaves them)
  CK(cudaMemsetAsync(s->node_id, 0, (size_t)s->n_pad * 4, ctx->stream));
  CK(cudaMemsetAsync(s->fv_pos, 0xff, (size_t)s->n_pad * 4, ctx->stream));
  s->has_fv = true;
  if (n == 0) return ASDM_OK;
  int rc;
  const size_t o_key = 0, o_leaf = align_up((size_t)n * 8, 256), o_cnt = o_leaf + align_up((size_t)n * 4, 256), sc_bytes = o_cnt + 256;
  if ((rc = ensure(ctx, ctx->scratch[SC_VOC], sc_bytes))) return rc;
  unsigned long long* d_key = (unsigned long long*)((char*)ctx->scratch[SC_VOC].p + o_key);
  int32_t* d_leaf = (int32_t*)((char*)ctx->scratch[SC_VOC].p + o_leaf);
  int32_t* d_cnt = (int32_t*)((char*)ctx->scratch[SC_VOC].p + o_cnt);
  VocDev vd{voc->rows, voc->meta, voc->root_k, voc->G};
  const int fpw = 32 / voc->G;
  bool staged = false;
  if (voc->kmax <= 32 && !ctx->opt_voc_direct) {  // child blocks staged through shared memory (cp.async)
    const int per_cta = VOC_WARPS * fpw;
    const size_t smem = (size_t)VOC_WARPS * voc_warp_words(voc->G) * sizeof(float);
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_voc_descend_staged, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // several CTAs per SM must fit (the descent is latency bound): ask for the largest shared-memory carve-out
    CK(cudaFuncSetAttribute(k_voc_descend_staged, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaMemsetAsync(d_cnt, 0, 4, ctx->stream));
    k_voc_descend_staged<<<(n + per_cta - 1) / per_cta, VOC_WARPS * 32, smem, ctx->stream>>>(
        vd, s->desc, n, voc->L - levelsup, d_leaf, s->node_id, s->fv_pos, d_key, ctx->opt_voc_exact, d_cnt);
    staged = true;
  } else {
    const int per_cta = VOC_WARPS_DIRECT * fpw;
    k_voc_descend<<<(n + per_cta - 1) / per_cta, VOC_WARPS_DIRECT * 32, (size_t)per_cta * ASDM_DESC_DIM * 4, ctx->stream>>>(
        vd, s->desc, n, voc->L - levelsup, d_leaf, s->node_id, s->fv_pos, d_key);
  }
  LAUNCHED(ctx);
  k_fv_rank<<<dim3((n + 127) / 128, (n + FV_JT - 1) / FV_JT), 128, 0, ctx->stream>>>(d_key, n, s->fv_pos);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  {  // the FeatureVector arrays were rewritten on this stream: what the other contexts of the family wait for
    const int rr = set_mark_ready(ctx, s, ctx->stream);
    if (rr) return rr;
  }
  const bool want_host = word_id || node_id || in_fv || bow_n || bow_id || bow_val;
  if (!want_host) return ASDM_OK;  // stays asynchronous: the searches that follow are ordered on the same stream
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, (size_t)n * 8 + 4, &h))) return rc;
  int32_t* h_leaf = (int32_t*)h;
  uint32_t* h_nid = (uint32_t*)h + n;
  int32_t* h_cnt = (int32_t*)h + 2 * (size_t)n;
  *h_cnt = 0;
  CK(cudaMemcpyAsync(h_leaf, d_leaf, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(h_nid, s->node_id, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (staged) CK(cudaMemcpyAsync(h_cnt, d_cnt, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->stat_voc_exact = *h_cnt;
  // marshalling for the caller's std::map containers: per-feature ids, then the BowVector in map (ascending word) order
  std::vector<std::pair<uint32_t, double>> words;  // (word, weight) of the features that count, in feature order
  words.reserve(n);
  for (int i = 0; i < n; i++) {
    const int lf = h_leaf[i];
    const double w = voc->h_weight[lf];
    if (word_id) word_id[i] = voc->h_word[lf];
    if (node_id) node_id[i] = (int32_t)h_nid[i];
    if (in_fv) in_fv[i] = w > 0 ? 1 : 0;
    if (w > 0) words.emplace_back((uint32_t)voc->h_word[lf], w);
  }
  synced(ctx);
  if (!(bow_n || bow_id || bow_val)) return ASDM_OK;
  std::stable_sort(words.begin(), words.end(),
                   [](const std::pair<uint32_t, double>& a, const std::pair<uint32_t, double>& b) { return a.first < b.first; });
  std::vector<uint32_t> ids;
  std::vector<double> vals;
  const bool tf = voc->weighting == 0 || voc->weighting == 1;
  for (size_t e = 0; e < words.size(); e++) {
    if (e > 0 && words[e].first == words[e - 1].first) {
      if (tf) vals.back() += words[e].second;  // BowVector::addWeight (BowVector.cpp:35-47); addIfNotExist keeps the first
    } else {
      ids.push_back(words[e].first);
      vals.push_back(words[e].second);
    }
  }
  const bool must = voc->scoring != 5;  // DotProductScoring alone does not normalise (ScoringObject.h:72-89)
  const bool l2 = voc->scoring == 1;
  if (tf && !ids.empty() && !must) {
    const double nd = (double)ids.size();
    for (double& x : vals) x /= nd;  // :1163-1169
  }
  if (must) {  // BowVector::normalize (BowVector.cpp:61-84)
    double norm = 0.0;
    if (!l2) {
      for (double x : vals) norm += fabs(x);
    } else {
      for (double x : vals) norm += x * x;
      norm = sqrt(norm);
    }
    if (norm > 0.0)
      for (double& x : vals) x /= norm;
  }
  if (bow_n) *bow_n = (int32_t)ids.size();
  if (bow_id) memcpy(bow_id, ids.data(), ids.size() * 4);
  if (bow_val) memcpy(bow_val, vals.data(), vals.size() * 8);
  return ASDM_OK;
}

