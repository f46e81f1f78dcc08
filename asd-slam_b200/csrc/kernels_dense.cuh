// kernels_dense.cuh -- BoW-gated dense search kernels (SearchByBoW x2, SearchForTriangulation).
//
//  k_cand_exact   : every gated (i,j) pair in exact fp32 (reference arithmetic), keeps d <= t_keep
//  k_rerank       : exact fp32 distance of the (i,j) pairs short-listed by the tcgen05 kernel
//  k_resolve_bow  : replays the reference's greedy query order as a parallel fixed point, ratio test,
//                   30-bin rotation histogram + ComputeThreeMaxima, writes the match lists
//  k_resolve_tri  : SearchForTriangulation selection (epipole + epipolar-line gates, last tie wins)
#pragma once
#include "internal.cuh"

// gate shared by all candidate producers: both features usable and in the same vocabulary node
__device__ __forceinline__ bool bow_query_ok(const PairDesc& pd, const BowParams& P, int i) {
  return i >= pd.rowA0 && i < pd.rowA1 && pd.maskA[i] == P.mask_sense && pd.posA[i] >= 0;
}
__device__ __forceinline__ bool bow_target_ok(const PairDesc& pd, const BowParams& P, int j) {
  return pd.maskB[j] == P.mask_sense && pd.posB[j] >= 0;
}

__device__ __forceinline__ int pair_slot(int pair, int n_stride, int i) { return pair * n_stride + i; }

// What the resolve kernels need of a listed pair besides (j, d): the target's scan position, the query's position in the
// reference's loop and the rotation bin of the pair -- gathered here, by the thread that already holds i and j, so that
// the resolve kernel never chases posA / posB / keypoint angles through global memory.
__device__ __forceinline__ int cand_meta(const PairDesc& pd, const BowParams& P, int i, int j) {
  const int bin = P.check_ori ? asdm_rot_bin(pd.kpA[i].angle, pd.kpB[j].angle) : 0;  // e.g. ORBmatcher.cc:245-250
  return (pd.posA[i] << 5) | bin;
}
__device__ __forceinline__ void cand_append(int32_t* cnt, Cand* cand, int cap, int qslot, int j, float d, int pos, int meta,
                                            int32_t* overflow) {
  const int s = atomicAdd(cnt + qslot, 1);
  if (s < cap) {
    *reinterpret_cast<int4*>(cand + (size_t)qslot * cap + s) = make_int4(j, __float_as_int(d), pos, meta);
  } else {
    *overflow = 1;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Exact dense candidates. Tile 64 queries x 64 targets per CTA, 256 threads, 4x4 pairs per thread; both
// tiles are staged K-transposed in shared memory so the inner loop is two LDS.128 per 48 fp32 ops. Every
// pair keeps its own accumulator and walks k = 0..127 in order: bit-identical to ORBmatcher.cc:1637-1644.
// Algorithmic work: 3*128 flops per pair (sub, mul, add; no FMA).
#define CE_TM 64
#define CE_TN 64
__global__ void __launch_bounds__(256) k_cand_exact(const PairDesc* __restrict__ pairs, BowParams P, int cap,
                                                    int n_stride, int32_t* __restrict__ cnt,
                                                    Cand* __restrict__ cand, int32_t* __restrict__ overflow) {
  extern __shared__ __align__(16) float sm[];
  float* As = sm;                          // [128][CE_TM]
  float* Bs = sm + ASDM_DESC_DIM * CE_TM;  // [128][CE_TN]
  const int pair = blockIdx.z;
  const PairDesc& pd = pairs[pair];
  const int i0 = blockIdx.x * CE_TM, j0 = blockIdx.y * CE_TN;
  if (i0 >= pd.rowA1 || i0 + CE_TM <= pd.rowA0 || j0 >= pd.nB) return;
  const int t = threadIdx.x;
  {
    const int row = t & 63, kc = t >> 6;  // consecutive threads take consecutive rows: conflict-free smem stores
    const bool okA = (i0 + row) < pd.nA, okB = (j0 + row) < pd.nB;
    const float4* ga = reinterpret_cast<const float4*>(pd.descA + (size_t)(i0 + row) * ASDM_DESC_DIM);
    const float4* gb = reinterpret_cast<const float4*>(pd.descB + (size_t)(j0 + row) * ASDM_DESC_DIM);
#pragma unroll
    for (int it = 0; it < 8; it++) {
      const int k4 = kc + 4 * it;
      const float4 va = okA ? __ldg(ga + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 vb = okB ? __ldg(gb + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
      As[(k4 * 4 + 0) * CE_TM + row] = va.x; As[(k4 * 4 + 1) * CE_TM + row] = va.y;
      As[(k4 * 4 + 2) * CE_TM + row] = va.z; As[(k4 * 4 + 3) * CE_TM + row] = va.w;
      Bs[(k4 * 4 + 0) * CE_TN + row] = vb.x; Bs[(k4 * 4 + 1) * CE_TN + row] = vb.y;
      Bs[(k4 * 4 + 2) * CE_TN + row] = vb.z; Bs[(k4 * 4 + 3) * CE_TN + row] = vb.w;
    }
  }
  __syncthreads();
  const int ty = t >> 4, tx = t & 15;
  float acc[4][4];
#pragma unroll
  for (int r = 0; r < 4; r++)
#pragma unroll
    for (int c = 0; c < 4; c++) acc[r][c] = 0.f;
#pragma unroll 4
  for (int k = 0; k < ASDM_DESC_DIM; k++) {
    const float4 a = *reinterpret_cast<const float4*>(As + k * CE_TM + ty * 4);
    const float4 b = *reinterpret_cast<const float4*>(Bs + k * CE_TN + tx * 4);
    const float av[4] = {a.x, a.y, a.z, a.w};
    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const float d = __fsub_rn(av[r], bv[c]);
        acc[r][c] = __fadd_rn(acc[r][c], __fmul_rn(d, d));
      }
  }
#pragma unroll
  for (int r = 0; r < 4; r++) {
    const int i = i0 + ty * 4 + r;
    if (i >= pd.nA) continue;
#pragma unroll
    for (int c = 0; c < 4; c++) {
      const int j = j0 + tx * 4 + c;
      if (j >= pd.nB) continue;
      if (!(acc[r][c] <= P.t_keep)) continue;
      if (!bow_query_ok(pd, P, i) || !bow_target_ok(pd, P, j)) continue;
      if (pd.nodeA[i] != pd.nodeB[j]) continue;
      cand_append(cnt, cand, cap, pair_slot(pair, n_stride, i), j, acc[r][c], pd.posB[j], cand_meta(pd, P, i, j), overflow);
    }
  }
}

// Exact rerank of the tensor-core short list: one CTA per list segment, warp-cooperative. A warp takes RR_BATCH hits at
// a time: lanes 0..7 read the hit records and evaluate the gates, then for every live hit the 32 lanes load its two
// 512-byte rows with ONE coalesced load each (lane l owns floats 4l..4l+3), square the differences and park them in
// shared memory; lane c finally adds hit c's 128 terms in index order with one fp32 accumulator -- the reference's
// summation (ORBmatcher.cc:1637-1644) to the bit, without a warp ever touching 32 different rows in one instruction.
// flags[1] is raised when a segment overflowed (the host grows the list and reruns the job).
#define RR_WARPS 8
#define RR_BATCH 8
#define RR_ROW 132
#define RR_MAX_SEG 2048
// Persistent and balanced: the hit list is segmented by (CTA, epilogue warp) of the shortlist kernel and the segments are
// far from equal, so every CTA first turns the segment counts into a prefix sum (shared memory) and the warps of the whole
// grid then take batches of RR_BATCH consecutive hits of the FLATTENED list, round-robin; a hit's segment is found by
// binary search in the prefix.
__global__ void __launch_bounds__(RR_WARPS * 32, 4) k_rerank(const PairDesc* __restrict__ pairs, BowParams P, int cap,
                                                          int n_stride, const TcHit* __restrict__ list,
                                                          const int32_t* __restrict__ seg_count, int n_seg, int seg_cap,
                                                          int32_t* __restrict__ cnt, Cand* __restrict__ cand,
                                                          int32_t* __restrict__ flags) {
  __shared__ __align__(16) float s_sq[RR_WARPS][RR_BATCH * RR_ROW];
  __shared__ int s_prefix[RR_MAX_SEG + 1];
  __shared__ int s_warp_tot[RR_WARPS];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  // inclusive scan of the (clamped) segment counts in ONE pass: thread t owns `per` consecutive segments (<= 8), adds them
  // up, the block scans the 256 partial sums (shuffles + one hop through shared memory) and the thread writes its
  // segments' prefixes. (Round by round, 256 segments at a time, this prologue was half of the kernel's instructions on a
  // short list -- every one of the 592 CTAs runs it.)
  {
    const int per = (n_seg + RR_WARPS * 32 - 1) / (RR_WARPS * 32);
    const int s0 = t * per;
    int v[RR_MAX_SEG / (RR_WARPS * 32)];
    int sum = 0;
#pragma unroll
    for (int k = 0; k < RR_MAX_SEG / (RR_WARPS * 32); k++) {
      v[k] = 0;
      if (k < per && s0 + k < n_seg) {
        int c = seg_count[s0 + k];
        if (c > seg_cap) {
          flags[1] = 1;
          c = seg_cap;
        }
        v[k] = c;
      }
      sum += v[k];
    }
    int x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp_tot[warp] = x;
    __syncthreads();
    int run = x - sum;  // exclusive prefix of this thread within the block
#pragma unroll
    for (int w = 0; w < RR_WARPS; w++) run += w < warp ? s_warp_tot[w] : 0;
#pragma unroll
    for (int k = 0; k < RR_MAX_SEG / (RR_WARPS * 32); k++)
      if (k < per && s0 + k < n_seg) {
        run += v[k];
        s_prefix[s0 + k + 1] = run;
      }
  }
  if (t == 0) s_prefix[0] = 0;
  __syncthreads();
  const int total = s_prefix[n_seg];
  float* sq = s_sq[warp];
  const int n_batches = (total + RR_BATCH - 1) / RR_BATCH;
  for (int b = blockIdx.x * RR_WARPS + warp; b < n_batches; b += gridDim.x * RR_WARPS) {
    const int f0 = b * RR_BATCH, nb = min(RR_BATCH, total - f0);
    // lanes 0..nb-1: one hit each -- where it lives, its record, the gates, the addresses of its two rows
    int hp = 0, hi = 0, hj = 0, hpos = 0;
    unsigned long long pa = 0, pb = 0;
    bool live = false;
    if (lane < nb) {
      const int f = f0 + lane;
      int lo = 0, hi_s = n_seg;  // largest s with prefix[s] <= f
      while (hi_s - lo > 1) {
        const int mid = (lo + hi_s) >> 1;
        if (s_prefix[mid] <= f) lo = mid;
        else hi_s = mid;
      }
      const TcHit h = list[(size_t)lo * seg_cap + (f - s_prefix[lo])];
      hp = h.pair; hi = h.i; hj = h.j;
      const PairDesc& pd = pairs[hp];
      if (hi < pd.nA && hj < pd.nB && bow_query_ok(pd, P, hi) && pd.maskB[hj] == P.mask_sense) {
        hpos = pd.posB[hj];
        live = hpos >= 0 && pd.nodeA[hi] == pd.nodeB[hj];
      }
      pa = reinterpret_cast<unsigned long long>(pd.descA + (size_t)hi * ASDM_DESC_DIM);
      pb = reinterpret_cast<unsigned long long>(pd.descB + (size_t)hj * ASDM_DESC_DIM);
    }
    const unsigned lives = __ballot_sync(0xffffffffu, live);
    if (!lives) continue;
    // the row loads go out four hits at a time before the first one is used: 8 x 512 B in flight per warp (all 16 at once
    // costs 96 registers and two resident CTAs less per SM -- measured slower)
#pragma unroll
    for (int h0 = 0; h0 < RR_BATCH; h0 += 4) {
      float4 a4[4], b4[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int c = h0 + u;
        const unsigned long long qa = __shfl_sync(0xffffffffu, pa, c), qb = __shfl_sync(0xffffffffu, pb, c);
        if ((lives >> c) & 1u) {
          a4[u] = __ldg(reinterpret_cast<const float4*>(qa) + lane);
          b4[u] = __ldg(reinterpret_cast<const float4*>(qb) + lane);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int c = h0 + u;
        if (!((lives >> c) & 1u)) continue;
        float4 s;
        float d;
        d = __fsub_rn(a4[u].x, b4[u].x); s.x = __fmul_rn(d, d);
        d = __fsub_rn(a4[u].y, b4[u].y); s.y = __fmul_rn(d, d);
        d = __fsub_rn(a4[u].z, b4[u].z); s.z = __fmul_rn(d, d);
        d = __fsub_rn(a4[u].w, b4[u].w); s.w = __fmul_rn(d, d);
        *reinterpret_cast<float4*>(sq + c * RR_ROW + 4 * lane) = s;
      }
    }
    __syncwarp();
    if (live) {
      const float4* row = reinterpret_cast<const float4*>(sq + lane * RR_ROW);
      float acc = 0.f;
#pragma unroll 8
      for (int k = 0; k < ASDM_DESC_DIM / 4; k++) {
        const float4 x = row[k];
        acc = __fadd_rn(acc, x.x); acc = __fadd_rn(acc, x.y); acc = __fadd_rn(acc, x.z); acc = __fadd_rn(acc, x.w);
      }
      if (acc <= P.t_keep) cand_append(cnt, cand, cap, pair_slot(hp, n_stride, hi), hj, acc, hpos, cand_meta(pairs[hp], P, hi, hj), flags);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------
// Greedy resolve for SearchByBoW (ORBmatcher.cc:156-297 and :533-666). One CTA per keyframe pair.
//
// The reference walks queries in FeatureVector order; an accepted query removes its target from every later
// query's scan (vpMapPointMatches[realIdxF] / vbMatched2). Let accept(i) be the decision of query i given
// the targets accepted by queries that precede it. The sequential run is the unique solution of
//   accept(i) = D_i({accept(i') : pos(i') < pos(i)}).
// We iterate all queries in parallel against owner[j] = min pos of a query currently accepting j; after
// iteration t the first t queries are final, and in practice two or three sweeps reach the fixed point.
// Only candidates with d <= t_keep are listed: a farther second-best cannot fail the ratio test
// (t_keep is chosen on the host so that nnratio * d > TH_LOW for every d > t_keep).
__device__ __forceinline__ bool cand_before(const Cand& x, int px, const Cand& y, int py) {
  return x.d < y.d || (x.d == y.d && px < py);
}

// Epilogue of the resolve kernels when the job publishes (PublishArgs): the CTA's finished row -- out_n entries, the
// rest of the gathered stride padded with -1 -- goes to the same slot of every rank's gathered arrays. Peer stores
// travel over NVLink while the other CTAs are still resolving their pairs. 16-byte stores when the rows allow it.
// Multicast destinations (an NVSwitch multicast mapping that every rank's array is bound to) take multimem.st: the
// row leaves this GPU once and the switch replicates it, so the egress is 1/N of the unicast version's.
__device__ __forceinline__ void pub_store4(void* p, int4 v, bool mc) {
  if (mc)
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(__int_as_float(v.x)),
                 "f"(__int_as_float(v.y)), "f"(__int_as_float(v.z)), "f"(__int_as_float(v.w))
                 : "memory");
  else
    *reinterpret_cast<int4*>(p) = v;
}
__device__ __forceinline__ void pub_store1(void* p, int v, bool mc) {
  if (mc)
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
  else
    *reinterpret_cast<int*>(p) = v;
}

__device__ __forceinline__ void publish_row(const PublishArgs& pub, int pair, const int32_t* __restrict__ matchp,
                                            const float* __restrict__ distp, int out_n, int nm, int c0 = 0, int c1 = 0) {
  const int t = threadIdx.x, nt = blockDim.x;
  const size_t slot = (size_t)pub.row0 + (size_t)pair * pub.row_step;
  const size_t row = pub.row_map ? (size_t)pub.row_map[pair] : slot;
  const bool mc = pub.multicast != 0;
  const bool wd = pub.no_dist == 0;
  // columns [lo, hi) of the destination row belong to this unit: the whole row (padded to the gathered stride with -1),
  // or the unit's own query rows when a pair is split into row blocks
  const int lo = pub.col_range ? c0 : 0, hi = pub.col_range ? c1 : pub.stride;
  const bool vec = (pub.stride & 3) == 0 && (lo & 3) == 0 &&
                   ((reinterpret_cast<uintptr_t>(matchp) | reinterpret_cast<uintptr_t>(distp)) & 15) == 0;
  if (vec) {
    const int hi4 = hi & ~3;
    for (int e = lo + 4 * t; e < hi4; e += 4 * nt) {
      int4 m = make_int4(-1, -1, -1, -1);
      float4 d = make_float4(-1.f, -1.f, -1.f, -1.f);
      if (e + 3 < out_n) {
        m = *reinterpret_cast<const int4*>(matchp + e);
        d = *reinterpret_cast<const float4*>(distp + e);
      } else if (e < out_n) {
        m.x = matchp[e]; d.x = distp[e];
        if (e + 1 < out_n) { m.y = matchp[e + 1]; d.y = distp[e + 1]; }
        if (e + 2 < out_n) { m.z = matchp[e + 2]; d.z = distp[e + 2]; }
      }
      const int4 di = make_int4(__float_as_int(d.x), __float_as_int(d.y), __float_as_int(d.z), __float_as_int(d.w));
      for (int p = 0; p < pub.n; p++) {
        pub_store4(pub.match[p] + row * pub.stride + e, m, mc);
        if (wd && ((pub.dist_mask >> p) & 1u)) pub_store4(pub.dist[p] + row * pub.stride + e, di, mc);
      }
    }
    for (int k = hi4 + t; k < hi; k += nt) {  // a block that ends off a 16-byte boundary (the last block of a pair)
      const int m = k < out_n ? matchp[k] : -1;
      const float d = k < out_n ? distp[k] : -1.f;
      for (int p = 0; p < pub.n; p++) {
        pub_store1(pub.match[p] + row * pub.stride + k, m, mc);
        if (wd && ((pub.dist_mask >> p) & 1u)) pub_store1(pub.dist[p] + row * pub.stride + k, __float_as_int(d), mc);
      }
    }
  } else {
    for (int k = lo + t; k < hi; k += nt) {
      const int m = k < out_n ? matchp[k] : -1;
      const float d = k < out_n ? distp[k] : -1.f;
      for (int p = 0; p < pub.n; p++) {
        pub_store1(pub.match[p] + row * pub.stride + k, m, mc);
        if (wd && ((pub.dist_mask >> p) & 1u)) pub_store1(pub.dist[p] + row * pub.stride + k, __float_as_int(d), mc);
      }
    }
  }
  if (t == 0)
    for (int p = 0; p < pub.n; p++) pub_store1(pub.nm[p] + slot, nm, mc);
  // No system fence here (there was one: 512 MEMBAR.SYS + ERRBAR per CTA, 8 us of a 30 us kernel). Nobody reads these rows
  // before this KERNEL has completed: the host after a stream synchronisation, the peers behind the group barrier, whose
  // signal is a later kernel of this stream that fences (k_group_signal) and stores with release.sys semantics.
}

__global__ void __launch_bounds__(512) k_resolve_bow_global(const PairDesc* __restrict__ pairs, BowParams P, int cap,
                                                     int n_stride, int m_stride, int out_stride,
                                                     const int32_t* __restrict__ cnt, Cand* __restrict__ cand,
                                                     int32_t* __restrict__ owner, int32_t* __restrict__ acc,
                                                     int32_t* __restrict__ match, float* __restrict__ dist,
                                                     int32_t* __restrict__ nmatches, const PublishArgs pub) {
  const int pair = blockIdx.x;
  const PairDesc& pd = pairs[pair];
  const int nA = pd.nA, nB = pd.nB;
  const int32_t* cntp = cnt + (size_t)pair * n_stride;
  Cand* candp = cand + (size_t)pair * n_stride * cap;
  int32_t* own0 = owner + (size_t)pair * 2 * m_stride;
  int32_t* accp = acc + (size_t)pair * n_stride;
  int32_t* matchp = match + (size_t)pair * out_stride;
  float* distp = dist + (size_t)pair * out_stride;
  __shared__ int s_changed;
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  __shared__ int s_count;
  const int t = threadIdx.x, nt = blockDim.x;

  // sort each query's list by (distance, target scan position): strict '<' in the reference means the first
  // candidate in scan order wins ties (ORBmatcher.cc:219-228)
  for (int i = t; i < nA; i += nt) {
    const int n = min(cntp[i], cap);
    Cand* L = candp + (size_t)i * cap;
    for (int a = 1; a < n; a++) {
      const Cand x = L[a];
      const int px = pd.posB[x.j];
      int b = a - 1;
      while (b >= 0 && cand_before(x, px, L[b], pd.posB[L[b].j])) {
        L[b + 1] = L[b];
        b--;
      }
      L[b + 1] = x;
    }
    accp[i] = -1;
  }
  const int out_n = (P.kind == KIND_BOW_KF_F) ? nB : nA;
  for (int k = t; k < out_n; k += nt) {
    matchp[k] = -1;
    distp[k] = -1.f;
  }
  for (int j = t; j < nB; j += nt) own0[j] = INT_MAX;
  if (t < ASDM_HISTO_LENGTH) s_hist[t] = 0;
  if (t == 0) s_count = 0;
  __syncthreads();

  for (int iter = 0;; iter++) {
    const int32_t* own = own0 + (size_t)(iter & 1) * m_stride;
    int32_t* own_next = own0 + (size_t)((iter & 1) ^ 1) * m_stride;
    if (t == 0) s_changed = 0;
    __syncthreads();
    for (int i = t; i < nA; i += nt) {
      const int n = min(cntp[i], cap);
      if (n == 0) continue;
      const Cand* L = candp + (size_t)i * cap;
      const int mypos = pd.posA[i];
      float d1 = ASDM_BEST_INIT_BOW, d2 = ASDM_BEST_INIT_BOW;
      int j1 = -1;
      for (int c = 0; c < n; c++) {
        const Cand x = L[c];
        if (!(x.d < ASDM_BEST_INIT_BOW)) break;
        if (own[x.j] < mypos) continue;  // taken by an earlier query
        if (j1 < 0) {
          j1 = x.j;
          d1 = x.d;
        } else {
          d2 = x.d;
          break;
        }
      }
      const bool pass = P.strict ? (d1 < P.th_accept) : (d1 <= P.th_accept);
      const int dec = (j1 >= 0 && pass && d1 < __fmul_rn(P.nnratio, d2)) ? j1 : -1;
      if (dec != accp[i]) {
        accp[i] = dec;
        s_changed = 1;
      }
    }
    for (int j = t; j < nB; j += nt) own_next[j] = INT_MAX;
    __syncthreads();
    for (int i = t; i < nA; i += nt) {
      const int a = accp[i];
      if (a >= 0) atomicMin(own_next + a, pd.posA[i]);
    }
    __syncthreads();
    if (!s_changed) break;
    __syncthreads();
  }

  // rotation consistency (ORBmatcher.cc:242-253, 274-294 / :616-626, 645-663)
  for (int i = t; i < nA; i += nt) {
    const int j = accp[i];
    if (j < 0) continue;
    atomicAdd(&s_count, 1);
    if (P.check_ori) atomicAdd(&s_hist[asdm_rot_bin(pd.kpA[i].angle, pd.kpB[j].angle)], 1);
  }
  __syncthreads();
  if (t == 0) {
    if (P.check_ori) {
      asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
      int pruned = 0;
      for (int b = 0; b < ASDM_HISTO_LENGTH; b++)
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) pruned += s_hist[b];
      s_count -= pruned;
    }
    nmatches[pair] = s_count;
  }
  __syncthreads();
  for (int i = t; i < nA; i += nt) {
    const int j = accp[i];
    if (j < 0) continue;
    if (P.check_ori) {
      const int b = asdm_rot_bin(pd.kpA[i].angle, pd.kpB[j].angle);
      if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) continue;
    }
    const Cand* L = candp + (size_t)i * cap;
    const int n = min(cntp[i], cap);
    float d = -1.f;
    for (int c = 0; c < n; c++)
      if (L[c].j == j) { d = L[c].d; break; }
    if (P.kind == KIND_BOW_KF_F) {
      matchp[j] = i;
      distp[j] = d;
    } else {
      matchp[i] = j;
      distp[i] = d;
    }
  }
  if (pub.n) {
    __syncthreads();  // the row is complete (block-wide) before it is copied out
    publish_row(pub, pair, matchp, distp, out_n, s_count);
  }
}

// ---------------------------------------------------------------------------------------------------------
// The same greedy resolve with its working set in shared memory (one CTA per keyframe pair): nine queries in ten have no
// candidate at all, so the CTA first compacts the queries that do into 48-byte records -- position in the reference's
// query order, candidate count and the RB_TOPK nearest candidates ranked as the reference's scan ranks them (distance,
// then FeatureVector position of the target) -- and replays the greedy order over those records only, against owner[]
// in shared memory. A query needs its best and second-best candidate among those not taken by earlier queries, so the
// short list decides unless (all but one of) its entries are taken; then the full candidate slots are scanned again.
// k_resolve_bow_global (above) remains for pairs whose records do not fit (rb_fits()).
#define RB_TOPK 4
struct BowRec {
  int32_t i, pos;   // query feature and its position in the reference's loop (FeatureVector order)
  int32_t n, pad;   // candidates listed
  int2 top[RB_TOPK];  // {target | rotation bin << 24, distance bits}, x = -1: none
};
__host__ __device__ inline size_t rb_smem_bytes(int n_b, int max_act) {
  return (size_t)max_act * (sizeof(BowRec) + 8) + (size_t)n_b * 4 + 64;
}

__global__ void __launch_bounds__(512) k_resolve_bow(const PairDesc* __restrict__ pairs, BowParams P, int cap,
                                                     int n_stride, int m_stride, int out_stride,
                                                     const int32_t* __restrict__ cnt, Cand* __restrict__ cand,
                                                     int32_t* __restrict__ match, float* __restrict__ dist,
                                                     int32_t* __restrict__ nmatches, int max_act,
                                                     int32_t* __restrict__ fallback, const PublishArgs pub) {
  extern __shared__ __align__(16) unsigned char rb_smem[];
  const int pair = blockIdx.x;
  const PairDesc& pd = pairs[pair];
  const int nA = pd.nA, nB = pd.nB;
  const int32_t* cntp = cnt + (size_t)pair * n_stride;
  const Cand* candp = cand + (size_t)pair * n_stride * cap;
  int32_t* matchp = match + (size_t)pair * out_stride;
  float* distp = dist + (size_t)pair * out_stride;
  BowRec* rec = reinterpret_cast<BowRec*>(rb_smem);
  int32_t* accj = reinterpret_cast<int32_t*>(rec + max_act);
  float* accd = reinterpret_cast<float*>(accj + max_act);
  int32_t* own = reinterpret_cast<int32_t*>(accd + max_act);  // [nB] smallest query position currently accepting target j
  __shared__ int s_nact, s_changed, s_count, s_pruned, s_over;
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  const int t = threadIdx.x, nt = blockDim.x, lane = t & 31;
  if (t == 0) { s_nact = 0; s_count = 0; s_pruned = 0; s_over = 0; }
  if (t < ASDM_HISTO_LENGTH) s_hist[t] = 0;
  const int out_n = (P.kind == KIND_BOW_KF_F) ? nB : nA;
  for (int k = t; k < out_n; k += nt) {
    matchp[k] = -1;
    distp[k] = -1.f;
  }
  for (int j = t; j < nB; j += nt) own[j] = INT_MAX;
  __syncthreads();
  // ---- queries with candidates -> records. Pass 1 compacts the queries that have any (one coalesced read of the counts,
  // positions handed out by a ballot per warp); pass 2 spreads those over all threads, each candidate one 16-byte load.
  int32_t* act_i = accj;  // [max_act] scratch: the active query of every slot (accj is initialised afterwards)
  for (int i0 = 0; i0 < nA; i0 += 8 * nt) {
    int c8[8];
#pragma unroll
    for (int u = 0; u < 8; u++) {  // eight independent loads in flight per thread
      const int i = i0 + u * nt + t;
      c8[u] = i < nA ? cntp[i] : 0;
    }
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int i = i0 + u * nt + t;
      const bool has = c8[u] > 0;
      const unsigned m = __ballot_sync(0xffffffffu, has);
      int base = 0;
      if (lane == 0 && m) base = atomicAdd(&s_nact, __popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (has) {
        const int slot = base + __popc(m & ((1u << lane) - 1u));
        if (slot < max_act) act_i[slot] = i;
        else s_over = 1;
      }
    }
  }
  __syncthreads();
  if (!s_over) {
    const int n_all = s_nact;
    for (int a = t; a < n_all; a += nt) {
      const int i = act_i[a], n = min(cntp[i], cap);
      const int4* L = reinterpret_cast<const int4*>(candp + (size_t)i * cap);
      int tj[RB_TOPK], tp[RB_TOPK];
      float td[RB_TOPK];
#pragma unroll
      for (int k = 0; k < RB_TOPK; k++) { tj[k] = -1; tp[k] = INT_MAX; td[k] = 0.f; }
      int pos_i = 0;
      for (int c0 = 0; c0 < n; c0 += 4) {
        int4 x[4];
#pragma unroll
        for (int u = 0; u < 4; u++) x[u] = c0 + u < n ? __ldg(L + c0 + u) : make_int4(-1, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < 4; u++) {
          if (x[u].x < 0) continue;
          pos_i = x[u].w >> 5;
          float dd = __int_as_float(x[u].y);
          int jj = x[u].x | ((x[u].w & 31) << 24), pp = x[u].z;
          bool ins = false;  // strict '<' in the reference: among equal distances the first target in scan order wins
#pragma unroll
          for (int k = 0; k < RB_TOPK; k++) {
            if (ins || tj[k] < 0 || dd < td[k] || (dd == td[k] && pp < tp[k])) {
              const float xd = td[k];
              const int xj = tj[k], xp = tp[k];
              td[k] = dd; tj[k] = jj; tp[k] = pp;
              dd = xd; jj = xj; pp = xp;
              ins = true;
            }
          }
        }
      }
      BowRec R;
      R.i = i; R.pos = pos_i; R.n = n; R.pad = 0;
#pragma unroll
      for (int k = 0; k < RB_TOPK; k++) R.top[k] = make_int2(tj[k], __float_as_int(td[k]));
      rec[a] = R;
    }
  }
  __syncthreads();
  if (!s_over)
    for (int a = t; a < s_nact; a += nt) accj[a] = -1;
  __syncthreads();
  if (s_over) {  // does not fit: the host sees the flag and resolves this job with k_resolve_bow_global
    if (t == 0) *fallback = 1;
    return;
  }
  const int n_act = s_nact;
  for (;;) {
    if (t == 0) s_changed = 0;
    __syncthreads();
    for (int a = t; a < n_act; a += nt) {
      const BowRec R = rec[a];
      unsigned avail = 0;
#pragma unroll
      for (int k = 0; k < RB_TOPK; k++) {
        const int j = R.top[k].x;
        const int o = own[j < 0 ? 0 : (j & 0xffffff)];
        avail |= (j >= 0 && o >= R.pos) ? (1u << k) : 0u;  // taken by an earlier query (vpMapPointMatches / vbMatched2)
      }
      const unsigned rest = avail & (avail - 1u);
      const int k1 = __ffs(avail) - 1, k2 = __ffs(rest) - 1;
      int2 E1 = R.top[0], E2 = R.top[0];
#pragma unroll
      for (int k = 1; k < RB_TOPK; k++) {
        E1 = k1 == k ? R.top[k] : E1;
        E2 = k2 == k ? R.top[k] : E2;
      }
      float d1 = ASDM_BEST_INIT_BOW, d2 = ASDM_BEST_INIT_BOW;
      int j1 = -1;  // target | rotation bin << 24
      if (k1 >= 0) { j1 = E1.x; d1 = __int_as_float(E1.y); }
      if (k2 >= 0) d2 = __int_as_float(E2.y);
      if (k2 < 0 && R.n > RB_TOPK) {
        // fewer than two of the short list are free: the reference's scan over everything listed (:219-228)
        d1 = ASDM_BEST_INIT_BOW; d2 = ASDM_BEST_INIT_BOW; j1 = -1;
        int p1 = INT_MAX, p2 = INT_MAX;
        const Cand* L = candp + (size_t)R.i * cap;
        for (int c = 0; c < R.n; c++) {
          const Cand x = L[c];
          if (own[x.j] < R.pos) continue;
          const int px = x.pos;
          if (x.d < d1 || (x.d == d1 && px < p1)) {
            d2 = d1; p2 = p1;
            d1 = x.d; p1 = px; j1 = x.j | ((x.pad & 31) << 24);
          } else if (x.d < d2 || (x.d == d2 && px < p2)) {
            d2 = x.d; p2 = px;
          }
        }
      }
      const bool pass = P.strict ? (d1 < P.th_accept) : (d1 <= P.th_accept);
      const int dec = (j1 >= 0 && pass && d1 < __fmul_rn(P.nnratio, d2)) ? j1 : -1;
      if (dec != accj[a]) {
        accj[a] = dec;
        s_changed = 1;
      }
      accd[a] = d1;
    }
    __syncthreads();
    for (int j = t; j < nB; j += nt) own[j] = INT_MAX;
    __syncthreads();
    for (int a = t; a < n_act; a += nt) {
      const int j = accj[a];
      if (j >= 0) atomicMin(own + (j & 0xffffff), rec[a].pos);
    }
    __syncthreads();
    const int changed = s_changed;
    __syncthreads();  // everybody has read the flag before thread 0 clears it for the next sweep
    if (!changed) break;
  }
  // rotation consistency (ORBmatcher.cc:242-253, 274-294 / :616-626, 645-663): one shared-memory add per distinct bin
  // of a warp (__match_any_sync), counts by shuffle
  int my_count = 0;
  for (int a0 = 0; a0 < n_act; a0 += nt) {
    const int a = a0 + t;
    int bin = -1;
    if (a < n_act && accj[a] >= 0) {
      my_count++;
      if (P.check_ori) bin = accj[a] >> 24;
    }
    if (P.check_ori) {
      const unsigned peers = __match_any_sync(0xffffffffu, bin);
      if (bin >= 0 && lane == __ffs(peers) - 1) atomicAdd(&s_hist[bin], __popc(peers));
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_count += __shfl_xor_sync(0xffffffffu, my_count, o);
  if (lane == 0 && my_count) atomicAdd(&s_count, my_count);
  __syncthreads();
  if (t == 0) {
    if (P.check_ori) {
      asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
      int pruned = 0;
      for (int b = 0; b < ASDM_HISTO_LENGTH; b++)
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) pruned += s_hist[b];
      s_pruned = pruned;
    }
    nmatches[pair] = s_count - s_pruned;
  }
  __syncthreads();
  for (int a = t; a < n_act; a += nt) {
    if (accj[a] < 0) continue;
    const int j = accj[a] & 0xffffff, i = rec[a].i;
    if (P.check_ori) {
      const int b = accj[a] >> 24;
      if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) continue;
    }
    if (P.kind == KIND_BOW_KF_F) {
      matchp[j] = i;
      distp[j] = accd[a];
    } else {
      matchp[i] = j;
      distp[i] = accd[a];
    }
  }
  if (pub.n) {
    __syncthreads();  // the row is complete (block-wide) before it is copied out
    publish_row(pub, pair, matchp, distp, out_n, s_count - s_pruned);
  }
}

// ---------------------------------------------------------------------------------------------------------
// SearchForTriangulation selection (ORBmatcher.cc:718-774). No cross-query state (vbMatched2 is never set),
// so each query is independent: among listed candidates (d <= TH_LOW) that pass the epipole gate (:743-747)
// and CheckDistEpipolarLine (:136-153), the running "dist > bestDist -> continue" keeps the LAST candidate in
// scan order among equal minima.
// scale_factors / level_sigma2: the level tables of BowParams, staged in shared memory by the caller (indexing a
// kernel-parameter array with a runtime index compiles to a chain of constant loads per candidate)
__device__ __forceinline__ bool tri_geo_ok(const PairDesc& pd, const float* scale_factors, const float* level_sigma2, int i, int j) {
  const asdm_kp k2 = pd.kpB[j];
  const float x2 = k2.x, y2 = k2.y;
  const int o2 = k2.octave;
  const float distex = __fsub_rn(pd.ex, x2), distey = __fsub_rn(pd.ey, y2);
  if (__fadd_rn(__fmul_rn(distex, distex), __fmul_rn(distey, distey)) < __fmul_rn(100.f, scale_factors[o2]))
    return false;
  const float x1 = pd.kpA[i].x, y1 = pd.kpA[i].y;
  const float* F = pd.F12;
  const float a = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[0]), __fmul_rn(y1, F[3])), F[6]);
  const float b = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[1]), __fmul_rn(y1, F[4])), F[7]);
  const float c = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[2]), __fmul_rn(y1, F[5])), F[8]);
  const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, x2), __fmul_rn(b, y2)), c);
  const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
  if (den == 0.f) return false;
  const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
  return (double)dsqr < __dmul_rn(3.84, (double)level_sigma2[o2]);
}

__global__ void __launch_bounds__(512) k_resolve_tri(const PairDesc* __restrict__ pairs, BowParams P, int cap,
                                                     int n_stride, int out_stride, const int32_t* __restrict__ cnt,
                                                     const Cand* __restrict__ cand, int32_t* __restrict__ acc,
                                                     int32_t* __restrict__ match, float* __restrict__ dist,
                                                     int32_t* __restrict__ nmatches, const PublishArgs pub) {
  const int pair = blockIdx.x;
  const PairDesc& pd = pairs[pair];
  const int nA = pd.nA;
  const int32_t* cntp = cnt + (size_t)pair * n_stride;
  const Cand* candp = cand + (size_t)pair * n_stride * cap;
  int32_t* accp = acc + (size_t)pair * n_stride;
  int32_t* matchp = match + (size_t)pair * out_stride;
  float* distp = dist + (size_t)pair * out_stride;
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  __shared__ int s_count;
  __shared__ float s_sf[ASDM_MAX_LEVELS], s_ls[ASDM_MAX_LEVELS];
  const int t = threadIdx.x, nt = blockDim.x;
  if (t < ASDM_HISTO_LENGTH) s_hist[t] = 0;
  if (t == 0) s_count = 0;
  if (t < ASDM_MAX_LEVELS) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int l = 0; l < ASDM_MAX_LEVELS; l++) {
      a = l == t ? P.scale_factors[l] : a;
      b = l == t ? P.level_sigma2[l] : b;
    }
    s_sf[t] = a;
    s_ls[t] = b;
  }
  __syncthreads();
  for (int i = t; i < nA; i += nt) {
    const int n = min(cntp[i], cap);
    const Cand* L = candp + (size_t)i * cap;
    float best = P.th_accept;  // :718
    int bj = -1, bpos = -1;
    for (int c = 0; c < n; c++) {
      const Cand x = L[c];
      if (x.d > P.th_accept) continue;
      const int px = pd.posB[x.j];
      // running "<=" minimum in scan order == smallest d, largest scan position among equals
      const bool better = (bj < 0) ? (x.d <= best) : (x.d < best || (x.d == best && px > bpos));
      if (!better) continue;
      if (!tri_geo_ok(pd, s_sf, s_ls, i, x.j)) continue;
      best = x.d; bj = x.j; bpos = px;
    }
    accp[i] = bj;
    matchp[i] = bj;
    distp[i] = bj >= 0 ? best : -1.f;
    if (bj >= 0) {
      atomicAdd(&s_count, 1);
      if (P.check_ori) atomicAdd(&s_hist[asdm_rot_bin(pd.kpA[i].angle, pd.kpB[bj].angle)], 1);
    }
  }
  __syncthreads();
  if (t == 0) {
    int pruned = 0;
    if (P.check_ori) {
      asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
      for (int b = 0; b < ASDM_HISTO_LENGTH; b++)
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) pruned += s_hist[b];
    }
    s_count -= pruned;
    nmatches[pair] = s_count;
  }
  __syncthreads();
  if (P.check_ori) {
    for (int i = t; i < nA; i += nt) {
      const int j = accp[i];
      if (j < 0) continue;
      const int b = asdm_rot_bin(pd.kpA[i].angle, pd.kpB[j].angle);
      if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) {
        matchp[i] = -1;
        distp[i] = -1.f;
      }
    }
  }
  if (pub.n) {
    __syncthreads();
    publish_row(pub, pair, matchp, distp, nA, s_count, pd.rowA0, pd.rowA1 >= nA ? pub.stride : pd.rowA1);  // the last block pads the row
  }
}
