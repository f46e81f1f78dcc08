// kernels_vocab.cuh -- DBoW2 vocabulary-tree descent on the device.
//
// Reference: TemplatedVocabulary<FSift::TDescriptor, FSift>::transform(feature, word_id, weight, nid, levelsup)
// (src/dbow2/include/TemplatedVocabulary.h:1216-1257) called for every row of Frame::mDescriptors by
// transform(features, BowVector&, FeatureVector&, levelsup) (:1125-1192) from Frame::ComputeBoW (Frame.cc:289-296) and
// KeyFrame::ComputeBoW (KeyFrame.cc:203-219), with FSift::distance (src/dbow2/DBoW2/FSift.cpp:86-101): every term
// (a_i - b_i) * (a_i - b_i) is an fp32 product, the 128 terms are added to a DOUBLE accumulator in index order.
//
// Layout: one 512-byte row per child SLOT (slot = position in the CSR of children, so the k children of a node are k
// consecutive rows = one contiguous block) and one 16-byte record per slot {child node id, first slot of ITS children,
// its fan-out, weight > 0}. A level of the descent therefore needs exactly one round trip to L2: the block and its k
// records are fetched together and everything the next level needs (where the chosen child's own block lives, whether it
// is a leaf, whether the word counts) is already in shared memory when the argmin is known. G lanes carry one feature
// (G = widest fan-out of the tree, at most 32), a warp carries 32/G features. The tree for k=10, L=5 is 57 MB: it lives
// in L2 (126 MB) after the first frame, so the descent is latency- not HBM-bound.
#pragma once
#include <cuda_pipeline.h>

#include "internal.cuh"

struct VocDev {
  const float* rows;  // [n_nodes-1][128] descriptor of the child in every slot
  const int4* meta;   // [n_nodes-1] {node id, first child slot, fan-out, weight > 0} of the child in every slot
  int32_t root_k;     // fan-out of the root (its children are slots 0 .. root_k-1)
  int32_t G;          // lanes per feature
};

#define VOC_WARPS 2
#define VOC_ROW 132  // words between rows in shared memory: 16-byte aligned, 8 consecutive rows cover the 32 banks once

// shared memory of one warp: fpw feature rows, fpw*G child rows, fpw*G slot records
__host__ __device__ inline int voc_warp_words(int G) {
  const int fpw = 32 / G;
  return fpw * VOC_ROW + fpw * G * VOC_ROW + fpw * G * 4;
}

// leaf[i]: leaf node reached by row i; node_id[i]: ancestor at level nid_level (0 = root when nid_level <= 0; the leaf
// itself when the branch ends above nid_level -- the reference reads an uninitialised NodeId there, :1148);
// pos[i]: 0 when the feature enters the FeatureVector (weight > 0, :1155), else -1; key[i]: sort key of k_fv_rank.
//
// Staged kernel, for trees whose widest fan-out fits one pass (k <= 32, i.e. every practical DBoW2 tree). Per level the
// warp copies the child block of each of its features into shared memory with 16-byte cp.async (one instruction = one
// 512-byte row, conflict-free), waits once, and every lane evaluates its own child out of shared memory with LDS.128:
// a fast fp32 pass that settles the argmin whenever the margin allows, the reference's 128-step fp32-product /
// double-sum chain otherwise (exact != 0 forces the chain; exact_levels counts the decisions that took it).
__global__ void __launch_bounds__(VOC_WARPS * 32)
k_voc_descend_staged(VocDev v, const float* __restrict__ desc, int n, int nid_level, int32_t* __restrict__ leaf,
                     uint32_t* __restrict__ node_id, int32_t* __restrict__ pos, unsigned long long* __restrict__ key,
                     int exact, int32_t* __restrict__ exact_levels) {
  extern __shared__ __align__(16) float sm[];
  const int G = v.G, fpw = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane / G, c = lane - g * G;
  const bool in_group = g < fpw;
  float* wf = sm + (size_t)warp * voc_warp_words(G);
  float* wrow = wf + fpw * VOC_ROW;
  int4* wmeta = reinterpret_cast<int4*>(wrow + fpw * G * VOC_ROW);
  const int f0 = (blockIdx.x * VOC_WARPS + warp) * fpw;
  for (int s = 0; s < fpw; s++) {
    const int f = f0 + s;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (f < n) x = __ldg(reinterpret_cast<const float4*>(desc + (size_t)f * ASDM_DESC_DIM) + lane);
    reinterpret_cast<float4*>(wf + s * VOC_ROW)[lane] = x;
  }
  const int f = f0 + g;
  bool done = !(in_group && f < n);
  const float4* fb = reinterpret_cast<const float4*>(wf + (in_group ? g : 0) * VOC_ROW);
  const float4* myrow = reinterpret_cast<const float4*>(wrow + (in_group ? lane : 0) * VOC_ROW);  // row g*G + c
  const int4* mymeta = wmeta + (in_group ? g * G : 0);
  const int src0 = g * G;
  int cur = 0, level = 0, off = 0, kp = done ? 0 : v.root_k, inc = 0, n_exact = 0;
  uint32_t nid = 0;
  bool nid_set = nid_level <= 0;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  while (__any_sync(0xffffffffu, !done)) {
    __syncwarp();  // the previous level's rows and records have been read by everybody
    for (int s = 0; s < fpw; s++) {
      const int off_s = __shfl_sync(0xffffffffu, off, s * G);
      const int kp_s = __shfl_sync(0xffffffffu, done ? 0 : kp, s * G);
      const float4* src = reinterpret_cast<const float4*>(v.rows + (size_t)off_s * ASDM_DESC_DIM) + lane;
      float4* dst = reinterpret_cast<float4*>(wrow + (size_t)s * G * VOC_ROW) + lane;
      for (int j0 = 0; j0 < kp_s; j0 += 8) {  // chunks of 8 rows off one pair of base registers (immediate offsets)
#pragma unroll
        for (int u = 0; u < 8; u++)
          if (j0 + u < kp_s)
            __pipeline_memcpy_async(dst + (j0 + u) * (VOC_ROW / 4), src + (j0 + u) * (ASDM_DESC_DIM / 4), 16);
      }
      if (lane < kp_s) __pipeline_memcpy_async(wmeta + s * G + lane, v.meta + off_s + lane, 16);
    }
    __pipeline_commit();
    __pipeline_wait_prior(0);
    __syncwarp();
    // Fast pass. Only the ARGMIN leaves a level, never a distance, so the 128-step double chain (about 38 cycles per
    // dependent DADD here) is needed only when two children are too close to call. F = the same fp32 differences,
    // squared and summed in fp32 over 8 independent chains: |F - D| <= 2^-19.7 D + 2^-141 against the reference's D
    // (fp32 products, sequential double sum; all terms non-negative). When the runner-up's F exceeds the winner's by
    // more than 2^-16 (relative) + 1e-30, the winner's D is strictly the smallest, which is all the reference's
    // first-wins scan can see. Anything else -- near ties, exact ties (duplicate children), inf, NaN -- takes the
    // exact chain below.
    float facc = __int_as_float(0x7f800000);
    if (!done && c < kp) {
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f, s5 = 0.f, s6 = 0.f, s7 = 0.f;
#pragma unroll 4
      for (int i = 0; i < ASDM_DESC_DIM / 4; i += 2) {
        const float4 a0 = fb[i], b0 = myrow[i], a1 = fb[i + 1], b1 = myrow[i + 1];
        float d;
        d = __fsub_rn(a0.x, b0.x); s0 = fmaf(d, d, s0);
        d = __fsub_rn(a0.y, b0.y); s1 = fmaf(d, d, s1);
        d = __fsub_rn(a0.z, b0.z); s2 = fmaf(d, d, s2);
        d = __fsub_rn(a0.w, b0.w); s3 = fmaf(d, d, s3);
        d = __fsub_rn(a1.x, b1.x); s4 = fmaf(d, d, s4);
        d = __fsub_rn(a1.y, b1.y); s5 = fmaf(d, d, s5);
        d = __fsub_rn(a1.z, b1.z); s6 = fmaf(d, d, s6);
        d = __fsub_rn(a1.w, b1.w); s7 = fmaf(d, d, s7);
      }
      facc = __fadd_rn(__fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3)), __fadd_rn(__fadd_rn(s4, s5), __fadd_rn(s6, s7)));
    }
    int bc = 0;
    bool amb = false;
    {
      const float FINF = __int_as_float(0x7f800000);
      float fbest = FINF, fsecond = FINF;
      bool bad = false;
#pragma unroll 2
      for (int s = 0; s < G; s++) {  // branch-free: slots past the fan-out read as +inf
        const float raw = __shfl_sync(0xffffffffu, facc, src0 + s);
        const bool valid = s < kp;
        const float ds = valid ? raw : FINF;
        bad |= valid && !(raw < FINF);  // inf or NaN
        const bool lt = ds < fbest;     // strict: the first of equals stays
        fsecond = lt ? fbest : fminf(fsecond, ds);
        fbest = lt ? ds : fbest;
        bc = lt ? s : bc;
      }
      amb = !done && (exact || bad || !(fsecond > fmaf(fbest, 1.0f + 1.52587890625e-5f, 1e-30f)));
    }
    if (__any_sync(0xffffffffu, amb)) {  // warp-uniform: the reference's arithmetic to the letter
      double acc = INF;
      if (!done && c < kp) {
        acc = 0.0;
#pragma unroll 4
        for (int i = 0; i < ASDM_DESC_DIM / 4; i++) {
          const float4 a = fb[i], b = myrow[i];
          float d = __fsub_rn(a.x, b.x);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
          d = __fsub_rn(a.y, b.y);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
          d = __fsub_rn(a.z, b.z);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
          d = __fsub_rn(a.w, b.w);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
        }
      }
      double best = INF;
      for (int s = 0; s < G; s++) {  // first-wins scan in child order (strict <, :1236-1245)
        const double ds = __shfl_sync(0xffffffffu, acc, src0 + s);
        if (s < kp) {
          if (s == 0) {
            best = ds;
            bc = 0;
          } else if (ds < best) {
            best = ds;
            bc = s;
          }
        }
      }
      if (amb && c == 0) n_exact++;
    }
    if (!done) {
      const int4 m = mymeta[bc];
      level++;
      cur = m.x;
      off = m.y;
      kp = m.z;
      inc = m.w;
      if (level == nid_level) {
        nid = (uint32_t)cur;
        nid_set = true;
      }
      done = kp == 0;  // isLeaf()
    }
  }
  if (in_group && f < n && c == 0) {
    if (!nid_set) nid = (uint32_t)cur;
    leaf[f] = cur;
    node_id[f] = nid;
    pos[f] = inc ? 0 : -1;
    key[f] = inc ? (((unsigned long long)nid << 32) | (unsigned)f) : ~0ull;
    if (n_exact) atomicAdd(exact_levels, n_exact);  // (feature, level) decisions that needed the exact chain
  }
}

// General kernel: any fan-out (nodes wider than G take several passes), rows and records read straight from global
// memory. Each lane walks its own child's 512-byte row with 16-byte loads.
#define VOC_WARPS_DIRECT 4
__global__ void __launch_bounds__(VOC_WARPS_DIRECT * 32)
k_voc_descend(VocDev v, const float* __restrict__ desc, int n, int nid_level, int32_t* __restrict__ leaf,
              uint32_t* __restrict__ node_id, int32_t* __restrict__ pos, unsigned long long* __restrict__ key) {
  extern __shared__ __align__(16) float sfeat[];  // [VOC_WARPS_DIRECT * fpw][128]
  const int G = v.G, fpw = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane / G, c = lane - g * G;
  const int slot = warp * fpw + (g < fpw ? g : 0);
  const int f0 = (blockIdx.x * VOC_WARPS_DIRECT + warp) * fpw;
  float* wf = sfeat + (size_t)warp * fpw * ASDM_DESC_DIM;
  for (int s = 0; s < fpw; s++) {
    const int f = f0 + s;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (f < n) x = __ldg(reinterpret_cast<const float4*>(desc + (size_t)f * ASDM_DESC_DIM) + lane);
    reinterpret_cast<float4*>(wf + s * ASDM_DESC_DIM)[lane] = x;
  }
  __syncwarp();
  const int f = f0 + g;
  bool done = !(g < fpw && f < n);
  const float4* fb = reinterpret_cast<const float4*>(sfeat + (size_t)slot * ASDM_DESC_DIM);
  const int src0 = g * G;
  int cur = 0, level = 0, off = 0, kp = done ? 0 : v.root_k, inc = 0;
  uint32_t nid = 0;
  bool nid_set = nid_level <= 0;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  while (__any_sync(0xffffffffu, !done)) {
    // widest fan-out among the live groups decides the number of passes (warp-uniform loop bound)
    int kmax = done ? 0 : kp;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
    double best = INF;
    int bc = 0;
    for (int base = 0; base < kmax; base += G) {
      const int cc = base + c;
      double acc = INF;
      if (!done && cc < kp) {
        const float4* pb = reinterpret_cast<const float4*>(v.rows + (size_t)(off + cc) * ASDM_DESC_DIM);
        acc = 0.0;
#pragma unroll 8
        for (int i = 0; i < ASDM_DESC_DIM / 4; i++) {
          const float4 a = fb[i], b = __ldg(pb + i);
          float d = __fsub_rn(a.x, b.x);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
          d = __fsub_rn(a.y, b.y);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
          d = __fsub_rn(a.z, b.z);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
          d = __fsub_rn(a.w, b.w);
          acc = __dadd_rn(acc, (double)__fmul_rn(d, d));
        }
      }
      // sequential first-wins scan of this pass, in child order (strict <, :1236-1245)
      for (int s = 0; s < G; s++) {
        const double ds = __shfl_sync(0xffffffffu, acc, src0 + s);
        const int cs = base + s;
        if (cs < kp) {
          if (cs == 0) {
            best = ds;
            bc = 0;
          } else if (ds < best) {
            best = ds;
            bc = cs;
          }
        }
      }
    }
    if (!done) {
      const int4 m = __ldg(v.meta + off + bc);
      level++;
      cur = m.x;
      off = m.y;
      kp = m.z;
      inc = m.w;
      if (level == nid_level) {
        nid = (uint32_t)cur;
        nid_set = true;
      }
      done = kp == 0;  // isLeaf()
    }
  }
  if (g < fpw && f < n && c == 0) {
    if (!nid_set) nid = (uint32_t)cur;
    leaf[f] = cur;
    node_id[f] = nid;
    pos[f] = inc ? 0 : -1;
    key[f] = inc ? (((unsigned long long)nid << 32) | (unsigned)f) : ~0ull;
  }
}

// FeatureVector position of every feature = rank of (node, index) among the included features: the order in which
// std::map<NodeId, vector<unsigned>> lists them (FeatureVector.cpp addFeature; ascending node, insertion = index order).
// grid = (ceil(n/128), ceil(n/FV_JT)); pos was initialised to 0 / -1 by k_voc_descend.
#define FV_JT 512
__global__ void __launch_bounds__(128)
k_fv_rank(const unsigned long long* __restrict__ key, int n, int32_t* __restrict__ pos) {
  __shared__ unsigned long long sk[FV_JT];
  const int j0 = blockIdx.y * FV_JT;
  for (int t = threadIdx.x; t < FV_JT; t += 128) sk[t] = (j0 + t < n) ? key[j0 + t] : ~0ull;
  __syncthreads();
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= n) return;
  const unsigned long long mine = key[i];
  if (mine == ~0ull) return;
  int cnt = 0;
#pragma unroll 8
  for (int t = 0; t < FV_JT; t++) cnt += sk[t] < mine;
  if (cnt) atomicAdd(pos + i, cnt);
}
