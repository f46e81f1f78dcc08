// kernels_vocab.cuh -- DBoW2 vocabulary-tree descent on the device.
//
// Reference: TemplatedVocabulary<FSift::TDescriptor, FSift>::transform(feature, word_id, weight, nid, levelsup)
// (src/dbow2/include/TemplatedVocabulary.h:1216-1257) called for every row of Frame::mDescriptors by
// transform(features, BowVector&, FeatureVector&, levelsup) (:1125-1192) from Frame::ComputeBoW (Frame.cc:289-296) and
// KeyFrame::ComputeBoW (KeyFrame.cc:203-219), with FSift::distance (src/dbow2/DBoW2/FSift.cpp:86-101): every term
// (a_i - b_i) * (a_i - b_i) is an fp32 product, the 128 terms are added to a DOUBLE accumulator in index order.
//
// Layout: the k children of an inner node sit in one block, dimension-major ([128][k] floats at child_off[p]*128), so
// the k lanes that evaluate the k children of one feature read consecutive words at every dimension. A warp carries
// 32/G features (G = widest fan-out of the tree, at most 32 lanes; wider nodes take several passes). The tree for
// k=10, L=5 is 57 MB: it lives in L2 (126 MB) after the first frame, so the descent is latency- not HBM-bound.
#pragma once
#include "internal.cuh"

struct VocDev {
  const int32_t* child_off;  // [n_nodes+1] CSR of children (push_back order of TemplatedVocabulary::load, :1486)
  const int32_t* child_idx;  // node id of every child slot
  const float* blk;          // child descriptors, one [128][kp] block per inner node
  const double* weight;      // [n_nodes] Node::weight
  int32_t n_nodes;
  int32_t G;  // lanes per feature
};

#define VOC_WARPS 4

// leaf[i]: leaf node reached by row i; node_id[i]: ancestor at level nid_level (0 = root when nid_level <= 0; the leaf
// itself when the branch ends above nid_level -- the reference reads an uninitialised NodeId there, :1148);
// pos[i]: 0 when the feature enters the FeatureVector (weight > 0, :1155), else -1; key[i]: sort key of k_fv_rank.
__global__ void __launch_bounds__(VOC_WARPS * 32)
k_voc_descend(VocDev v, const float* __restrict__ desc, int n, int nid_level, int32_t* __restrict__ leaf,
              uint32_t* __restrict__ node_id, int32_t* __restrict__ pos, unsigned long long* __restrict__ key) {
  extern __shared__ float sfeat[];  // [VOC_WARPS * fpw][128]
  const int G = v.G, fpw = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane / G, c = lane - g * G;
  const int slot = warp * fpw + (g < fpw ? g : 0);
  const int f0 = (blockIdx.x * VOC_WARPS + warp) * fpw;
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
  const float* fb = sfeat + (size_t)slot * ASDM_DESC_DIM;
  const int src0 = g * G;
  int cur = 0, level = 0;
  uint32_t nid = 0;
  bool nid_set = nid_level <= 0;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  while (__any_sync(0xffffffffu, !done)) {
    int off = 0, kp = 0;
    if (!done) {
      off = __ldg(v.child_off + cur);
      kp = __ldg(v.child_off + cur + 1) - off;
    }
    // widest fan-out among the live groups decides the number of passes (warp-uniform loop bound)
    int kmax = kp;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
    double best = INF;
    int bc = 0;
    for (int base = 0; base < kmax; base += G) {
      const int cc = base + c;
      double acc = INF;
      if (!done && cc < kp) {
        const float* pb = v.blk + (size_t)off * ASDM_DESC_DIM + cc;
        acc = 0.0;
#pragma unroll 16
        for (int i = 0; i < ASDM_DESC_DIM; i++) {
          const float d = __fsub_rn(fb[i], __ldg(pb + (size_t)i * kp));
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
      level++;
      cur = __ldg(v.child_idx + off + bc);
      if (level == nid_level) {
        nid = (uint32_t)cur;
        nid_set = true;
      }
      done = __ldg(v.child_off + cur + 1) == __ldg(v.child_off + cur);  // isLeaf()
    }
  }
  if (g < fpw && f < n && c == 0) {
    if (!nid_set) nid = (uint32_t)cur;
    const bool inc = __ldg(v.weight + cur) > 0.0;
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
