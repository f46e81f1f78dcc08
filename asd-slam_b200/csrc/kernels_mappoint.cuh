// kernels_mappoint.cuh -- MapPoint::ComputeDistinctiveDescriptors (reference MapPoint.cc:271-338), batched over the
// map points of a keyframe insertion (LocalMapping.cc:245, 535, 628).
//
// Per map point with N observed descriptors: the N x N matrix of exact DescriptorDistance values (:306-318), per
// row the median vDists[0.5*(N-1)] of the sorted row INCLUDING its zero diagonal (:323-327), and the first row whose
// median is strictly below the running best (100 initially) wins (:321-334).
//
//  k_distinct_small : one warp per map point, N <= 32, distance matrix in shared memory (the common case: a map point
//                     is seen from a handful of keyframes)
//  k_distinct_large : one CTA per map point, any N, distance matrix in a global scratch slab
//
// The median needs no sort: it is the value v of the row with  #{x < v} <= m < #{x < v} + #{x == v},  m = int(0.5*(N-1)).
#pragma once
#include "internal.cuh"

struct DistinctArgs {
  const float* const* set_desc;  // [n_sets] device pointer of each keyframe's fp32 rows
  const int32_t* obs_off;        // [n_points+1]
  const int32_t* obs_set;        // [n_obs]
  const int32_t* obs_row;        // [n_obs]
  int32_t* best;                 // [n_points] position in the point's list, -1 = no observation
  float* best_median;            // [n_points]
};

// (i, j), i < j, of the p-th pair in row-major upper-triangle order
__device__ __forceinline__ void distinct_pair(int p, int N, int& i, int& j) {
  const float b = 2.f * N - 1.f;
  i = (int)((b - sqrtf(fmaxf(b * b - 8.f * p, 0.f))) * 0.5f);
  i = max(0, min(i, N - 2));
  while (i > 0 && i * (2 * N - i - 1) / 2 > p) i--;
  while ((i + 1) * (2 * N - i - 2) / 2 <= p) i++;
  j = p - i * (2 * N - i - 1) / 2 + i + 1;
}

__device__ __forceinline__ const float* distinct_row(const DistinctArgs& a, int e) {
  return a.set_desc[a.obs_set[e]] + (size_t)a.obs_row[e] * ASDM_DESC_DIM;
}

// median of row i of the N x N matrix D (row stride ld) by rank counting
__device__ __forceinline__ float distinct_median(const float* __restrict__ D, int ld, int N, int i) {
  const int m = (int)(0.5 * (double)(N - 1));  // :327
  const float* r = D + (size_t)i * ld;
  for (int j = 0; j < N; j++) {
    const float v = r[j];
    int less = 0, eq = 0;
    for (int k = 0; k < N; k++) {
      less += r[k] < v ? 1 : 0;
      eq += r[k] == v ? 1 : 0;
    }
    if (less <= m && m < less + eq) return v;
  }
  return __int_as_float(0x7f800000);  // unreachable for finite distances
}

#define DS_WARPS 8
#define DS_LD 33
__global__ void __launch_bounds__(DS_WARPS * 32) k_distinct_small(DistinctArgs a, const int32_t* __restrict__ points,
                                                                  int n_small) {
  __shared__ float sD[DS_WARPS][32 * DS_LD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = blockIdx.x * DS_WARPS + warp;
  if (w >= n_small) return;
  const int pt = points[w];
  const int e0 = a.obs_off[pt], N = a.obs_off[pt + 1] - e0;
  float* D = sD[warp];
  const unsigned full = 0xffffffffu;
  if (lane < N) D[lane * DS_LD + lane] = 0.f;
  const int npairs = N * (N - 1) / 2;
  for (int p = lane; p < npairs; p += 32) {
    int i, j;
    distinct_pair(p, N, i, j);
    const float d = asdm_exact_dist(distinct_row(a, e0 + i), distinct_row(a, e0 + j));  // :312, i < j as in the reference
    D[i * DS_LD + j] = d;
    D[j * DS_LD + i] = d;
  }
  __syncwarp();
  float med = __int_as_float(0x7f800000);
  if (lane < N) med = distinct_median(D, DS_LD, N, lane);
  // first row with the smallest median, accepted only below 100 (:321, :329-333)
  float bm = med;
  int bi = lane;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(full, bm, o);
    const int oi = __shfl_xor_sync(full, bi, o);
    if (om < bm || (om == bm && oi < bi)) {
      bm = om;
      bi = oi;
    }
  }
  if (lane == 0) {
    const bool hit = bm < 100.f;
    a.best[pt] = hit ? bi : 0;
    a.best_median[pt] = hit ? bm : 100.f;
  }
}

// One CTA per large map point; D lives in global scratch at d_off[blockIdx.x] (N*N floats).
__global__ void __launch_bounds__(256) k_distinct_large(DistinctArgs a, const int32_t* __restrict__ points,
                                                        const int64_t* __restrict__ d_off, float* __restrict__ scratch) {
  const int pt = points[blockIdx.x];
  const int e0 = a.obs_off[pt], N = a.obs_off[pt + 1] - e0;
  float* D = scratch + d_off[blockIdx.x];
  const int t = threadIdx.x, nt = blockDim.x;
  for (int i = t; i < N; i += nt) D[(size_t)i * N + i] = 0.f;
  const long npairs = (long)N * (N - 1) / 2;
  for (long p = t; p < npairs; p += nt) {
    // row-major upper triangle, decoded with 64-bit arithmetic for large N
    long i = (long)(((2.0 * N - 1.0) - sqrt(fmax((2.0 * N - 1.0) * (2.0 * N - 1.0) - 8.0 * (double)p, 0.0))) * 0.5);
    i = max(0L, min(i, (long)N - 2));
    while (i > 0 && i * (2L * N - i - 1) / 2 > p) i--;
    while ((i + 1) * (2L * N - i - 2) / 2 <= p) i++;
    const long j = p - i * (2L * N - i - 1) / 2 + i + 1;
    const float d = asdm_exact_dist(distinct_row(a, e0 + (int)i), distinct_row(a, e0 + (int)j));
    D[i * N + j] = d;
    D[j * N + i] = d;
  }
  __syncthreads();
  __shared__ float s_med[256];
  __shared__ int s_idx[256];
  float bm = __int_as_float(0x7f800000);
  int bi = 0x7fffffff;
  for (int i = t; i < N; i += nt) {  // ascending i per thread: strict '<' keeps the first
    const float med = distinct_median(D, N, N, i);
    if (med < bm) {
      bm = med;
      bi = i;
    }
  }
  s_med[t] = bm;
  s_idx[t] = bi;
  __syncthreads();
  for (int o = nt >> 1; o > 0; o >>= 1) {
    if (t < o) {
      const float om = s_med[t + o];
      const int oi = s_idx[t + o];
      if (om < s_med[t] || (om == s_med[t] && oi < s_idx[t])) {
        s_med[t] = om;
        s_idx[t] = oi;
      }
    }
    __syncthreads();
  }
  if (t == 0) {
    const bool hit = s_med[0] < 100.f;
    a.best[pt] = hit ? s_idx[0] : 0;
    a.best_median[pt] = hit ? s_med[0] : 100.f;
  }
}
