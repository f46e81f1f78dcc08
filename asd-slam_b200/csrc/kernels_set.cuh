// kernels_set.cuh -- one-time preparation of a device-resident descriptor set (Frame.h:147, KeyFrame.h:199):
//  k_pack_desc  : fp32 rows -> bf16 tensor-core copy + squared row norms + range flag
//  k_build_grid : Frame::AssignFeaturesToGrid / PosInGrid (Frame.cc:123-138, 276-286; KeyFrame.cc:116-143) as a CSR
//                 bucket whose cells list keypoints in insertion (= index) order.
#pragma once
#include "internal.cuh"

// One warp per row; lane l owns floats 4l..4l+3 (one LDG.128, one STG.64). Algorithmic bytes per row:
// 512 read + 256 + 4 written.
__global__ void __launch_bounds__(256) k_pack_desc(const float* __restrict__ desc, int n, int n_pad,
                                                   __nv_bfloat16* __restrict__ bf, float* __restrict__ norm2,
                                                   int32_t* __restrict__ flags) {
  const int lane = threadIdx.x & 31;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n_pad) return;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row < n) v = __ldg(reinterpret_cast<const float4*>(desc + (size_t)row * ASDM_DESC_DIM) + lane);
  float s = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 packed;
  packed.x = *reinterpret_cast<uint32_t*>(&lo);
  packed.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(bf + (size_t)row * ASDM_DESC_DIM + lane * 4) = packed;
  if (lane == 0) {
    norm2[row] = s;
    // The tensor-core shortlist bound (DESIGN.md "error bound") needs finite rows with a sane norm.
    const bool ok = (s == 0.f) || (s >= 9.5367431640625e-07f && s <= 1048576.f);  // [2^-20, 2^20]
    if (!ok) flags[0] = 1;
  }
}

// Single CTA. smem: cols*rows+1 counters.
__global__ void __launch_bounds__(1024) k_build_grid(const asdm_kp* __restrict__ kps, int n, asdm_grid g,
                                                     int32_t* __restrict__ cell_of_tmp,
                                                     int32_t* __restrict__ cell_start,
                                                     int32_t* __restrict__ cell_items, int4* __restrict__ cell_kp) {
  extern __shared__ int32_t s_cnt[];  // [nc + 1]
  __shared__ int32_t s_warp[32];
  const int nc = g.cols * g.rows;
  const int t = threadIdx.x, nt = blockDim.x;
  for (int c = t; c <= nc; c += nt) s_cnt[c] = 0;
  __syncthreads();
  for (int i = t; i < n; i += nt) {
    const asdm_kp kp = kps[i];
    // Frame.cc:278-279
    const float fx = roundf(__fmul_rn(__fsub_rn(kp.x, g.min_x), g.inv_w));
    const float fy = roundf(__fmul_rn(__fsub_rn(kp.y, g.min_y), g.inv_h));
    int cell = -1;
    if (fx >= 0.f && fx < (float)g.cols && fy >= 0.f && fy < (float)g.rows) {
      cell = (int)fx * g.rows + (int)fy;
      atomicAdd(&s_cnt[cell], 1);
    }
    cell_of_tmp[i] = cell;
  }
  __syncthreads();
  // exclusive scan of s_cnt[0..nc) in chunks of blockDim
  int carry = 0;
  for (int base = 0; base < nc; base += nt) {
    const int c = base + t;
    const int v = c < nc ? s_cnt[c] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if ((t & 31) >= o) x += y;
    }
    if ((t & 31) == 31) s_warp[t >> 5] = x;
    __syncthreads();
    if (t < 32) {
      int w = t < (nt >> 5) ? s_warp[t] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (t >= o) w += y;
      }
      s_warp[t] = w;
    }
    __syncthreads();
    const int warp_off = (t >> 5) ? s_warp[(t >> 5) - 1] : 0;
    const int total = s_warp[(nt >> 5) - 1];
    const int excl = carry + warp_off + x - v;
    __syncthreads();
    if (c < nc) {
      cell_start[c] = excl;
      s_cnt[c] = excl;  // becomes the fill cursor
    }
    carry += total;
  }
  if (t == 0) cell_start[nc] = carry;
  __syncthreads();
  for (int i = t; i < n; i += nt) {
    const int cell = cell_of_tmp[i];
    if (cell >= 0) cell_items[atomicAdd(&s_cnt[cell], 1)] = i;
  }
  __syncthreads();
  // mGrid[ix][iy] lists keypoints in push_back order = ascending index (Frame.cc:128-137)
  for (int c = t; c < nc; c += nt) {
    const int b = cell_start[c], e = s_cnt[c];
    for (int a = b + 1; a < e; a++) {
      const int x = cell_items[a];
      int k = a - 1;
      while (k >= b && cell_items[k] > x) {
        cell_items[k + 1] = cell_items[k];
        k--;
      }
      cell_items[k + 1] = x;
    }
  }
  __syncthreads();
  // the same bucket with the keypoint fields inline (k_proj_fused reads one 16-byte record per candidate)
  for (int a = t; a < carry; a += nt) {
    const int i = cell_items[a];
    const asdm_kp kp = kps[i];
    cell_kp[a] = make_int4(__float_as_int(kp.x), __float_as_int(kp.y), kp.octave, i);
  }
}

// Batched k_pack_desc for asdm_set_create_batch: one launch packs every set of a batch (a wave of keyframes that
// arrived together). blockIdx.y = set; rows beyond n are zeroed here (desc, bf16 copy, norm), so the batch needs no
// memset per set.
struct PackItem {
  float* desc;
  __nv_bfloat16* bf;
  float* norm2;
  int32_t* flags;
  int32_t n, n_pad;
  int32_t* h_flag;  // page-locked, device-visible copy of flags[0] (written straight over PCIe: no copy-back per set)
};

__global__ void __launch_bounds__(256) k_pack_batch(const PackItem* __restrict__ items) {
  const PackItem it = items[blockIdx.y];
  const int lane = threadIdx.x & 31;
  for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < it.n_pad; row += (gridDim.x * blockDim.x) >> 5) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    float4* src = reinterpret_cast<float4*>(it.desc + (size_t)row * ASDM_DESC_DIM) + lane;
    if (row < it.n) v = *src;
    else *src = v;
    float s = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 packed;
    packed.x = *reinterpret_cast<uint32_t*>(&lo);
    packed.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(it.bf + (size_t)row * ASDM_DESC_DIM + lane * 4) = packed;
    if (lane == 0) {
      it.norm2[row] = s;
      const bool ok = (s == 0.f) || (s >= 9.5367431640625e-07f && s <= 1048576.f);  // [2^-20, 2^20], as k_pack_desc
      if (!ok) {
        it.flags[0] = 1;
        if (it.h_flag) *it.h_flag = 1;
      }
    }
  }
}
