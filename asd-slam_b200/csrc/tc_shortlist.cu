// tc_shortlist.cu -- K1: the N x M distance matrix of a keyframe pair as a dense contraction on the 5th-gen tensor
// cores (tcgen05.mma, bf16 operands staged by TMA, fp32 accumulators in TMEM), reduced on the fly to the short list
// of (i, j) whose squared distance can be <= t_keep.  k_rerank then recomputes those few pairs in the reference's
// exact fp32 arithmetic, so the tensor-core result is only ever a filter (DESIGN.md "error bound").
//
//   ||a-b||^2 = |a|^2 + |b|^2 - 2 a.b ;  acc = bf16(a).bf16(b) in fp32 ;  |acc - a.b| <= 1.25 * 2^-8 * (|a|^2+|b|^2)
//   d <= t_keep  =>  acc >= KAPPA*(|a|^2 + |b|^2) - t_keep/2 * (1+1e-4),   KAPPA = 0.495
//
// Persistent kernel, one CTA per SM, 10 warps: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane),
// warps 2..9 = epilogue (two per TMEM lane quarter, 128 columns each). A work unit keeps one 256-row tile of set B resident in shared
// memory (double buffered across units) and streams 128-row tiles of set A through a 3-stage ring; accumulators are
// double buffered in TMEM (2 x 256 columns), so the epilogue of tile t overlaps the MMAs of tile t+1. Each epilogue
// warp pulls its 32 lanes x 128 columns into registers with four back-to-back tcgen05.ld, waits once, and hands the
// accumulator back to the MMA warp BEFORE it looks at the values, so TMEM is held for one load latency per tile
// (1437 -> 1632 TFLOP/s on C5; without that the MMA warp stalled on t_empty: profiles/r01e vs r01f).
// Per tile: 8 x tcgen05.mma 128x256x16 = 8.39 MFLOP; epilogue = 32768 fp32 reads from TMEM, one max per value.
#include <cuda.h>

#include "internal.cuh"

namespace {

constexpr int TM = 128;        // rows of A per tile (UMMA M)
constexpr int TN = 256;        // rows of B per tile (UMMA N)
constexpr int KD = ASDM_DESC_DIM;
constexpr int MAX_A_STAGES = 6;
constexpr int A_BYTES = TM * KD * 2;   // 32 KB
constexpr int B_BYTES = TN * KD * 2;   // 64 KB
constexpr int KBLK_A = TM * 128;       // bytes of one 64-element K block of an A tile (128 B per row)
constexpr int KBLK_B = TN * 128;
constexpr int tc_smem(int a_stages, int b_bufs) { return b_bufs * B_BYTES + a_stages * A_BYTES + 256; }  // tiles + barriers
constexpr float KAPPA = 0.495f;

// instruction descriptor (cute::UMMA::InstrDescriptor): D=f32, A=B=bf16, both K-major, N=256, M=128
constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const void* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// K-major, 128-byte swizzle: 8-row groups are 1024 B apart (SBO), LBO unused, descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

__device__ __forceinline__ float max_of_32(const uint32_t (&v)[32]) {
  float m[11];
#pragma unroll
  for (int k = 0; k < 10; k++)
    m[k] = max3(__uint_as_float(v[3 * k]), __uint_as_float(v[3 * k + 1]), __uint_as_float(v[3 * k + 2]));
  m[10] = fmaxf(__uint_as_float(v[30]), __uint_as_float(v[31]));
  const float a = max3(m[0], m[1], m[2]), b = max3(m[3], m[4], m[5]), c = max3(m[6], m[7], m[8]);
  return max3(max3(a, b, c), m[9], m[10]);
}

struct __align__(8) TcBarriers {
  uint64_t a_full[MAX_A_STAGES], a_empty[MAX_A_STAGES];
  uint64_t b_full[2], b_empty[2];
  uint64_t t_full[2], t_empty[2];
  uint32_t tmem_base;
};

// The slow path of the epilogue, entered by a whole warp when some lane of it saw max(32 columns) >= its row bound.
// Emits (pair, i, j) for every column at or above the row bound; the query-side gates are already folded into the
// bound (+inf for unusable rows), the target-side gates (mask, node, j < nB) are re-checked by k_rerank.
__device__ __noinline__ int tc_emit_hits(uint32_t mask, int pair, int i, int jbase, TcHit* __restrict__ seg, int seg_cnt,
                                         int seg_cap) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  while (__any_sync(full, mask != 0u)) {
    const bool has = mask != 0u;
    const int c = __ffs((int)mask) - 1;
    mask &= mask - 1u;
    const unsigned b = __ballot_sync(full, has);
    if (has) {
      const int pos = seg_cnt + __popc(b & lt);
      if (pos < seg_cap) seg[pos] = TcHit{pair, i, jbase + c};
    }
    seg_cnt += __popc(b);
  }
  return seg_cnt;
}

// Row / column bounds of the filter, once per keyframe of a job: rb[i] = KAPPA*|x_i|^2 for a usable feature (mask ==
// mask_sense and present in the FeatureVector), +inf otherwise; cmin[c] = min of rb over the 128-row chunk c. The
// epilogue then needs ONE independent load per tile (a dependent mask -> pos -> norm chain cost three L2 round
// trips per tile and kept the tensor pipe at 50 %: profiles/r01d).
__global__ void __launch_bounds__(128) k_tc_row_bounds(const TcSetDesc* __restrict__ sets,
                                                       const int32_t* __restrict__ chunk_set, int mask_sense,
                                                       float* __restrict__ rb, float* __restrict__ cmin) {
  const TcSetDesc sd = sets[chunk_set[blockIdx.x]];
  const int i = (blockIdx.x - sd.chunk0) * 128 + threadIdx.x;
  float v = __int_as_float(0x7f800000);
  if (i < sd.n && sd.mask[i] == mask_sense && sd.pos[i] >= 0) v = KAPPA * sd.norm2[i];
  rb[(size_t)blockIdx.x * 128 + threadIdx.x] = v;
  __shared__ float s_min[4];
  float m = v;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) cmin[blockIdx.x] = fminf(fminf(s_min[0], s_min[1]), fminf(s_min[2], s_min[3]));
}

// EPI_WARPS = 4 or 8 (with 8, the two warps of a TMEM lane quarter split the 256 columns); A_STAGES x 32 KB ring of
// streamed A tiles, B_BUFS x 64 KB resident B tiles; NO_EPI (timing experiments only) skips the TMEM reads.
template <int EPI_WARPS, int A_STAGES, int B_BUFS, bool NO_EPI, bool EARLY>
__global__ void __launch_bounds__(64 + 32 * EPI_WARPS, 1)
k_tc_shortlist(const PairDesc* __restrict__ pairs, const uint8_t* __restrict__ tmaps, const TcWork* __restrict__ work,
               int n_work, float t_half, TcHit* __restrict__ list, int seg_cap, int32_t* __restrict__ seg_count,
               int32_t* __restrict__ list_total) {
  constexpr int EPI_COLS = TN / (EPI_WARPS / 4);
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* smB = smem;                      // B_BUFS x 64 KB
  uint8_t* smA = smem + B_BUFS * B_BYTES;   // A_STAGES x 32 KB
  TcBarriers& bars = *reinterpret_cast<TcBarriers*>(smem + B_BUFS * B_BYTES + A_STAGES * A_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) __trap();  // the 128-byte swizzle of the tiles assumes 1024 B alignment
    for (int s = 0; s < A_STAGES; s++) {
      mbar_init(&bars.a_full[s], 1);
      mbar_init(&bars.a_empty[s], 1);
    }
    for (int s = 0; s < 2; s++) {
      mbar_init(&bars.b_full[s], 1);
      mbar_init(&bars.b_empty[s], 1);
      mbar_init(&bars.t_full[s], 1);
      mbar_init(&bars.t_empty[s], EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars.tmem_base)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars.tmem_base;

  if (warp == 0) {
    // ================================================================ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t aphase = 0;
      int ul = 0;
      for (int u = blockIdx.x; u < n_work; u += gridDim.x, ul++) {
        const TcWork w = work[u];
        const PairDesc& pd = pairs[w.pair];
        const void* tmA = tmaps + (size_t)pd.setA * 128;
        const void* tmB = tmaps + (size_t)pd.setB * 128;
        const int bb = ul % B_BUFS;
        mbar_wait(&bars.b_empty[bb], ((ul / B_BUFS) & 1) ^ 1);
        mbar_expect_tx(&bars.b_full[bb], B_BYTES);
#pragma unroll
        for (int kb = 0; kb < 2; kb++)
#pragma unroll
          for (int h = 0; h < 2; h++)
            tma_load_2d(smB + bb * B_BYTES + kb * KBLK_B + h * (128 * 128), tmB, kb * 64, w.n_tile * TN + h * 128,
                        &bars.b_full[bb]);
        for (int t = 0; t < w.m_tiles; t++) {
          mbar_wait(&bars.a_empty[stage], aphase ^ 1);
          mbar_expect_tx(&bars.a_full[stage], A_BYTES);
#pragma unroll
          for (int kb = 0; kb < 2; kb++)
            tma_load_2d(smA + stage * A_BYTES + kb * KBLK_A, tmA, kb * 64, (w.m_tile0 + t) * TM, &bars.a_full[stage]);
          if (++stage == A_STAGES) {
            stage = 0;
            aphase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer
    if (lane == 0) {
      int stage = 0, acc = 0;
      uint32_t aphase = 0, accphase = 0;
      int ul = 0;
      for (int u = blockIdx.x; u < n_work; u += gridDim.x, ul++) {
        const int m_tiles = work[u].m_tiles;
        const int bb = ul % B_BUFS;
        mbar_wait(&bars.b_full[bb], (ul / B_BUFS) & 1);
        const uint32_t b_addr = smem_u32(smB + bb * B_BYTES);
        for (int t = 0; t < m_tiles; t++) {
          mbar_wait(&bars.t_empty[acc], accphase ^ 1);
          mbar_wait(&bars.a_full[stage], aphase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smA + stage * A_BYTES);
          const uint32_t d_tmem = tmem_base + (uint32_t)acc * TN;
#pragma unroll
          for (int kb = 0; kb < 2; kb++)
#pragma unroll
            for (int kk = 0; kk < 4; kk++)
              umma_bf16(d_tmem, umma_desc(a_addr + kb * KBLK_A + kk * 32), umma_desc(b_addr + kb * KBLK_B + kk * 32), IDESC,
                        (kb | kk) ? 1u : 0u);
          umma_commit(&bars.a_empty[stage]);  // smem slot reusable once these MMAs have read it
          umma_commit(&bars.t_full[acc]);     // accumulator ready for the epilogue
          if (++stage == A_STAGES) {
            stage = 0;
            aphase ^= 1;
          }
          acc ^= 1;
          if (acc == 0) accphase ^= 1;
        }
        umma_commit(&bars.b_empty[bb]);
      }
    }
  } else {
    // ================================================================ epilogue: EPI_WARPS warps; warp e reads TMEM
    // lanes 32*(warp%4).. (the hardware's lane-quarter rule) and columns [chalf*EPI_COLS, +EPI_COLS)
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int chalf = ew >> 2;
    const int row = quarter * 32 + lane;
    const int seg_id = blockIdx.x * EPI_WARPS + ew;
    TcHit* __restrict__ seg = list + (size_t)seg_id * seg_cap;
    int seg_cnt = 0;
    int acc = 0;
    uint32_t accphase = 0;
    const float INF = __int_as_float(0x7f800000);
    for (int u = blockIdx.x; u < n_work; u += gridDim.x) {
      const TcWork w = work[u];
      const PairDesc& pd = pairs[w.pair];
      const float* __restrict__ rbA = pd.rbA + (size_t)w.m_tile0 * TM + row;
      const int j0 = w.n_tile * TN + chalf * EPI_COLS;
      const int c0 = j0 >> 7;
      float minb = __ldg(pd.cminB + c0);
      if (EPI_COLS == 256) minb = fminf(minb, __ldg(pd.cminB + c0 + 1));
      const float colb = minb - t_half;  // +inf when no column of the range is usable
      // the row bound of the next tile is fetched one tile ahead: a single independent load whose latency hides
      // behind this tile's TMEM reads
      float ai_next = __ldg(rbA);
      for (int t = 0; t < w.m_tiles; t++) {
        const int i = (w.m_tile0 + t) * TM + row;
        const float thr_row = ai_next + colb;  // +inf for padding / unusable rows
        ai_next = (t + 1 < w.m_tiles) ? __ldg(rbA + (size_t)(t + 1) * TM) : INF;
        mbar_wait(&bars.t_full[acc], accphase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * TN + chalf * EPI_COLS);
        if (EARLY) {
          // all 128 columns of this warp into registers at once, then hand the accumulator back to the MMA warp
          // BEFORE looking at the values: TMEM is held for one load latency per tile instead of the whole epilogue
          static_assert(!EARLY || EPI_COLS == 128, "early release keeps 128 columns in registers");
          uint32_t v[4][32];
#pragma unroll
          for (int ch = 0; ch < 4; ch++) tmem_ld32(taddr + ch * 32, v[ch]);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars.t_empty[acc]);
#pragma unroll
          for (int ch = 0; ch < 4; ch++) {
            if (__any_sync(0xffffffffu, max_of_32(v[ch]) >= thr_row)) {
              uint32_t mask = 0;
#pragma unroll
              for (int c = 0; c < 32; c++) mask |= (__uint_as_float(v[ch][c]) >= thr_row) ? (1u << c) : 0u;
              seg_cnt = tc_emit_hits(mask, w.pair, i, j0 + ch * 32, seg, seg_cnt, seg_cap);
            }
          }
        } else {
          uint32_t v[2][32];
          if (!NO_EPI) tmem_ld32(taddr, v[0]);
#pragma unroll 2
          for (int ch = 0; ch < (NO_EPI ? 0 : EPI_COLS / 32); ch++) {
            tmem_ld_wait();
            if (ch + 1 < EPI_COLS / 32) tmem_ld32(taddr + (ch + 1) * 32, v[(ch + 1) & 1]);
            const uint32_t(&cur)[32] = v[ch & 1];
            if (__any_sync(0xffffffffu, max_of_32(cur) >= thr_row)) {
              uint32_t mask = 0;
#pragma unroll
              for (int c = 0; c < 32; c++) mask |= (__uint_as_float(cur[c]) >= thr_row) ? (1u << c) : 0u;
              seg_cnt = tc_emit_hits(mask, w.pair, i, j0 + ch * 32, seg, seg_cnt, seg_cap);
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars.t_empty[acc]);
        }
        acc ^= 1;
        if (acc == 0) accphase ^= 1;
      }
    }
    if (lane == 0) {
      seg_count[seg_id] = seg_cnt;  // > seg_cap means this segment overflowed: the host grows the list and reruns
      if (seg_cnt) atomicAdd(list_total, seg_cnt);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

}  // namespace

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int asdm_tc_build_tmap(asdm_ctx* ctx, asdm_set* set) {
  std::lock_guard<std::mutex> lks(set->mu);  // a set may be prepared for jobs of several contexts at once
  if (!set->tmap_bf.empty()) return ASDM_OK;
  if (!ctx->tmap_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) {
      cudaGetLastError();
      return ASDM_ERR_CUDA;
    }
    ctx->tmap_encode = fn;
  }
  CUtensorMap tm;
  const cuuint64_t dims[2] = {(cuuint64_t)KD, (cuuint64_t)set->n_pad};
  const cuuint64_t strides[1] = {(cuuint64_t)KD * 2};
  const cuuint32_t box[2] = {64, 128};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = ((PFN_encodeTiled)ctx->tmap_encode)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, set->desc_bf, dims, strides,
                                                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return ASDM_ERR_CUDA;
  set->tmap_bf.resize(128);
  memcpy(set->tmap_bf.data(), &tm, 128);
  return ASDM_OK;
}

cudaError_t asdm_tc_row_bounds_launch(asdm_ctx* ctx, const TcSetDesc* d_sets, const int32_t* d_chunk_set, int n_chunks,
                                      int mask_sense, float* d_rb, float* d_cmin) {
  if (n_chunks <= 0) return cudaSuccess;
  k_tc_row_bounds<<<n_chunks, 128, 0, ctx->stream>>>(d_sets, d_chunk_set, mask_sense, d_rb, d_cmin);
  return cudaGetLastError();
}

template <int EPI_WARPS, int A_STAGES, int B_BUFS, bool NO_EPI, bool EARLY>
static cudaError_t tc_launch(asdm_ctx* ctx, const PairDesc* d_pairs, const void* d_tmaps, const TcWork* d_work, int n_work,
                             float t_half, TcHit* d_list, int seg_cap, int32_t* d_seg_count, int32_t* d_list_total) {
  // the attribute belongs to the (function, device) pair: a process may drive several GPUs (asdm_group_create)
  static unsigned long long attr_set = 0;  // bit d: done on device d
  if (ctx->device >= 64 || !((attr_set >> ctx->device) & 1ull)) {
    cudaError_t e = cudaFuncSetAttribute(k_tc_shortlist<EPI_WARPS, A_STAGES, B_BUFS, NO_EPI, EARLY>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem(A_STAGES, B_BUFS));
    if (e != cudaSuccess) return e;
    if (ctx->device < 64) attr_set |= 1ull << ctx->device;
  }
  const int grid = n_work < ctx->sm_count ? n_work : ctx->sm_count;
  k_tc_shortlist<EPI_WARPS, A_STAGES, B_BUFS, NO_EPI, EARLY><<<grid, 64 + 32 * EPI_WARPS, tc_smem(A_STAGES, B_BUFS), ctx->stream>>>(
      d_pairs, (const uint8_t*)d_tmaps, d_work, n_work, t_half, d_list, seg_cap, d_seg_count, d_list_total);
  return cudaGetLastError();
}

cudaError_t asdm_tc_shortlist_launch(asdm_ctx* ctx, const PairDesc* d_pairs, const void* d_tmaps, const TcWork* d_work,
                                     int n_work, float t_keep, TcHit* d_list, int seg_cap, int32_t* d_seg_count,
                                     int32_t* d_list_total) {
  const float t_half = 0.5f * t_keep * 1.0001f + 1e-6f;
#define TC_ARGS ctx, d_pairs, d_tmaps, d_work, n_work, t_half, d_list, seg_cap, d_seg_count, d_list_total
  switch (ctx->opt_tc_variant) {  // tuning experiments (asdm_set_option "tc_variant"); 0 is the shipped configuration
    case 1: return tc_launch<8, 5, 1, false, false>(TC_ARGS);
    case 2: return tc_launch<8, 2, 2, false, false>(TC_ARGS);
    case 3: return tc_launch<8, 3, 2, true, false>(TC_ARGS);  // no epilogue: timing of the TMA + MMA pipeline only
    case 5: return tc_launch<8, 3, 2, false, false>(TC_ARGS);  // accumulator released after the whole epilogue
    default: break;
  }
  if (ctx->opt_tc_epi_warps == 8) return tc_launch<8, 3, 2, false, true>(TC_ARGS);
  return tc_launch<4, 3, 2, false, false>(TC_ARGS);
#undef TC_ARGS
}
