// test_group.cc -- the multi-GPU group API of libasdmatch_cuda.so driven from plain C++ (no Python, no torch, no NCCL in
// the data path): the two batched workloads of BASELINE.json -- the loop-closure sweep (SearchByBoW(KF, KF) per candidate
// pair, LoopClosing.cc:291-318) and the triangulation batch of a new keyframe (SearchForTriangulation per covisible
// keyframe, LocalMapping.cc:335-370, pairs cut into row blocks) -- over a group of N ranks must equal, bit for bit, the
// same job run on one GPU through asdm_job_*.
//
//   test_group [n_ranks=2] [n_desc=1200] [n_pairs=24]
// Ranks take GPUs 0..n-1 when the box has that many, else they share the GPUs round-robin (two ranks on one GPU exercise
// everything but NVLink itself). Exit code 0 = all comparisons identical.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <random>
#include <vector>

#include "asdmatch.h"

static asdm_group* g_group = nullptr;
#define CHECK(call)                                                                          \
  do {                                                                                       \
    int rc__ = (call);                                                                       \
    if (rc__ != 0) {                                                                         \
      fprintf(stderr, "%s:%d: %s -> %d (%s)\n", __FILE__, __LINE__, #call, rc__,             \
              g_group ? asdm_group_last_error(g_group) : "");                                \
      exit(2);                                                                               \
    }                                                                                        \
  } while (0)

struct KeyFrame {
  std::vector<float> desc;
  std::vector<asdm_kp> kps;
  std::vector<uint8_t> valid, has_mp;
  std::vector<uint32_t> node_ids;
  std::vector<int32_t> off, idx;
  asdm_featvec fv;
};

static void make_keyframes(int n_kf, int n, int n_nodes, std::vector<KeyFrame>& kfs) {
  std::mt19937 rng(0xA5D00005u);
  std::normal_distribution<float> N01(0.f, 1.f);
  std::uniform_real_distribution<float> U(0.f, 1.f);
  const int pool = 3 * n;
  std::vector<float> land((size_t)pool * 128);
  std::vector<int> land_node((size_t)pool);
  for (int p = 0; p < pool; p++) {
    double s = 0;
    for (int k = 0; k < 128; k++) {
      land[(size_t)p * 128 + k] = N01(rng);
      s += (double)land[(size_t)p * 128 + k] * land[(size_t)p * 128 + k];
    }
    const float inv = (float)(1.0 / sqrt(s));
    for (int k = 0; k < 128; k++) land[(size_t)p * 128 + k] *= inv;
    land_node[(size_t)p] = (int)(rng() % (unsigned)n_nodes);
  }
  kfs.resize((size_t)n_kf);
  for (KeyFrame& kf : kfs) {
    kf.desc.resize((size_t)n * 128);
    kf.kps.resize((size_t)n);
    kf.valid.resize((size_t)n);
    kf.has_mp.resize((size_t)n);
    std::vector<std::vector<int32_t>> by_node((size_t)n_nodes);
    for (int i = 0; i < n; i++) {
      const int p = (int)(rng() % (unsigned)pool);
      double s = 0;
      float* d = &kf.desc[(size_t)i * 128];
      for (int k = 0; k < 128; k++) {
        d[k] = land[(size_t)p * 128 + k] + 0.03f * N01(rng);
        s += (double)d[k] * d[k];
      }
      const float inv = (float)(1.0 / sqrt(s));
      for (int k = 0; k < 128; k++) d[k] *= inv;
      kf.kps[(size_t)i] = asdm_kp{U(rng) * 1241.f, U(rng) * 376.f, U(rng) * 360.f, (int32_t)(rng() % 8)};
      kf.valid[(size_t)i] = U(rng) < 0.9f;
      kf.has_mp[(size_t)i] = U(rng) < 0.5f;
      by_node[(size_t)land_node[(size_t)p]].push_back(i);
    }
    kf.off.push_back(0);
    for (int k = 0; k < n_nodes; k++) {
      if (by_node[(size_t)k].empty()) continue;
      kf.node_ids.push_back((uint32_t)(10 + 3 * k));
      kf.idx.insert(kf.idx.end(), by_node[(size_t)k].begin(), by_node[(size_t)k].end());
      kf.off.push_back((int32_t)kf.idx.size());
    }
    kf.fv = asdm_featvec{(int32_t)kf.node_ids.size(), kf.node_ids.data(), kf.off.data(), kf.idx.data()};
  }
}

static int compare(const char* what, const std::vector<int32_t>& a, const std::vector<int32_t>& b, const std::vector<float>& da,
                   const std::vector<float>& db, const std::vector<int32_t>& na, const std::vector<int32_t>& nb) {
  const bool same = a == b && na == nb && da.size() == db.size() && memcmp(da.data(), db.data(), da.size() * 4) == 0;
  long matches = 0;
  for (int32_t v : na) matches += v;
  printf("%-58s %s  (%ld matches over %zu pairs)\n", what, same ? "identical" : "DIFFERENT", matches, na.size());
  return same ? 0 : 1;
}

int main(int argc, char** argv) {
  const int W = argc > 1 ? atoi(argv[1]) : 2, n = argc > 2 ? atoi(argv[2]) : 1200, n_pairs = argc > 3 ? atoi(argv[3]) : 24;
  const int n_kf = 9;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
    fprintf(stderr, "no CUDA device: the matcher has no CPU fallback\n");
    return 3;
  }
  std::vector<int> devices((size_t)W);
  for (int r = 0; r < W; r++) devices[(size_t)r] = r % ndev;
  std::vector<KeyFrame> kfs;
  make_keyframes(n_kf, n, 5, kfs);
  const asdm_grid grid{0.f, 0.f, 1241.f, 376.f, 64.f / 1241.f, 48.f / 376.f, 64, 48};
  std::vector<asdm_set_desc> descs((size_t)n_kf);
  for (int k = 0; k < n_kf; k++) descs[(size_t)k] = asdm_set_desc{kfs[(size_t)k].desc.data(), n, 0, kfs[(size_t)k].kps.data(), &grid, &kfs[(size_t)k].fv};
  std::mt19937 rng(7);
  std::vector<int32_t> ia((size_t)n_pairs), ib((size_t)n_pairs);
  for (int p = 0; p < n_pairs; p++) {
    ia[(size_t)p] = (int32_t)(rng() % n_kf);
    ib[(size_t)p] = (int32_t)((ia[(size_t)p] + 1 + rng() % (n_kf - 1)) % n_kf);
  }
  std::vector<float> F12((size_t)n_pairs * 9, 0.f), ep((size_t)n_pairs * 2, -1.0e4f);
  for (int p = 0; p < n_pairs; p++) {  // epipolar lines y2 = y1 with wide-open gates: the descriptor rule decides
    F12[(size_t)p * 9 + 5] = 1.f;
    F12[(size_t)p * 9 + 7] = -1.f;
  }
  float sf[8], ls[8];
  for (int l = 0; l < 8; l++) {
    sf[l] = powf(1.2f, (float)l);
    ls[l] = 1.0e12f;
  }
  std::vector<const uint8_t*> valid((size_t)n_kf), has_mp((size_t)n_kf);
  for (int k = 0; k < n_kf; k++) {
    valid[(size_t)k] = kfs[(size_t)k].valid.data();
    has_mp[(size_t)k] = kfs[(size_t)k].has_mp.data();
  }
  auto fill_desc = [&](asdm_job_desc& d, int kind, asdm_set* const* sets) {
    memset(&d, 0, sizeof d);
    d.kind = kind;
    d.n_kf = n_kf;
    d.sets = sets;
    d.masks = kind == ASDM_JOB_BOW_KF_KF ? valid.data() : has_mp.data();
    d.n_pairs = n_pairs;
    d.ia = ia.data();
    d.ib = ib.data();
    d.nnratio = kind == ASDM_JOB_BOW_KF_KF ? 0.85f : 0.6f;
    d.check_ori = kind == ASDM_JOB_BOW_KF_KF ? 1 : 0;
    d.F12 = F12.data();
    d.epipoles = ep.data();
    d.scale_factors = sf;
    d.level_sigma2 = ls;
    d.n_levels = 8;
  };

  // ---- one GPU, asdm_job_*: the results to reproduce
  asdm_ctx* ctx = nullptr;
  CHECK(asdm_create(0, &ctx));
  std::vector<asdm_set*> sets1((size_t)n_kf);
  CHECK(asdm_set_create_batch(ctx, descs.data(), n_kf, sets1.data()));
  std::vector<int32_t> ref_m[2], ref_n[2];
  std::vector<float> ref_d[2];
  const int kinds[2] = {ASDM_JOB_BOW_KF_KF, ASDM_JOB_TRIANGULATION};
  for (int t = 0; t < 2; t++) {
    asdm_job_desc d;
    fill_desc(d, kinds[t], sets1.data());
    asdm_job* job = nullptr;
    CHECK(asdm_job_prepare(ctx, &d, &job));
    const int S = asdm_job_out_stride(job);
    ref_m[t].resize((size_t)n_pairs * S);
    ref_d[t].resize((size_t)n_pairs * S);
    ref_n[t].resize((size_t)n_pairs);
    CHECK(asdm_job_run(job));
    CHECK(asdm_job_fetch(job, ref_m[t].data(), ref_d[t].data(), ref_n[t].data()));
    CHECK(asdm_job_destroy(job));
  }
  for (asdm_set* s : sets1) asdm_set_release(s);
  CHECK(asdm_destroy(ctx));

  // ---- the group
  asdm_group* g = nullptr;
  CHECK(asdm_group_create(devices.data(), W, (size_t)512 << 20, &g));
  g_group = g;
  int bad = 0;
  for (int round = 0; round < 2; round++) {  // twice: the second batch recycles the window range of the first
    std::vector<asdm_set*> gsets((size_t)n_kf * W);
    CHECK(asdm_group_sets_create(g, descs.data(), n_kf, gsets.data()));
    for (int t = 0; t < 2; t++) {
      for (int variant = 0; variant < 2; variant++) {
        asdm_gjob_desc gd;
        fill_desc(gd.job, kinds[t], gsets.data());
        gd.blocks_per_pair = (kinds[t] == ASDM_JOB_TRIANGULATION && variant == 1) ? 3 : 1;
        gd.gather = variant == 1 ? ASDM_GATHER_DEVICE : 0;
        asdm_gjob* job = nullptr;
        CHECK(asdm_group_job_prepare(g, &gd, &job));
        const int S = asdm_group_job_stride(job), U = asdm_group_job_units(job);
        std::vector<int32_t> m((size_t)n_pairs * S), nm((size_t)n_pairs);
        std::vector<float> dd((size_t)n_pairs * S);
        char what[128];
        if (variant == 0) {  // the convenience call: everything on the host, ordinary arrays
          CHECK(asdm_group_job_fetch(job, m.data(), dd.data(), nm.data()));
          snprintf(what, sizeof what, "round %d %s over %d ranks, fetch", round, t ? "triangulation" : "loop sweep", W);
          bad += compare(what, m, ref_m[t], dd, ref_d[t], nm, ref_n[t]);
        } else {  // rows stored by the kernels into page-locked host arrays AND into every rank's GPU
          void *hm = nullptr, *hd = nullptr, *hn = nullptr;
          CHECK(asdm_group_host_alloc(g, (size_t)n_pairs * S * 4, &hm));
          CHECK(asdm_group_host_alloc(g, (size_t)n_pairs * S * 4, &hd));
          CHECK(asdm_group_host_alloc(g, (size_t)U * 4, &hn));
          CHECK(asdm_group_job_run(job, (int32_t*)hm, (float*)hd, (int32_t*)hn));
          CHECK(asdm_group_job_wait(job));
          memcpy(m.data(), hm, m.size() * 4);
          memcpy(dd.data(), hd, dd.size() * 4);
          CHECK(asdm_group_job_counts(job, (const int32_t*)hn, nm.data()));
          snprintf(what, sizeof what, "round %d %s over %d ranks, %d units, host rows", round, t ? "triangulation" : "loop sweep", W, U);
          bad += compare(what, m, ref_m[t], dd, ref_d[t], nm, ref_n[t]);
          for (int l = 0; l < W; l++) {  // what every rank's GPU holds after the device-side barrier
            void *gm = nullptr, *gdist = nullptr, *gn = nullptr;
            CHECK(asdm_group_job_gathered(job, l, &gm, &gdist, &gn));
            std::vector<int32_t> un((size_t)U);
            cudaSetDevice(devices[(size_t)l]);
            if (cudaMemcpy(m.data(), gm, m.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess ||
                cudaMemcpy(dd.data(), gdist, dd.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess ||
                cudaMemcpy(un.data(), gn, un.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess) {
              fprintf(stderr, "cudaMemcpy of the gathered records failed\n");
              return 2;
            }
            CHECK(asdm_group_job_counts(job, un.data(), nm.data()));
            snprintf(what, sizeof what, "round %d %s, records gathered on the GPU of rank %d", round, t ? "triangulation" : "loop sweep", l);
            bad += compare(what, m, ref_m[t], dd, ref_d[t], nm, ref_n[t]);
          }
          // Pipelined runs: the gathered records rotate through three generations so that the exchange of a run overlaps
          // the next run. Learn the three generations (one run each), poison them on every rank, queue four runs back to
          // back without a host wait in between, then every generation on every rank must hold the reference again.
          std::vector<std::vector<void*>> gen_m(3), gen_d(3), gen_n(3);
          for (int k = 0; k < 3; k++) {
            CHECK(asdm_group_job_run(job, nullptr, nullptr, nullptr));
            CHECK(asdm_group_job_wait(job));
            for (int l = 0; l < W; l++) {
              void *gm = nullptr, *gdist = nullptr, *gn = nullptr;
              CHECK(asdm_group_job_gathered(job, l, &gm, &gdist, &gn));
              gen_m[(size_t)k].push_back(gm);
              gen_d[(size_t)k].push_back(gdist);
              gen_n[(size_t)k].push_back(gn);
            }
          }
          if (gen_m[0][0] == gen_m[1][0] || gen_m[1][0] == gen_m[2][0] || gen_m[0][0] == gen_m[2][0]) {
            printf("%s: the gathered records do not rotate through three generations  DIFFERENT\n", what);
            bad++;
          }
          for (int k = 0; k < 3; k++)
            for (int l = 0; l < W; l++) {
              cudaSetDevice(devices[(size_t)l]);
              cudaMemset(gen_m[(size_t)k][(size_t)l], 0x5a, m.size() * 4);
              cudaMemset(gen_d[(size_t)k][(size_t)l], 0x5a, dd.size() * 4);
              cudaMemset(gen_n[(size_t)k][(size_t)l], 0x5a, (size_t)U * 4);
            }
          for (int l = 0; l < W; l++) {
            cudaSetDevice(devices[(size_t)l]);
            cudaDeviceSynchronize();
          }
          for (int k = 0; k < 4; k++) CHECK(asdm_group_job_run(job, nullptr, nullptr, nullptr));
          CHECK(asdm_group_job_wait(job));
          for (int k = 0; k < 3; k++)
            for (int l = 0; l < W; l++) {
              std::vector<int32_t> un((size_t)U);
              cudaSetDevice(devices[(size_t)l]);
              if (cudaMemcpy(m.data(), gen_m[(size_t)k][(size_t)l], m.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess ||
                  cudaMemcpy(dd.data(), gen_d[(size_t)k][(size_t)l], dd.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess ||
                  cudaMemcpy(un.data(), gen_n[(size_t)k][(size_t)l], un.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess) {
                fprintf(stderr, "cudaMemcpy of the gathered records failed\n");
                return 2;
              }
              CHECK(asdm_group_job_counts(job, un.data(), nm.data()));
              snprintf(what, sizeof what, "round %d %s, pipelined runs, generation %d on the GPU of rank %d", round,
                       t ? "triangulation" : "loop sweep", k, l);
              bad += compare(what, m, ref_m[t], dd, ref_d[t], nm, ref_n[t]);
            }
          CHECK(asdm_group_host_free(g, hm));
          CHECK(asdm_group_host_free(g, hd));
          CHECK(asdm_group_host_free(g, hn));
        }
        CHECK(asdm_group_job_destroy(job));
      }
    }
    for (asdm_set* s : gsets) asdm_set_release(s);
  }
  CHECK(asdm_group_destroy(g));
  printf(bad ? "FAILED: %d comparisons differ\n" : "all comparisons identical\n", bad);
  return bad ? 1 : 0;
}
