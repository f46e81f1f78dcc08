// ref_voc_driver.cc -- flat C entry points around the reference's OWN vocabulary code, compiled from where it lies
// (/root/reference/src/dbow2: include/TemplatedVocabulary.h, DBoW2/FSift.cpp, BowVector.cpp, FeatureVector.cpp,
// ScoringObject.cpp) against the from-scratch OpenCV sliver in oracle/ref_shim.  TEST INFRASTRUCTURE ONLY: it pins
// oracle/asd_oracle.cc (ora_voc_transform) and generates tests/golden fixtures; output goes to oracle/_ref/ only.
//
// A vocabulary is obtained in the two ways the reference offers:
//   ref_voc_train  TemplatedVocabulary::create (:557-600, the reference's k-means++ tree, rand() seeded here) followed
//                  by save -> load through the in-memory cv::FileStorage, i.e. the state System.cc:114-121 ends up in;
//   ref_voc_load   TemplatedVocabulary::load (:1455-1504) of a document built from flat arrays (descriptors printed
//                  with 9 significant digits, which FSift::fromString reads back exactly).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "TemplatedVocabulary.h"
#include "FSift.h"

#ifdef ASD_VOC_ADAPTOR
// the same harness around the product's drop-in vocabulary (asd-slam_b200/host/asdm_vocabulary.h): transform() then
// runs on the GPU through the C ABI, everything else is still the reference's code
#include "asdm_vocabulary.h"
typedef asdm::DeviceVocabulary<DBoW2::FSift::TDescriptor, DBoW2::FSift> RefVocabulary;
#else
typedef DBoW2::TemplatedVocabulary<DBoW2::FSift::TDescriptor, DBoW2::FSift> RefVocabulary;  // ORBVocabulary.h:34
#endif

namespace {

// the tree is protected state; TemplatedVocabulary is designed to be subclassed
struct Voc : RefVocabulary {
  Voc(int k, int L, DBoW2::WeightingType w, DBoW2::ScoringType s) : RefVocabulary(k, L, w, s) {}
  int n_nodes() const { return (int)m_nodes.size(); }
  int n_children(int i) const { return (int)m_nodes[i].children.size(); }
  int child(int i, int c) const { return (int)m_nodes[i].children[c]; }
  double weight(int i) const { return m_nodes[i].weight; }
  int word(int i) const { return (int)m_nodes[i].word_id; }
  const cv::Mat& desc(int i) const { return m_nodes[i].descriptor; }
};

cv::Mat row_mat(const float* p) {
  cv::Mat m(1, 128, CV_32F);
  memcpy(m.ptr<float>(0), p, 128 * sizeof(float));
  return m;
}

}  // namespace

extern "C" {

// desc: n x 128 training descriptors split evenly over n_images "images" (the idf weights depend on the split)
void* ref_voc_train(const float* desc, int n, int n_images, int k, int L, int weighting, int scoring, unsigned seed) {
  std::vector<std::vector<cv::Mat>> training((size_t)n_images);
  for (int i = 0; i < n; i++) training[(size_t)i * n_images / n].push_back(row_mat(desc + (size_t)i * 128));
  srand(seed);
  Voc trained(k, L, (DBoW2::WeightingType)weighting, (DBoW2::ScoringType)scoring);
  trained.create(training, k, L);
  cv::FileStorage fs;
  trained.save(fs, "vocabulary");
  Voc* v = new Voc(k, L, (DBoW2::WeightingType)weighting, (DBoW2::ScoringType)scoring);
  v->load(fs, "vocabulary");
  return v;
}

// parent[i] (i >= 1) = parent of node i; nodes are handed to load() in ascending id order, so children lists come out
// in ascending id order (load pushes back in file order, :1486)
void* ref_voc_load(int k, int L, int weighting, int scoring, int n_nodes, const int32_t* parent, const float* desc,
                   const double* weight, int n_words, const int32_t* word_node) {
  auto root = cv::FsDoc::make(cv::FsDoc::MAP);
  auto voc = cv::FsDoc::make(cv::FsDoc::MAP);
  voc->put("k", cv::FsDoc::of((long long)k));
  voc->put("L", cv::FsDoc::of((long long)L));
  voc->put("scoringType", cv::FsDoc::of((long long)scoring));
  voc->put("weightingType", cv::FsDoc::of((long long)weighting));
  auto nodes = cv::FsDoc::make(cv::FsDoc::SEQ);
  for (int i = 1; i < n_nodes; i++) {
    auto nd = cv::FsDoc::make(cv::FsDoc::MAP);
    nd->put("nodeId", cv::FsDoc::of((long long)i));
    nd->put("parentId", cv::FsDoc::of((long long)parent[i]));
    nd->put("weight", cv::FsDoc::of(weight[i]));
    std::string s;
    char buf[32];
    for (int d = 0; d < 128; d++) {
      snprintf(buf, sizeof buf, "%.9g ", desc[(size_t)i * 128 + d]);
      s += buf;
    }
    nd->put("descriptor", cv::FsDoc::of(s));
    nodes->items.push_back(nd);
  }
  voc->put("nodes", nodes);
  auto words = cv::FsDoc::make(cv::FsDoc::SEQ);
  for (int w = 0; w < n_words; w++) {
    auto wd = cv::FsDoc::make(cv::FsDoc::MAP);
    wd->put("wordId", cv::FsDoc::of((long long)w));
    wd->put("nodeId", cv::FsDoc::of((long long)word_node[w]));
    words->items.push_back(wd);
  }
  voc->put("words", words);
  root->put("vocabulary", voc);
  cv::FileStorage fs(root);
  Voc* v = new Voc(k, L, (DBoW2::WeightingType)weighting, (DBoW2::ScoringType)scoring);
  v->load(fs, "vocabulary");
  return v;
}

void ref_voc_free(void* h) { delete (Voc*)h; }

// sizes[0] = n_nodes, [1] = k, [2] = L, [3] = weighting, [4] = scoring, [5] = words
void ref_voc_sizes(void* h, int32_t* sizes) {
  Voc* v = (Voc*)h;
  sizes[0] = v->n_nodes();
  sizes[1] = v->getBranchingFactor();
  sizes[2] = v->getDepthLevels();
  sizes[3] = (int)v->getWeightingType();
  sizes[4] = (int)v->getScoringType();
  sizes[5] = (int)v->size();
}

// flat export: child_off[n_nodes+1], child_idx[n_nodes-1], desc[n_nodes*128] (root row zero), weight, word_id
void ref_voc_export(void* h, int32_t* child_off, int32_t* child_idx, float* desc, double* weight, int32_t* word_id) {
  Voc* v = (Voc*)h;
  const int nn = v->n_nodes();
  int e = 0;
  for (int i = 0; i < nn; i++) {
    child_off[i] = e;
    for (int c = 0; c < v->n_children(i); c++) child_idx[e++] = v->child(i, c);
    weight[i] = v->weight(i);
    word_id[i] = v->word(i);
    if (!v->desc(i).empty()) memcpy(desc + (size_t)i * 128, v->desc(i).ptr<float>(0), 128 * sizeof(float));
    else memset(desc + (size_t)i * 128, 0, 128 * sizeof(float));
  }
  child_off[nn] = e;
}

// TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1125-1192) as Frame::ComputeBoW
// calls it (Frame.cc:289-296). Outputs: the BowVector (ascending word id), the FeatureVector flattened to per-feature
// node ids (fv_node[i] = -1 when feature i is in no node) with its iteration order (fv_order: feature indices in the
// order the map lists them), and word_id[i] from the single-feature transform (:1048-1060).
int ref_voc_transform(void* h, const float* desc, int n, int levelsup, int32_t* word_id, int32_t* fv_node,
                      int32_t* fv_order, int32_t* n_in_fv, uint32_t* bow_id, double* bow_val) {
  Voc* v = (Voc*)h;
  std::vector<cv::Mat> feats;
  feats.reserve(n);
  for (int i = 0; i < n; i++) feats.push_back(row_mat(desc + (size_t)i * 128));
  DBoW2::BowVector bow;
  DBoW2::FeatureVector fv;
  v->transform(feats, bow, fv, levelsup);
  for (int i = 0; i < n; i++) {
    fv_node[i] = -1;
    if (word_id) word_id[i] = (int32_t)v->transform(feats[i]);
  }
  int e = 0;
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it)
    for (size_t t = 0; t < it->second.size(); t++) {
      fv_node[it->second[t]] = (int32_t)it->first;
      fv_order[e++] = (int32_t)it->second[t];
    }
  *n_in_fv = e;
  int k = 0;
  for (DBoW2::BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it, ++k) {
    bow_id[k] = it->first;
    bow_val[k] = it->second;
  }
  return k;
}

// FSift::distance (FSift.cpp:86-101)
double ref_fsift_distance(const float* a, const float* b) { return DBoW2::FSift::distance(row_mat(a), row_mat(b)); }

}  // extern "C"
