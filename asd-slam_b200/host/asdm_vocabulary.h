// asdm_vocabulary.h -- drop-in for the reference's vocabulary object whose tree descent runs on the B200.
//
// The reference builds `new ORBVocabulary()` (System.cc:114; ORBVocabulary = DBoW2::TemplatedVocabulary<
// FSift::TDescriptor, FSift>, ORBVocabulary.h:34), loads it (System.cc:120) and calls
//     mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)          Frame.cc:294, KeyFrame.cc:203-219
// That transform overload is VIRTUAL in the reference (TemplatedVocabulary.h:133-135), so the integration only changes
// the `new`:      mpVocabulary = new asdm::DeviceVocabulary<DBoW2::FSift::TDescriptor, DBoW2::FSift>();
// load()/create() still run the reference's code; the first transform flattens m_nodes (protected state this subclass
// may read) and uploads it once (asdm_vocab_create). Two entry points:
//   * transform(features, bow, fv, levelsup)  -- the reference signature; the rows are uploaded to a temporary set;
//   * transform(DeviceSet&, bow, fv, levelsup) -- for Frame::ComputeBoW once the frame's descriptors are resident
//     (INTEGRATION.md): no descriptor upload, and the FeatureVector is attached to the set on the device.
// Header only; includes the reference's own TemplatedVocabulary.h (src/dbow2/include must be on the include path).
#pragma once
#include <cstring>
#include <vector>

#include "TemplatedVocabulary.h"
#include "asdm_device.h"

namespace asdm {

template <class TDescriptor, class F>
class DeviceVocabulary : public DBoW2::TemplatedVocabulary<TDescriptor, F> {
  typedef DBoW2::TemplatedVocabulary<TDescriptor, F> Base;

 public:
  using Base::Base;
  using Base::transform;  // the single-feature and BowVector-only overloads stay the reference's
  DeviceVocabulary() : Base() {}
  ~DeviceVocabulary() override {
    if (dev_) asdm_vocab_release(dev_);
  }
  DeviceVocabulary(const DeviceVocabulary&) = delete;
  DeviceVocabulary& operator=(const DeviceVocabulary&) = delete;

  using Base::load;
  void load(const cv::FileStorage& fs, const std::string& name = "vocabulary") override {  // :1455-1504
    InvalidateDevice();
    Base::load(fs, name);
  }

  // call after create() changed the tree (load() does it itself; the upload is otherwise done lazily, once)
  void InvalidateDevice() const {
    if (dev_) asdm_vocab_release(dev_);
    dev_ = nullptr;
  }

  // reference signature (TemplatedVocabulary.h:1125-1192)
  void transform(const std::vector<TDescriptor>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv,
                 int levelsup) const override {
    const int n = (int)features.size();
    std::vector<float> rows((size_t)n * ASDM_DESC_DIM);
    for (int i = 0; i < n; i++)
      std::memcpy(rows.data() + (size_t)i * ASDM_DESC_DIM, features[i].template ptr<float>(0), ASDM_DESC_DIM * sizeof(float));
    DeviceSet tmp;
    Check(asdm_set_create(Context(), rows.data(), n, 0, nullptr, nullptr, &tmp.set), "asdm_set_create");
    transform(tmp, v, fv, levelsup);
  }

  // resident descriptors: the call Frame::ComputeBoW / KeyFrame::ComputeBoW make after the integration
  void transform(DeviceSet& d, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const {
    v.clear();
    fv.clear();
    if (this->empty()) return;  // TemplatedVocabulary.h:1132
    EnsureDevice();
    const int n = asdm_set_size(d.set);
    std::vector<int32_t> word((size_t)n), node((size_t)n);
    std::vector<uint8_t> in_fv((size_t)n);
    std::vector<uint32_t> bow_id((size_t)n);
    std::vector<double> bow_val((size_t)n);
    int32_t nb = 0;
    Check(asdm_vocab_transform(Context(), dev_, d.set, levelsup, word.data(), node.data(), in_fv.data(), &nb,
                               bow_id.data(), bow_val.data()),
          "asdm_vocab_transform");
    d.has_fv = true;  // attached on the device: AttachFeatureVector has nothing left to do
    // std::map fill in ascending key order with an end() hint: O(1) per insertion
    for (int k = 0; k < nb; k++) v.insert(v.end(), DBoW2::BowVector::value_type(bow_id[k], bow_val[k]));
    for (int i = 0; i < n; i++)
      if (in_fv[i]) fv.addFeature((unsigned)node[i], (unsigned)i);  // FeatureVector.cpp addFeature: ascending i per node
  }

 private:
  void EnsureDevice() const {
    if (dev_) return;
    const int nn = (int)this->m_nodes.size();
    std::vector<int32_t> off((size_t)nn + 1, 0), idx, word((size_t)nn);
    std::vector<float> desc((size_t)nn * ASDM_DESC_DIM, 0.f);
    std::vector<double> weight((size_t)nn);
    idx.reserve(nn);
    for (int i = 0; i < nn; i++) {
      const auto& nd = this->m_nodes[i];
      for (size_t c = 0; c < nd.children.size(); c++) idx.push_back((int32_t)nd.children[c]);
      off[i + 1] = (int32_t)idx.size();
      weight[i] = nd.weight;
      word[i] = (int32_t)nd.word_id;
      if (i > 0 && !nd.descriptor.empty())
        std::memcpy(desc.data() + (size_t)i * ASDM_DESC_DIM, nd.descriptor.template ptr<float>(0), ASDM_DESC_DIM * sizeof(float));
    }
    asdm_vocab_desc d;
    d.k = this->m_k;
    d.L = this->m_L;
    d.weighting = (int32_t)this->m_weighting;
    d.scoring = (int32_t)this->m_scoring;
    d.n_nodes = nn;
    d.child_off = off.data();
    d.child_idx = idx.data();
    d.desc = desc.data();
    d.weight = weight.data();
    d.word_id = word.data();
    Check(asdm_vocab_create(Context(), &d, &dev_), "asdm_vocab_create");
  }
  mutable asdm_vocab* dev_ = nullptr;
};

}  // namespace asdm
