// cv_fs_shim.h -- an in-memory stand-in for cv::FileStorage / cv::FileNode, written from scratch so that the UNMODIFIED
// reference vocabulary code (/root/reference/src/dbow2/include/TemplatedVocabulary.h: save :1358-1450, load :1455-1504)
// compiles and runs here without OpenCV 3.2.  TEST INFRASTRUCTURE ONLY -- never part of the product.
//
// The document is a tree of maps / sequences / scalars kept in memory (FileStorage::doc); nothing touches the disk.
// Writing follows OpenCV's streaming protocol: inside a map, strings alternate between key and value; "{" / "{:" opens
// a map, "}" closes it, "[" / "[:" opens a sequence, "]" closes it.
#pragma once
#include <cstdlib>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace cv {

struct FsDoc {
  enum Kind { NONE, INT, REAL, STR, MAP, SEQ } kind = NONE;
  long long i = 0;
  double r = 0;
  std::string s;
  std::vector<std::pair<std::string, std::shared_ptr<FsDoc>>> fields;  // MAP (few keys: linear lookup)
  std::vector<std::shared_ptr<FsDoc>> items;                           // SEQ
  static std::shared_ptr<FsDoc> make(Kind k) {
    auto p = std::make_shared<FsDoc>();
    p->kind = k;
    return p;
  }
  static std::shared_ptr<FsDoc> of(long long v) { auto p = make(INT); p->i = v; return p; }
  static std::shared_ptr<FsDoc> of(double v) { auto p = make(REAL); p->r = v; return p; }
  static std::shared_ptr<FsDoc> of(const std::string& v) { auto p = make(STR); p->s = v; return p; }
  void put(const std::string& key, std::shared_ptr<FsDoc> v) { fields.emplace_back(key, std::move(v)); }
};

class FileNode {
 public:
  FileNode() {}
  explicit FileNode(std::shared_ptr<FsDoc> d) : d_(std::move(d)) {}
  FileNode operator[](const std::string& key) const {
    if (d_ && d_->kind == FsDoc::MAP)
      for (auto& f : d_->fields)
        if (f.first == key) return FileNode(f.second);
    return FileNode();
  }
  FileNode operator[](const char* key) const { return (*this)[std::string(key)]; }
  FileNode operator[](int i) const {
    if (d_ && d_->kind == FsDoc::SEQ && i >= 0 && (size_t)i < d_->items.size()) return FileNode(d_->items[i]);
    return FileNode();
  }
  size_t size() const {
    if (!d_) return 0;
    if (d_->kind == FsDoc::SEQ) return d_->items.size();
    if (d_->kind == FsDoc::MAP) return d_->fields.size();
    return d_->kind == FsDoc::NONE ? 0 : 1;
  }
  bool empty() const { return !d_ || d_->kind == FsDoc::NONE; }
  operator int() const { return (int)num(); }
  operator float() const { return (float)num(); }
  operator double() const { return num(); }
  operator std::string() const { return d_ && d_->kind == FsDoc::STR ? d_->s : std::string(); }

 private:
  double num() const {
    if (!d_) return 0;
    if (d_->kind == FsDoc::INT) return (double)d_->i;
    if (d_->kind == FsDoc::REAL) return d_->r;
    return 0;
  }
  std::shared_ptr<FsDoc> d_;
};

class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage() : doc(FsDoc::make(FsDoc::MAP)) { stack_.push_back(doc); }
  // "files" do not exist in this shim: a named storage opens empty (writing) or unopened (reading)
  FileStorage(const std::string&, int flags) : FileStorage() { opened_ = flags == WRITE; }
  FileStorage(const char*, int flags) : FileStorage() { opened_ = flags == WRITE; }
  explicit FileStorage(std::shared_ptr<FsDoc> d) : doc(std::move(d)) { stack_.push_back(doc); }
  bool isOpened() const { return opened_; }
  void release() {}
  FileNode operator[](const std::string& key) const { return FileNode(doc)[key]; }
  FileNode operator[](const char* key) const { return FileNode(doc)[key]; }

  // ---- streaming writer
  void write_string(const std::string& s) {
    FsDoc& top = *stack_.back();
    if (s == "}" || s == "]") {
      if (stack_.size() > 1) stack_.pop_back();
      return;
    }
    const bool open_map = s == "{" || s == "{:", open_seq = s == "[" || s == "[:";
    if (top.kind == FsDoc::MAP && !have_key_ && !open_map && !open_seq) {
      key_ = s;
      have_key_ = true;
      return;
    }
    if (open_map || open_seq) {
      auto child = FsDoc::make(open_map ? FsDoc::MAP : FsDoc::SEQ);
      attach(child);
      stack_.push_back(child);
      return;
    }
    attach(FsDoc::of(s));
  }
  void write_value(std::shared_ptr<FsDoc> v) { attach(std::move(v)); }

  std::shared_ptr<FsDoc> doc;

 private:
  void attach(std::shared_ptr<FsDoc> v) {
    FsDoc& top = *stack_.back();
    if (top.kind == FsDoc::MAP) {
      top.put(key_, std::move(v));
      have_key_ = false;
    } else {
      top.items.push_back(std::move(v));
    }
  }
  bool opened_ = true;
  bool have_key_ = false;
  std::string key_;
  std::vector<std::shared_ptr<FsDoc>> stack_;
};

inline FileStorage& operator<<(FileStorage& fs, const char* s) { fs.write_string(s); return fs; }
inline FileStorage& operator<<(FileStorage& fs, const std::string& s) { fs.write_string(s); return fs; }
inline FileStorage& operator<<(FileStorage& fs, int v) { fs.write_value(FsDoc::of((long long)v)); return fs; }
inline FileStorage& operator<<(FileStorage& fs, unsigned v) { fs.write_value(FsDoc::of((long long)v)); return fs; }
inline FileStorage& operator<<(FileStorage& fs, float v) { fs.write_value(FsDoc::of((double)v)); return fs; }
inline FileStorage& operator<<(FileStorage& fs, double v) { fs.write_value(FsDoc::of(v)); return fs; }

}  // namespace cv
