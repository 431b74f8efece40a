// cow_vec.hpp -- copy-on-write vector shared by the problem object (lp_model.hpp) and the engine's
// host mirror (engine_cuda.cu): every open B&B node is a copy of its parent (util.cpp:25-42), and
// a copy costs a reference count until one side writes.
#pragma once

#include <algorithm>
#include <memory>
#include <vector>

// Copy-on-write vector.
template <class T>
class CowVec {
  std::shared_ptr<std::vector<T>> p_;

 public:
  size_t size() const { return p_ ? p_->size() : 0; }
  bool empty() const { return size() == 0; }
  const T &operator[](size_t i) const { return (*p_)[i]; }
  const T *data() const { return p_ ? p_->data() : nullptr; }
  const T *begin() const { return data(); }
  const T *end() const { return data() + size(); }
  const T &back() const { return p_->back(); }
  void clear() { p_.reset(); }
  // the vector, detached from every other holder (copied when shared)
  std::vector<T> &mut() {
    if (!p_) p_ = std::make_shared<std::vector<T>>();
    else if (p_.use_count() > 1) p_ = std::make_shared<std::vector<T>>(*p_);
    return *p_;
  }
  T &w(size_t i) { return mut()[i]; }
  void push_back(const T &v) { mut().push_back(v); }
  // a detached vector of n elements whose old contents are NOT kept (full overwrite follows)
  std::vector<T> &fresh(size_t n) {
    if (!p_ || p_.use_count() > 1) p_ = std::make_shared<std::vector<T>>(n);
    else p_->resize(n);
    return *p_;
  }
  void assign(const T *a, const T *b) {
    std::vector<T> &v = fresh((size_t)(b - a));
    std::copy(a, b, v.begin());
  }
};

