// cow_vec.hpp -- copy-on-write array shared by the problem object (lp_model.hpp) and the engine's
// host mirror (engine_cuda.cu): every open B&B node is a copy of its parent (util.cpp:25-42), and
// a copy costs a reference count until one side writes.  One heap block per array (header +
// elements): a tree of 10^5 open nodes is millions of these, so the number of allocations counts.
#pragma once

#include <atomic>
#include <cstddef>
#include <cstdlib>
#include <new>
#include <type_traits>
#include <utility>

template <class T>
class CowVec {
  struct alignas(16) Hdr {
    std::atomic<int> rc;
    size_t n, cap;
  };
  Hdr *h_ = nullptr;

  static T *elems(Hdr *h) { return reinterpret_cast<T *>(h + 1); }
  static Hdr *alloc(size_t cap) {
    Hdr *h = static_cast<Hdr *>(std::malloc(sizeof(Hdr) + (cap ? cap : 1) * sizeof(T)));
    if (!h) throw std::bad_alloc();
    new (&h->rc) std::atomic<int>(1);
    h->n = 0;
    h->cap = cap ? cap : 1;
    return h;
  }
  static void release(Hdr *h) {
    if (h && h->rc.fetch_sub(1, std::memory_order_acq_rel) == 1) {
      if (!std::is_trivially_destructible<T>::value)
        for (size_t i = 0; i < h->n; i++) elems(h)[i].~T();
      std::free(h);
    }
  }
  // a block of capacity cap holding copies of the first keep elements of this one
  Hdr *clone(size_t cap, size_t keep) const {
    Hdr *h = alloc(cap);
    const T *src = h_ ? elems(h_) : nullptr;
    for (size_t i = 0; i < keep; i++) new (elems(h) + i) T(src[i]);
    h->n = keep;
    return h;
  }
  void detach(size_t min_cap) {
    const size_t n = size();
    if (!h_) { h_ = alloc(min_cap); return; }
    if (h_->rc.load(std::memory_order_acquire) > 1 || h_->cap < min_cap) {
      Hdr *h = clone(min_cap > n ? min_cap : n, n);
      release(h_);
      h_ = h;
    }
  }

 public:
  // what mut() hands out: the detached elements
  struct Span {
    T *p;
    size_t n;
    T *begin() const { return p; }
    T *end() const { return p + n; }
    T *data() const { return p; }
    size_t size() const { return n; }
    T &operator[](size_t i) const { return p[i]; }
  };

  CowVec() = default;
  CowVec(const CowVec &o) : h_(o.h_) {
    if (h_) h_->rc.fetch_add(1, std::memory_order_relaxed);
  }
  CowVec(CowVec &&o) noexcept : h_(o.h_) { o.h_ = nullptr; }
  CowVec &operator=(const CowVec &o) {
    if (o.h_) o.h_->rc.fetch_add(1, std::memory_order_relaxed);
    release(h_);
    h_ = o.h_;
    return *this;
  }
  CowVec &operator=(CowVec &&o) noexcept {
    if (this != &o) {
      release(h_);
      h_ = o.h_;
      o.h_ = nullptr;
    }
    return *this;
  }
  ~CowVec() { release(h_); }

  size_t size() const { return h_ ? h_->n : 0; }
  bool empty() const { return size() == 0; }
  const T &operator[](size_t i) const { return elems(h_)[i]; }
  const T *data() const { return h_ ? elems(h_) : nullptr; }
  const T *begin() const { return data(); }
  const T *end() const { return data() + size(); }
  const T &back() const { return elems(h_)[h_->n - 1]; }
  void clear() {
    release(h_);
    h_ = nullptr;
  }
  // the elements, detached from every other holder (copied when shared)
  Span mut() {
    detach(size());
    return Span{elems(h_), h_->n};
  }
  T &w(size_t i) {
    detach(size());
    return elems(h_)[i];
  }
  void push_back(const T &v) {
    const size_t n = size();
    const T copy(v);  // v may live in the block that is about to move
    if (!h_ || h_->rc.load(std::memory_order_acquire) > 1 || h_->cap < n + 1) detach(n + 1 > 2 * n ? n + 1 : 2 * n);
    new (elems(h_) + n) T(copy);
    h_->n = n + 1;
  }
  void resize(size_t n, const T &v = T()) {
    detach(n);
    T *e = elems(h_);
    for (size_t i = h_->n; i < n; i++) new (e + i) T(v);
    if (!std::is_trivially_destructible<T>::value)
      for (size_t i = n; i < h_->n; i++) e[i].~T();
    h_->n = n;
  }
  // n elements, detached, whose old contents are NOT kept (a full overwrite follows); trivial
  // types are left uninitialised
  T *fresh(size_t n) {
    if (h_ && h_->rc.load(std::memory_order_acquire) == 1 && h_->cap >= n) {
      resize(n);
      return elems(h_);
    }
    release(h_);
    h_ = alloc(n);
    if (!std::is_trivially_default_constructible<T>::value)
      for (size_t i = 0; i < n; i++) new (elems(h_) + i) T();
    h_->n = n;
    return elems(h_);
  }
  void assign(const T *a, const T *b) {
    T *d = fresh((size_t)(b - a));
    for (const T *p = a; p != b; ++p, ++d) *d = *p;
  }
};
