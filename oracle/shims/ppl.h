// TEST INFRASTRUCTURE ONLY (oracle build shim) -- not part of the product.
// Serial stand-in for the MSVC Parallel Patterns Library used by the reference
// (VolumeKdtree_recover.h:16,19; call sites VolumeKdTree_recover.cpp:175,243-251,294,338,607).
// Task scheduling only; running the tasks in order gives the same results because every
// active parallel region writes disjoint outputs.
#pragma once
#include <functional>
namespace concurrency {
template <typename F1, typename F2> inline void parallel_invoke(const F1& a, const F2& b) { a(); b(); }
template <typename I, typename F> inline void parallel_for(I first, I last, const F& f) {
  for (I i = first; i < last; ++i) f(i);
}
template <typename T> class combinable {
  T v_;
 public:
  template <typename Init> explicit combinable(Init init) : v_(init()) {}
  T& local() { return v_; }
  template <typename Op> T combine(Op) { return v_; }
};
}  // namespace concurrency
