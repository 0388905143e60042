// Test stub: the four fields of pangolin::Image<T> the hot path touches
// (reference thirdparty/Pangolin/include/pangolin/image/image.h:420-423).
#pragma once
#include <cstddef>
#include <cstdint>
namespace pangolin {
template <class T>
struct ManagedImage {
  size_t pitch = 0;
  T* ptr = nullptr;
  size_t w = 0, h = 0;
};
}  // namespace pangolin
