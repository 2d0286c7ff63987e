// Test-infrastructure stub (NOT product code): the reference's util.h:5 includes
// <blaze/Math.h> only for torch_to_blaze(); blaze is absent on this image and the
// graph game never calls it.  This minimal stand-in lets util.h parse.
#pragma once
#include <cstddef>
namespace blaze {
template <class T> class DynamicVector {
 public:
  DynamicVector(std::size_t, const T*) {}
};
}  // namespace blaze
