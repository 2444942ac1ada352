// Force-included (-include) ahead of every reference translation unit by oracle/build_ref.py.
// TEST INFRASTRUCTURE ONLY.  The reference sources are compiled where they lie under
// /root/reference, unmodified; this header only restores an overload that torch >= 2.x dropped:
// the sources call AT_DISPATCH_*(tensor.type(), ...) (src/correlation_kernels.cu:145,176,
// src/altcorr_kernel.cu:309, thirdparty/lietorch/lietorch/include/dispatch.h:39), and the
// dispatch macro resolves the type through ::detail::scalar_type(), which no longer accepts
// the deprecated at::DeprecatedTypeProperties that Tensor::type() returns.
#pragma once
#include <torch/extension.h>

namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) { return t.scalarType(); }
}  // namespace detail
