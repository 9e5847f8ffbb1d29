// Force-included (-include) when compiling the reference's DCN sources against torch 2.11.
// The reference dispatches with AT_DISPATCH_FLOATING_TYPES(input.type(), ...) (modulated_deform_conv_cuda.cu:93);
// the macro's helper no longer accepts the deprecated type-properties object, so the overload is restored here.
#pragma once
#include <ATen/ATen.h>
#include <ATen/Dispatch.h>
namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) { return t.scalarType(); }
}
