// TEST INFRASTRUCTURE ONLY.  Force-included (-include) when oracle/Makefile compiles the reference's
// UNMODIFIED fbgemm_gpu/src/sparse_ops_cpu.cpp against this image's torch 2.11 headers.  The 2021
// source passes `tensor.type()` (DeprecatedTypeProperties) to AT_DISPATCH_* and names
// optTypeMetaToScalarType without its namespace; both still exist in torch, only the glue is gone.
#pragma once
#include <ATen/ATen.h>
#include <ATen/Dispatch.h>

namespace detail {
inline c10::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) { return t.scalarType(); }
}  // namespace detail
using c10::optTypeMetaToScalarType;
