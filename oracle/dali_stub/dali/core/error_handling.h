// TEST INFRASTRUCTURE (oracle/). Minimal stand-in for NVIDIA DALI's header of the same name — written for this repo,
// NOT a copy of DALI — so that the reference's operator source
//   /root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc (+ .h)
// compiles UNMODIFIED, from where it lies, into oracle/_ref/libdg_ref.so (DALI itself is not installed here).
// Only the API surface that file touches exists; see oracle/dali_stub/dali/pipeline/operator/operator.h.
#pragma once
#include <stdexcept>
#include <string>
#define DALI_FAIL(msg) throw std::runtime_error(std::string(msg))
