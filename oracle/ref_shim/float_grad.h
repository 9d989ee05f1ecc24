// Stand-in for the MISSING third-party header `float_grad.h`
// (rogerhh/auto-jvp-example @ 06f0e55, a private submodule absent from the
// reference archive -- SURVEY.md F1).  It only declares the NAMES that the
// reference's non-JVP translation units still mention, so that the reference's
// own primal / VJP / binning kernels can be compiled, unmodified and in place,
// into oracle/_ref/ as a GPU-side parity oracle.  No arithmetic lives here; the
// JVP kernels of the reference stay unbuildable.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <type_traits>
#include <cuda_runtime.h>
template <typename T> struct is_float_grad : std::false_type {};
template <typename T> struct FloatGrad {
    T d, g;
    __host__ __device__ FloatGrad() {}
    __host__ __device__ FloatGrad(T x) : d(x), g() {}
};
template <typename T> struct FloatGradArray { T* data_arr_ = nullptr; T* grad_arr_ = nullptr; };
