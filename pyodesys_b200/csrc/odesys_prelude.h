// Shared prelude of every generated ODE-system header.
//
// Provides the few names the generated ``struct odesys_anyode::OdeSys`` (and user supplied
// ``namespace_override`` snippets written for the reference's template) rely on:
//   AnyODE::Status   - callback return codes, same values as
//                      external/anyode/include/anyode/anyode.hpp:125
//   AnyODE::ignore   - silences unused-argument warnings (anyode.hpp:20-22)
// Compiles under nvcc (device + host) and under plain g++ (CPU oracle).
#pragma once
#include <math.h>
#ifdef __CUDACC__
#define ODESYS_FN __host__ __device__ __forceinline__
#define ODESYS_RESTRICT __restrict__
#else
#define ODESYS_FN inline
#define ODESYS_RESTRICT __restrict__
#endif

namespace AnyODE {
enum Status : int { success = 0, recoverable_error = 1, unrecoverable_error = -1 };
template <class T> ODESYS_FN void ignore(const T &) {}
}  // namespace AnyODE

template <int N> ODESYS_FN double odesys_ipow(double x) {
    if constexpr (N == 0) {
        return 1.0;
    } else if constexpr (N == 1) {
        return x;
    } else if constexpr (N % 2 == 0) {
        const double h = odesys_ipow<N / 2>(x);
        return h * h;
    } else {
        return x * odesys_ipow<N - 1>(x);
    }
}
