// TEST INFRASTRUCTURE (oracle): functor base the reference's Dim2/Dim5 derive from.
#pragma once
#include <mcbooster/GTypes.h>
#include <mcbooster/Vector4R.h>
namespace mcbooster {
struct IFunctionArray {
    GInt_t dim{0};
    __host__ __device__ virtual void operator()(const GInt_t n, Vector4R **particles, GReal_t *variables) = 0;
    virtual ~IFunctionArray() = default;
};
} // namespace mcbooster
