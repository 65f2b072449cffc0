// TEST INFRASTRUCTURE (oracle): accept/reject flag functor (index, weight) -> bool with a
// counter-based uniform draw. Call sites: /root/reference/src/PDFs/physics/Amp4BodyBase.cu:17-21.
#pragma once
#include <mcbooster/GTypes.h>
namespace mcbooster {
__host__ __device__ inline uint64_t shim_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
__host__ __device__ inline double shim_uniform(uint64_t seed, uint64_t idx, uint64_t stream) {
    uint64_t h = shim_splitmix64(shim_splitmix64(seed ^ (stream * 0xD1B54A32D192ED03ULL)) + idx);
    return (h >> 11) * (1.0 / 9007199254740992.0);
}
struct FlagAcceptReject {
    GReal_t wmax;
    FlagAcceptReject(GReal_t m) : wmax(m) {}
    __host__ __device__ GBool_t operator()(GLong_t idx, GReal_t weight) const {
        return weight > wmax * shim_uniform(0x5eedULL, static_cast<uint64_t>(idx), 77);
    }
};
} // namespace mcbooster
