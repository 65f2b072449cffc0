// MathUtils.h - the reference's compensated host-side sum (goofit/MathUtils.h:7-19, src/goofit/MathUtils.cpp:11-30): Neumaier's
// variant of Kahan summation, the correction applied once at the end.
#pragma once
#include <cmath>
#include <vector>

#include <goofit/detail/Core.h>

namespace GooFit {
class MathUtils final {
  public:
    static fptype doNeumaierSummation(const std::vector<fptype> &vals) {
        fptype sum = 0.0, lost = 0.0;
        for(fptype v : vals) {
            const fptype t = sum + v;
            // whichever of the two operands is smaller in magnitude loses its low-order bits in t
            lost += std::abs(sum) >= std::abs(v) ? (sum - t) + v : (v - t) + sum;
            sum = t;
        }
        return vals.size() == 1 ? vals[0] : sum + lost;
    }
};
} // namespace GooFit
