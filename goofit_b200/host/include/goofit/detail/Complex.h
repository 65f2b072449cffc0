// detail/Complex.h — include/goofit/detail/Complex.h of the reference: the complex type of the host API and its `_i` literals. Upstream's
// fpcomplex is thrust::complex<fptype>; on the host side of this build it is std::complex<double> (detail/Core.h), same members.
#pragma once
#include <goofit/detail/Core.h>

namespace GooFit {
namespace literals {
namespace complex_literals {
inline fpcomplex operator"" _i(unsigned long long d) { return {0.0, static_cast<fptype>(d)}; }
inline fpcomplex operator"" _i(long double d) { return {0.0, static_cast<fptype>(d)}; }
} // namespace complex_literals
} // namespace literals
} // namespace GooFit
