// utilities/Uncertain.h — the value-with-uncertainty helper user code prints fit results with (include/goofit/utilities/Uncertain.h of
// the reference, where it wraps github.com/clausjensbymadsen/uncertain-values): same members and operators, written against
// <iostream> alone so that it needs no {fmt}. Sums and differences add the absolute uncertainties, products and quotients the
// relative ones (passed on, as upstream does, as the uncertainty of the result); a number scales both.
#pragma once
#include <cmath>
#include <cstdio>
#include <iosfwd>
#include <ostream>
#include <stdexcept>

#include <goofit/detail/Core.h>

namespace GooFit {

class Uncertain {
  public:
    Uncertain(fptype value, fptype uncertainty = 0)
        : val_(value)
        , unc_(uncertainty) {
        if(unc_ < 0) throw std::invalid_argument("Uncertainty can not be negative.");
    }
    Uncertain(const Variable &v)
        : Uncertain(v.getValue(), v.getError()) {}

    fptype get_value() const { return val_; }
    fptype get_uncertainty() const { return unc_; }
    fptype get_relative_uncertainty() const { return unc_ / val_; }

    bool operator==(const Uncertain &o) const { return val_ == o.val_ && unc_ == o.unc_; }
    bool operator!=(const Uncertain &o) const { return !(*this == o); }

    Uncertain operator+(const Uncertain &o) const { return {val_ + o.val_, unc_ + o.unc_}; }
    Uncertain operator-(const Uncertain &o) const { return {val_ - o.val_, unc_ + o.unc_}; }
    Uncertain operator*(const Uncertain &o) const { return {val_ * o.val_, get_relative_uncertainty() + o.get_relative_uncertainty()}; }
    Uncertain operator/(const Uncertain &o) const { return {val_ / o.val_, get_relative_uncertainty() + o.get_relative_uncertainty()}; }
    Uncertain operator*(fptype f) const { return {val_ * f, unc_ * f}; }
    Uncertain operator/(const fptype &f) const { return {val_ / f, unc_ / f}; }

  private:
    fptype val_, unc_;
};

inline Uncertain operator*(fptype f, const Uncertain &u) { return {u.get_value() * f, u.get_uncertainty() * f}; }

/// "value +/- uncertainty" with as many decimals as the uncertainty's leading digits need (two significant digits up to 2.5, none above)
inline std::ostream &operator<<(std::ostream &os, Uncertain u) {
    const fptype v = u.get_value(), e = u.get_uncertainty();
    char buf[96];
    if(e > 0 && e < 2.5) {
        int digits = (int)std::ceil(-std::log10(e) + std::log10(2.5));
        std::snprintf(buf, sizeof buf, "%.*f +/- %.*f", digits, v, digits, e);
        return os << buf;
    }
    if(e >= 2.5) {
        std::snprintf(buf, sizeof buf, "%.0f +/- %.0f", v, e);
        return os << buf;
    }
    return os << v << " +/- " << e;
}

} // namespace GooFit

#if defined(FMT_VERSION) && FMT_VERSION >= 90000
#include <fmt/ostream.h>
template <>
struct fmt::formatter<GooFit::Uncertain> : fmt::ostream_formatter {};
#endif
