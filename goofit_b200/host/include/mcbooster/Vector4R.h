// mcbooster/Vector4R.h — a four-vector (E, px, py, pz) with the accessors user code calls on the particles GenerateSig returns
// (see GTypes.h for what this directory is)
#pragma once
#include <cmath>
#include <mcbooster/GTypes.h>

namespace mcbooster {
class Vector4R {
  public:
    Vector4R() = default;
    Vector4R(GReal_t e, GReal_t px, GReal_t py, GReal_t pz)
        : v_{e, px, py, pz} {}
    void set(GReal_t e, GReal_t px, GReal_t py, GReal_t pz) {
        v_[0] = e;
        v_[1] = px;
        v_[2] = py;
        v_[3] = pz;
    }
    void set(int i, GReal_t d) { v_[i] = d; }
    GReal_t get(int i) const { return v_[i]; }
    GReal_t mass2() const { return v_[0] * v_[0] - v_[1] * v_[1] - v_[2] * v_[2] - v_[3] * v_[3]; }
    GReal_t mass() const {
        const GReal_t m2 = mass2();
        return m2 > 0 ? std::sqrt(m2) : 0.0;
    }
    GReal_t d3mag() const { return std::sqrt(v_[1] * v_[1] + v_[2] * v_[2] + v_[3] * v_[3]); }
    Vector4R operator+(const Vector4R &o) const { return {v_[0] + o.v_[0], v_[1] + o.v_[1], v_[2] + o.v_[2], v_[3] + o.v_[3]}; }
    Vector4R operator-(const Vector4R &o) const { return {v_[0] - o.v_[0], v_[1] - o.v_[1], v_[2] - o.v_[2], v_[3] - o.v_[3]}; }

  private:
    GReal_t v_[4]{0, 0, 0, 0};
};
} // namespace mcbooster
