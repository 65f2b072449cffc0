// TEST INFRASTRUCTURE (oracle): Lorentz 4-vector with the EvtGen-style interface the reference
// calls (set(E,px,py,pz), get(0)=E; operator* = Minkowski product; dot/cross/d3mag spatial).
// Call sites: /root/reference/include/goofit/PDFs/physics/detail/Dim5.h:22-110, Dim2.h:22-35,
// /root/reference/include/goofit/PDFs/physics/UserUtils.h:18-21.
#pragma once
#include <cmath>
#include <mcbooster/GTypes.h>
#ifndef __host__
#define __host__
#define __device__
#endif
namespace mcbooster {
class Vector4R {
  public:
    __host__ __device__ Vector4R() : v{0, 0, 0, 0} {}
    __host__ __device__ Vector4R(GReal_t e, GReal_t px, GReal_t py, GReal_t pz) : v{e, px, py, pz} {}
    __host__ __device__ void set(int i, GReal_t d) { v[i] = d; }
    __host__ __device__ void set(GReal_t e, GReal_t px, GReal_t py, GReal_t pz) {
        v[0] = e; v[1] = px; v[2] = py; v[3] = pz;
    }
    __host__ __device__ GReal_t get(int i) const { return v[i]; }
    __host__ __device__ Vector4R &operator*=(GReal_t c) { v[0] *= c; v[1] *= c; v[2] *= c; v[3] *= c; return *this; }
    __host__ __device__ Vector4R &operator/=(GReal_t c) {
        GReal_t ci = 1.0 / c; v[0] *= ci; v[1] *= ci; v[2] *= ci; v[3] *= ci; return *this;
    }
    __host__ __device__ Vector4R &operator+=(const Vector4R &o) { for(int i = 0; i < 4; i++) v[i] += o.v[i]; return *this; }
    __host__ __device__ Vector4R &operator-=(const Vector4R &o) { for(int i = 0; i < 4; i++) v[i] -= o.v[i]; return *this; }
    __host__ __device__ GReal_t mass2() const { return v[0] * v[0] - v[1] * v[1] - v[2] * v[2] - v[3] * v[3]; }
    __host__ __device__ GReal_t mass() const { GReal_t m2 = mass2(); return m2 > 0.0 ? sqrt(m2) : 0.0; }
    __host__ __device__ GReal_t d3mag() const { return sqrt(v[1] * v[1] + v[2] * v[2] + v[3] * v[3]); }
    __host__ __device__ GReal_t dot(const Vector4R &o) const { return v[1] * o.v[1] + v[2] * o.v[2] + v[3] * o.v[3]; }
    __host__ __device__ Vector4R cross(const Vector4R &o) const {
        return Vector4R(0.0, v[2] * o.v[3] - v[3] * o.v[2], v[3] * o.v[1] - v[1] * o.v[3], v[1] * o.v[2] - v[2] * o.v[1]);
    }
    // boost this vector by the velocity of p4 (standard active boost)
    __host__ __device__ void applyBoostTo(const Vector4R &p4, bool inverse = false) {
        GReal_t e = p4.get(0);
        GReal_t s = inverse ? -1.0 : 1.0;
        applyBoostTo(s * p4.get(1) / e, s * p4.get(2) / e, s * p4.get(3) / e);
    }
    __host__ __device__ void applyBoostTo(GReal_t bx, GReal_t by, GReal_t bz) {
        GReal_t b2 = bx * bx + by * by + bz * bz;
        if(b2 <= 0.0) return;
        GReal_t gamma = 1.0 / sqrt(1.0 - b2);
        GReal_t bp = bx * v[1] + by * v[2] + bz * v[3];
        GReal_t g2 = (gamma - 1.0) / b2;
        v[1] = v[1] + g2 * bp * bx + gamma * bx * v[0];
        v[2] = v[2] + g2 * bp * by + gamma * by * v[0];
        v[3] = v[3] + g2 * bp * bz + gamma * bz * v[0];
        v[0] = gamma * (v[0] + bp);
    }
  private:
    GReal_t v[4];
};
__host__ __device__ inline Vector4R operator+(const Vector4R &a, const Vector4R &b) { return Vector4R(a) += b; }
__host__ __device__ inline Vector4R operator-(const Vector4R &a, const Vector4R &b) { return Vector4R(a) -= b; }
__host__ __device__ inline Vector4R operator*(GReal_t c, const Vector4R &a) { return Vector4R(a) *= c; }
__host__ __device__ inline Vector4R operator*(const Vector4R &a, GReal_t c) { return Vector4R(a) *= c; }
__host__ __device__ inline Vector4R operator/(const Vector4R &a, GReal_t c) { return Vector4R(a) /= c; }
__host__ __device__ inline GReal_t operator*(const Vector4R &a, const Vector4R &b) {
    return a.get(0) * b.get(0) - a.get(1) * b.get(1) - a.get(2) * b.get(2) - a.get(3) * b.get(3);
}
} // namespace mcbooster
