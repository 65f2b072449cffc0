// gf_math.cuh — small arithmetic helpers shared by the three- and four-body cache kernels
#pragma once
// Branch-free reciprocal for the cache kernels (gf_pair_prepare / gf_rbw_term of the three-body waves, the Breit-Wigner family of the four-body lineshapes): the hardware seed (MUFU.RCP64H, ~2^-20) and one
// cubic Newton step, e = 1 - x r, r <- r (1 + e + e^2): relative error ~e^3 + rounding, i.e. within an ulp or two of the
// correctly rounded quotient, for finite non-zero normal x - which every denominator on that path is (s > 0 inside the Dalitz
// boundary, 1 + z^2, A^2 + B^2). The IEEE division of `a / b` costs the same arithmetic plus a range check, a reconvergence
// region and a slow-path call per division; with ~80 divisions per event those, not the FP64 pipe, bounded k_amp3_waves.
#ifndef GF_FAST_RCP
#define GF_FAST_RCP 1
#endif
__device__ __forceinline__ double gf_rcp(double x) {
#if !GF_FAST_RCP
    return 1.0 / x;
#endif
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    const double t = fma(e, e, e);
    r = fma(r, t, r);
    const double e2 = fma(-x, r, 1.0); // one linear step more: the seed is only guaranteed to ~2^-20
    return fma(r, e2, r);
}
