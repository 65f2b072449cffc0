// tests/ref_root stand-in (see TH1.h): the random-number interface of TRandom on top of std::mt19937_64. The streams are NOT ROOT's
// (TRandom is an LCG, TRandom3 MT19937 with its own seeding): the examples only draw toy data from them, and every comparison with
// the reference in this repo feeds both sides the same events from a file.
#pragma once
#include <cmath>
#include <random>
class TRandom {
  public:
    explicit TRandom(unsigned long seed = 65539)
        : g_(seed) {}
    virtual ~TRandom() = default;
    void SetSeed(unsigned long seed = 0) { g_.seed(seed ? seed : std::random_device{}()); }
    double Rndm() { return std::generate_canonical<double, 53>(g_); }
    double Uniform(double hi = 1.0) { return hi * Rndm(); }
    double Uniform(double lo, double hi) { return lo + (hi - lo) * Rndm(); }
    double Gaus(double mean = 0.0, double sigma = 1.0) { return std::normal_distribution<double>(mean, sigma)(g_); }
    double Exp(double tau) { return -tau * std::log(1.0 - Rndm()); }
    int Poisson(double mean) { return std::poisson_distribution<int>(mean)(g_); }
    unsigned int Integer(unsigned int imax) { return (unsigned int)(imax * Rndm()); }
    /// Landau-distributed number by inversion of a tabulated-free approximation is beyond a stand-in: the Moyal approximation
    double Landau(double mpv = 0.0, double sigma = 1.0) {
        double z = std::normal_distribution<double>(0.0, 1.0)(g_);
        return mpv + sigma * (-std::log(z * z + 1e-300)); // a Moyal variate: -log(chi2_1)
    }

  protected:
    std::mt19937_64 g_;
};
inline TRandom *gRandomStandIn() {
    static TRandom r;
    return &r;
}
#define gRandom (gRandomStandIn())
