// TEST INFRASTRUCTURE (oracle): deterministic N-body phase-space generator with the PhaseSpace
// interface the reference calls (Amp4Body.cu:243-252, NormEvents_4Body_Base.cu:32-41,
// Amp3Body.cu:440-450, Amp4Body_TD.cu:831-842). Algorithm: Raubold-Lynch / GENBOD sequential
// two-body decays with sorted uniforms; one counter-based RNG stream per event so the result is
// independent of threading. NOT MCBooster's generator ("parity unpinned" for the generator).
#pragma once
#include <algorithm>
#include <cmath>
#include <mcbooster/GContainers.h>
#include <mcbooster/functors/FlagAcceptReject.h>
#include <stdexcept>
#include <vector>
namespace mcbooster {
class PhaseSpace {
  public:
    PhaseSpace(GReal_t mother, const std::vector<GReal_t> &masses, GLong_t nevents, GLong_t seed_or_offset = 0)
        : fMother(mother), fMasses(masses), fN(nevents), fSeed(static_cast<uint64_t>(seed_or_offset)), fNAcc(0) {
        if(masses.size() < 2 || masses.size() > 8) throw std::runtime_error("PhaseSpace shim: 2..8 daughters");
        GReal_t s = 0; for(auto m : masses) s += m;
        if(s >= mother) throw std::runtime_error("PhaseSpace shim: not enough energy");
        fDaughters.resize(masses.size());
    }
    void SetSeed(GLong_t s) { fSeed = static_cast<uint64_t>(s); }
    GLong_t GetSeed() const { return static_cast<GLong_t>(fSeed); }
    GLong_t GetNAccepted() const { return fNAcc; }
    const BoolVector_d &GetAccRejFlags() const { return fFlags; }
    Particles_d &GetDaughters(int i) { return fDaughters[i]; }
    const RealVector_d &GetWeights() const { return fWeights; }
    void FreeResources() { for(auto &d : fDaughters) { d.clear(); d.shrink_to_fit(); } fWeights.clear(); fFlags.clear(); }

    static GReal_t pdk(GReal_t a, GReal_t b, GReal_t c) {
        GReal_t x = (a - b - c) * (a + b + c) * (a + b - c) * (a - b + c);
        return sqrt(x > 0 ? x : 0) / (2 * a);
    }
    void Generate(const Vector4R &mother) {
        const int nd = static_cast<int>(fMasses.size());
        for(auto &d : fDaughters) d.resize(fN);
        fWeights.resize(fN);
        std::vector<Vector4R *> dp(nd);
        for(int k = 0; k < nd; k++) dp[k] = thrust::raw_pointer_cast(fDaughters[k].data());
        GReal_t *wp = thrust::raw_pointer_cast(fWeights.data());
        GReal_t tecm = mother.mass(); for(auto m : fMasses) tecm -= m;
        // maximum weight: product of the largest possible two-body momenta
        GReal_t emmax = tecm + fMasses[0], emmin = 0, wtmax = 1;
        for(int n = 1; n < nd; n++) { emmin += fMasses[n - 1]; emmax += fMasses[n]; wtmax *= pdk(emmax, emmin, fMasses[n]); }
        fWtMax = wtmax;
        const uint64_t seed = fSeed;
        const std::vector<GReal_t> mass = fMasses;
#pragma omp parallel for
        for(long ev = 0; ev < static_cast<long>(fN); ev++) {
            GReal_t rno[8], invMas[8], pd[8];
            uint64_t ctr = 0;
            rno[0] = 0;
            for(int n = 1; n < nd - 1; n++) rno[n] = shim_uniform(seed, static_cast<uint64_t>(ev), ctr++);
            std::sort(rno + 1, rno + nd - 1);
            rno[nd - 1] = 1;
            GReal_t sum = 0;
            for(int n = 0; n < nd; n++) { sum += mass[n]; invMas[n] = rno[n] * tecm + sum; }
            GReal_t wt = 1.0 / wtmax;
            for(int n = 0; n < nd - 1; n++) { pd[n] = pdk(invMas[n + 1], invMas[n], mass[n + 1]); wt *= pd[n]; }
            Vector4R p[8];
            p[0].set(sqrt(pd[0] * pd[0] + mass[0] * mass[0]), 0, pd[0], 0);
            int i = 1;
            while(true) {
                p[i].set(sqrt(pd[i - 1] * pd[i - 1] + mass[i] * mass[i]), 0, -pd[i - 1], 0);
                GReal_t cZ = 2 * shim_uniform(seed, ev, ctr++) - 1, sZ = sqrt(1 - cZ * cZ);
                GReal_t angY = 2 * M_PI * shim_uniform(seed, ev, ctr++), cY = cos(angY), sY = sin(angY);
                for(int j = 0; j <= i; j++) {
                    GReal_t x = p[j].get(1), y = p[j].get(2);
                    p[j].set(1, cZ * x - sZ * y); p[j].set(2, sZ * x + cZ * y); // rotate about z
                    x = p[j].get(1); GReal_t z = p[j].get(3);
                    p[j].set(1, cY * x - sY * z); p[j].set(3, sY * x + cY * z); // rotate about y
                }
                if(i == nd - 1) break;
                GReal_t beta = pd[i] / sqrt(pd[i] * pd[i] + invMas[i] * invMas[i]);
                for(int j = 0; j <= i; j++) p[j].applyBoostTo(0, beta, 0);
                i++;
            }
            for(int n = 0; n < nd; n++) { p[n].applyBoostTo(mother); dp[n][ev] = p[n]; }
            wp[ev] = wt;
        }
    }
    void Unweight() {
        fFlags.resize(fN);
        GLong_t acc = 0;
        for(GLong_t ev = 0; ev < fN; ev++) {
            bool f = fWeights[ev] > shim_uniform(fSeed, static_cast<uint64_t>(ev), 1000);
            fFlags[ev] = f; acc += f;
        }
        fNAcc = acc;
    }
  private:
    GReal_t fMother; std::vector<GReal_t> fMasses; GLong_t fN; uint64_t fSeed; GLong_t fNAcc; GReal_t fWtMax{1};
    std::vector<Particles_d> fDaughters; RealVector_d fWeights; BoolVector_d fFlags;
};
} // namespace mcbooster
