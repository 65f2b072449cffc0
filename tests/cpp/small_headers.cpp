// host-only check of the small user-facing headers: goofit/utilities/Uncertain.h, goofit/detail/Complex.h, goofit/detail/Abort.h
#include <goofit/Variable.h>
#include <goofit/detail/Abort.h>
#include <goofit/detail/Complex.h>
#include <goofit/PDFs/physics/Amp3Body.h>
#include <goofit/utilities/Uncertain.h>

#include <cmath>
#include <sstream>

using namespace GooFit;

int main() {
    using namespace GooFit::literals::complex_literals;
    fpcomplex z = 2.0 + 3_i;
    if(z.real() != 2.0 || z.imag() != 3.0) return 1;
    if((1.5_i).imag() != 1.5) return 2;

    Variable v("v", 1.2345, 0.0123, 0, 2);
    Uncertain u(v);
    if(u.get_value() != 1.2345 || u.get_uncertainty() != 0.0123) return 3;
    std::ostringstream a;
    a << u; // ceil(-log10(0.0123) + log10(2.5)) = 3 decimals
    if(a.str() != "1.234 +/- 0.012" && a.str() != "1.235 +/- 0.012") return 4;
    std::ostringstream b;
    b << Uncertain(1234.4, 25.0);
    if(b.str() != "1234 +/- 25") return 5;
    Uncertain s = Uncertain(1.0, 0.1) + Uncertain(2.0, 0.2);
    if(std::fabs(s.get_value() - 3.0) > 1e-15 || std::fabs(s.get_uncertainty() - 0.3) > 1e-15) return 6;
    Uncertain p = Uncertain(2.0, 0.2) * Uncertain(4.0, 0.4); // relative uncertainties add: 0.1 + 0.1
    if(p.get_value() != 8.0 || std::fabs(p.get_uncertainty() - 0.2) > 1e-15) return 7;
    Uncertain k = 100 * Uncertain(0.5, 0.01);
    if(k.get_value() != 50.0 || std::fabs(k.get_uncertainty() - 1.0) > 1e-15) return 8;
    if(!(Uncertain(1, 1) == Uncertain(1, 1)) || !(Uncertain(1, 1) != Uncertain(1, 2))) return 9;
    try {
        Uncertain bad(1.0, -1.0);
        return 10;
    } catch(const std::invalid_argument &) {
    }
    try {
        GooFit::abort(__FILE__, __LINE__, "checked on purpose");
        return 11;
    } catch(const GeneralError &) {
    }
    // GenerateSig's return type: flat columns -> upstream's tuple of MCBooster host containers (the receiver owns the sets)
    GeneratedSignal G;
    G.particles = {{1, 2, 3, 4, 5, 6, 7, 8}, {9, 10, 11, 12, 13, 14, 15, 16}};
    G.variables = {{0.1, 0.2}, {0.3, 0.4}, {0.5, 0.6}};
    G.weights   = {0.7, 0.8};
    G.flags     = {true, false};
    mcbooster::ParticlesSet_h particles;
    mcbooster::VariableSet_h variables;
    mcbooster::RealVector_h weights;
    mcbooster::BoolVector_h flags;
    std::tie(particles, variables, weights, flags) = toMCBoosterContainers<mcbooster::BoolVector_h>(G);
    if(particles.size() != 2 || variables.size() != 3 || weights.size() != 2 || flags.size() != 2) return 12;
    if((*particles[1])[1].get(0) != 13 || (*particles[1])[1].get(3) != 16 || (*particles[0])[0].get(1) != 2) return 13;
    if((*variables[2])[1] != 0.6 || weights[0] != 0.7 || !flags[0] || flags[1]) return 14;
    for(auto *p : particles) delete p;
    for(auto *v : variables) delete v;
    auto asReal = toMCBoosterContainers<mcbooster::RealVector_h>(G); // Amp3Body / Amp4Body: the flags as 0 / 1 reals
    if(std::get<3>(asReal)[0] != 1.0 || std::get<3>(asReal)[1] != 0.0) return 15;
    for(auto *p : std::get<0>(asReal)) delete p;
    for(auto *v : std::get<1>(asReal)) delete v;
    return 0;
}
