// physics4.cpp — Python bindings of the four-body amplitude classes (reference: python/PDFs/physics/{Amp4Body.cu,
// Amplitude.cpp,Lineshapes.cpp,SpinFactors.cpp,DalitzPlotHelpers.cpp}).
#include <pybind11/complex.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <goofit/detail/Physics4.h>

namespace py = pybind11;
using namespace pybind11::literals;
using namespace GooFit;

template <class LS>
static void bind_ls(py::module &m, const char *name) {
    py::class_<LS, Lineshape>(m, name).def(
        py::init<std::string, Variable, Variable, unsigned int, unsigned int, FF, fptype>(), "name"_a, "mass"_a, "width"_a,
        "L"_a, "Mpair"_a, "FormFac"_a = FF::BL_Prime, "radius"_a = 1.5);
}

// (particles[nd, N, 4] as (E, px, py, pz), variables[nv, N], weights[N], flags[N]) -- python/PDFs/physics/Amp4Body.cu:36-74
py::tuple generated_to_numpy(const GeneratedSignal &G, int ncomp) {
    const size_t nd = G.particles.size(), nv = G.variables.size(), n = G.weights.size();
    py::array_t<fptype> particles({nd, n, (size_t)ncomp}), variables({nv, n}), weights((py::ssize_t)n);
    py::array_t<bool> flags((py::ssize_t)n);
    auto P = particles.mutable_unchecked<3>();
    auto V = variables.mutable_unchecked<2>();
    auto W = weights.mutable_unchecked<1>();
    auto F = flags.mutable_unchecked<1>();
    for(size_t d = 0; d < nd; d++)
        for(size_t i = 0; i < n; i++)
            for(int k = 0; k < ncomp; k++) P(d, i, k) = G.particles[d][4 * i + k];
    for(size_t v = 0; v < nv; v++)
        for(size_t i = 0; i < n; i++) V(v, i) = G.variables[v][i];
    for(size_t i = 0; i < n; i++) {
        W(i) = G.weights[i];
        F(i) = G.flags[i];
    }
    return py::make_tuple(particles, variables, weights, flags);
}

void init_physics4(py::module &m) {
    py::enum_<DP4Pair>(m, "DP4Pair")
        .value("M_12", M_12).value("M_34", M_34).value("M_13", M_13).value("M_14", M_14).value("M_23", M_23).value("M_24", M_24)
        .value("M_12_3", M_12_3).value("M_13_2", M_13_2).value("M_23_1", M_23_1).value("M_12_4", M_12_4).value("M_14_2", M_14_2)
        .value("M_24_1", M_24_1).value("M_13_4", M_13_4).value("M_14_3", M_14_3).value("M_34_1", M_34_1).value("M_23_4", M_23_4)
        .value("M_24_3", M_24_3).value("M_34_2", M_34_2)
        .export_values();
    py::enum_<FF>(m, "FF").value("One", FF::One).value("BL", FF::BL).value("BL_Prime", FF::BL_Prime).value("BL2", FF::BL2);
    py::class_<Lineshape, AmpComponent>(m, "Lineshape");
    auto mls = m.def_submodule("Lineshapes");
    bind_ls<Lineshapes::RBW>(mls, "RBW");
    bind_ls<Lineshapes::SBW>(mls, "SBW");
    bind_ls<Lineshapes::LASS>(mls, "LASS");
    bind_ls<Lineshapes::Bugg>(mls, "Bugg");
    bind_ls<Lineshapes::Bugg3>(mls, "Bugg3");
    bind_ls<Lineshapes::One>(mls, "One");
    bind_ls<Lineshapes::NonRes>(mls, "NonRes");
    bind_ls<Lineshapes::Flatte>(mls, "Flatte");
    py::class_<Lineshapes::GLASS, Lineshape>(mls, "GLASS")
        .def(py::init<std::string, Variable, Variable, unsigned int, unsigned int, FF, fptype, std::vector<Variable>>(), "name"_a, "mass"_a,
             "width"_a, "L"_a, "Mpair"_a, "FormFac"_a, "radius"_a, "AdditionalVars"_a);
    py::class_<Lineshapes::kMatrix, Lineshape>(mls, "kMatrix")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, Variable, Variable, std::vector<Variable>, std::vector<Variable>,
                      std::vector<Variable>, std::vector<Variable>, std::vector<Variable>, std::vector<Variable>, unsigned int, unsigned int, FF,
                      fptype>(),
             "name"_a, "a_r"_a, "a_i"_a, "sA0"_a, "sA"_a, "s0_prod"_a, "s0_scatt"_a, "beta_r"_a, "beta_i"_a, "f_prod_r"_a, "f_prod_i"_a, "fscat"_a,
             "poles"_a, "L"_a, "Mpair"_a, "FormFac"_a = FF::BL_Prime, "radius"_a = 1.5);
    py::class_<Lineshapes::GSpline, Lineshape>(mls, "GSpline")
        .def(py::init<std::string, Variable, Variable, unsigned int, unsigned int, FF, fptype, std::vector<Variable>, Lineshapes::spline_t>(),
             "name"_a, "mass"_a, "width"_a, "L"_a, "Mpair"_a, "FormFac"_a, "radius"_a, "AdditionalVars"_a, "SplineInfo"_a)
        .def_static("curvatures", &Lineshapes::GSpline::curvatures, "vars"_a, "SplineInfo"_a);
    py::class_<Lineshapes::FOCUS, Lineshape> focus(mls, "FOCUS");
    py::enum_<Lineshapes::FOCUS::Mod>(focus, "Mod")
        .value("Kpi", Lineshapes::FOCUS::Mod::Kpi).value("KEta", Lineshapes::FOCUS::Mod::KEta).value("I32", Lineshapes::FOCUS::Mod::I32);
    focus.def(py::init<std::string, Lineshapes::FOCUS::Mod, Variable, Variable, unsigned int, unsigned int, FF, fptype>(), "name"_a, "mod"_a,
              "mass"_a, "width"_a, "L"_a, "Mpair"_a, "FormFac"_a = FF::BL_Prime, "radius"_a = 1.5);
    py::enum_<SF_4Body>(m, "SF_4Body")
        .value("DtoPP1_PtoSP2_StoP3P4", SF_4Body::DtoPP1_PtoSP2_StoP3P4)
        .value("DtoPP1_PtoVP2_VtoP3P4", SF_4Body::DtoPP1_PtoVP2_VtoP3P4)
        .value("DtoV1V2_V1toP1P2_V2toP3P4_S", SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_S)
        .value("DtoV1V2_V1toP1P2_V2toP3P4_P", SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_P)
        .value("DtoV1V2_V1toP1P2_V2toP3P4_D", SF_4Body::DtoV1V2_V1toP1P2_V2toP3P4_D)
        .value("DtoAP1_AtoVP2_VtoP3P4", SF_4Body::DtoAP1_AtoVP2_VtoP3P4)
        .value("DtoAP1_AtoVP2Dwave_VtoP3P4", SF_4Body::DtoAP1_AtoVP2Dwave_VtoP3P4)
        .value("DtoVS_VtoP1P2_StoP3P4", SF_4Body::DtoVS_VtoP1P2_StoP3P4)
        .value("DtoV1P1_V1toV2P2_V2toP3P4", SF_4Body::DtoV1P1_V1toV2P2_V2toP3P4)
        .value("DtoAP1_AtoSP2_StoP3P4", SF_4Body::DtoAP1_AtoSP2_StoP3P4)
        .value("DtoTP1_TtoVP2_VtoP3P4", SF_4Body::DtoTP1_TtoVP2_VtoP3P4)
        .value("FF_12_34_L1", SF_4Body::FF_12_34_L1)
        .value("FF_12_34_L2", SF_4Body::FF_12_34_L2)
        .value("FF_123_4_L1", SF_4Body::FF_123_4_L1)
        .value("FF_123_4_L2", SF_4Body::FF_123_4_L2)
        .value("ONE", SF_4Body::ONE);
    py::class_<SpinFactor, AmpComponent>(m, "SpinFactor")
        .def(py::init<std::string, SF_4Body, fptype, unsigned int, unsigned int, unsigned int, unsigned int>(), "name"_a, "SF"_a,
             "mD0"_a, "P0"_a, "P1"_a, "P2"_a, "P3"_a)
        .def(py::init<std::string, SF_4Body, unsigned int, unsigned int, unsigned int, unsigned int>(), "name"_a, "SF"_a, "P0"_a,
             "P1"_a, "P2"_a, "P3"_a);
    py::class_<Amplitude, AmpComponent>(m, "Amplitude")
        .def(py::init<std::string, Variable, Variable, std::vector<Lineshape *>, std::vector<SpinFactor *>, unsigned int>(),
             "uniqueDecayStr"_a, "ar"_a, "ai"_a, "LS"_a, "SF"_a, "nPerm"_a = 1, py::keep_alive<1, 5>(), py::keep_alive<1, 6>())
        .def("lineshapeParametersChanged", &Amplitude::lineshapeParametersChanged);
    py::class_<DecayInfo4>(m, "DecayInfo4")
        .def(py::init<>())
        .def_readwrite("particle_masses", &DecayInfo4::particle_masses)
        .def_readwrite("meson_radius", &DecayInfo4::meson_radius)
        .def_readwrite("amplitudes", &DecayInfo4::amplitudes)
        .def_readwrite("amplitudes_B", &DecayInfo4::amplitudes_B)
        .def("add_amplitude", [](DecayInfo4 &d, Amplitude *a) { d.amplitudes.push_back(a); }, py::keep_alive<1, 2>())
        .def("add_amplitude_B", [](DecayInfo4 &d, Amplitude *a) { d.amplitudes_B.push_back(a); }, py::keep_alive<1, 2>());
    py::class_<Amp4BodyBase, GooPdf> a4b(m, "Amp4BodyBase");
    py::enum_<Amp4BodyBase::NormMode>(a4b, "NormMode").value("Gram", Amp4BodyBase::NormMode::Gram).value("PerEvent", Amp4BodyBase::NormMode::PerEvent);
    a4b.def("setNormMode", &Amp4BodyBase::setNormMode, "mode"_a)
        .def("getNormMode", &Amp4BodyBase::getNormMode)
        .def("setForceIntegrals", &Amp4BodyBase::setForceIntegrals, "f"_a = true)
        .def("numLineshapes", &Amp4BodyBase::numLineshapes)
        .def("numSpinFactors", &Amp4BodyBase::numSpinFactors)
        .def("numAmplitudes", &Amp4BodyBase::numAmplitudes)
        .def("setNormEvents",
             [](Amp4BodyBase &a, py::array_t<fptype, py::array::c_style | py::array::forcecast> arr) {
                 if(arr.ndim() != 2 || arr.shape(0) != 5) throw GeneralError("setNormEvents expects an array of shape (5, M)");
                 std::vector<std::vector<fptype>> cols(5);
                 for(int k = 0; k < 5; k++) cols[k].assign(arr.data(k, 0), arr.data(k, 0) + arr.shape(1));
                 a.setNormEvents(cols);
             },
             "cols5"_a)
        .def("getNormEvents",
             [](Amp4BodyBase &a) {
                 auto &c = a.getNormEvents();
                 size_t n = c.empty() ? 0 : c[0].size();
                 py::array_t<fptype> r({(size_t)5, n});
                 for(int k = 0; k < 5 && n; k++) std::copy(c[k].begin(), c[k].end(), r.mutable_data(k, 0));
                 return r;
             })
        .def("getCachedAmp", [](const Amp4BodyBase &a, size_t i) { return py::array(py::cast(a.getCachedAmp(i))); }, "a"_a);
    py::class_<Amp4Body, Amp4BodyBase>(m, "Amp4Body")
        .def(py::init<std::string, std::vector<Observable>, DecayInfo4, GooPdf *, unsigned int>(), "name"_a, "observables"_a,
             "decay"_a, "eff"_a, "MCeventsNorm"_a = 5000000, py::keep_alive<1, 4>(), py::keep_alive<1, 5>())
        .def("setDataSize", &Amp4Body::setDataSize, "dataSize"_a, "evtSize"_a = 6)
        .def("setGenerationOffset", &Amp4Body::setGenerationOffset, "off"_a)
        .def("getGenerationOffset", &Amp4Body::getGenerationOffset)
        .def(
            "GenerateSig",
            [](Amp4Body &self, size_t numEvents, int seed) { return generated_to_numpy(self.GenerateSigColumns(numEvents, seed), 4); },
            "numEvents"_a, "seed"_a = 0)
        .def("getMCevents", &Amp4Body::getMCevents);
    // time-dependent: python/PDFs/physics/{Amp4Body_TD.cu:18-38,TruthResolution.cpp,ThreeGaussResolution.cpp,DalitzPlotHelpers.cpp}
    py::class_<MixingTimeResolution, GooPdf>(m, "MixingTimeResolution").def("normalization", &MixingTimeResolution::normalization);
    py::class_<TruthResolution, MixingTimeResolution>(m, "TruthResolution").def(py::init<>());
    py::class_<ThreeGaussResolution, MixingTimeResolution>(m, "ThreeGaussResolution")
        .def(py::init<Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable,
                      Variable, Variable>(),
             "cf"_a, "tf"_a, "cb"_a, "cs"_a, "tb"_a, "ts"_a, "ob"_a, "os"_a, "sb"_a, "common_bias_offset"_a, "common_width_scale"_a,
             "common_bias_offset2"_a, "common_width_scale2"_a);
    // time-dependent Dalitz amplitude (python/PDFs/physics/Amp3Body_TD.cpp, DalitzPlotHelpers.cpp upstream)
    py::class_<DecayInfo3t, DecayInfo3>(m, "DecayInfo3t")
        .def(py::init<Variable, Variable, Variable, Variable, Variable>(), "tau"_a, "xmixing"_a, "ymixing"_a, "deltax"_a, "deltay"_a)
        .def_readonly("_tau", &DecayInfo3t::_tau)
        .def_readonly("_xmixing", &DecayInfo3t::_xmixing)
        .def_readonly("_ymixing", &DecayInfo3t::_ymixing)
        .def_readonly("_deltax", &DecayInfo3t::_deltax)
        .def_readonly("_deltay", &DecayInfo3t::_deltay);
    py::class_<Amp3Body_TD, Amp3BodyBase>(m, "Amp3Body_TD")
        .def(py::init<std::string, Observable, Observable, Observable, Observable, EventNumber, DecayInfo3t, MixingTimeResolution *, GooPdf *,
                      Observable *, Observable *>(),
             "n"_a, "dtime"_a, "sigmat"_a, "m12"_a, "m13"_a, "eventNumber"_a, "decay"_a, "r"_a, "eff"_a, "mistag"_a = nullptr,
             "charmtag"_a = nullptr, py::keep_alive<1, 8>(), py::keep_alive<1, 9>(), py::keep_alive<1, 10>())
        .def("setDataSize", &Amp3Body_TD::setDataSize, "dataSize"_a, "evtSize"_a = 5, "offset"_a = 0)
        .def("setForceIntegrals", &Amp3Body_TD::setForceIntegrals, "f"_a = true)
        .def("setD0Fraction", &Amp3Body_TD::setD0Fraction, "d0fraction"_a)
        .def("getD0Fraction", &Amp3Body_TD::getD0Fraction)
        .def("getCachedWaveA", [](const Amp3Body_TD &a, size_t i) { return py::array(py::cast(a.getCachedWaveA(i))); }, "i"_a)
        .def("getCachedWaveB", [](const Amp3Body_TD &a, size_t i) { return py::array(py::cast(a.getCachedWaveB(i))); }, "i"_a);
    py::class_<DecayInfo4t, DecayInfo4>(m, "DecayInfo4t")
        .def(py::init<Variable, Variable, Variable, Variable>(), "tau"_a, "xmixing"_a, "ymixing"_a, "SqWStoRSrate"_a)
        .def_readonly("_tau", &DecayInfo4t::_tau)
        .def_readonly("_xmixing", &DecayInfo4t::_xmixing)
        .def_readonly("_ymixing", &DecayInfo4t::_ymixing)
        .def_readonly("_SqWStoRSrate", &DecayInfo4t::_SqWStoRSrate);
    // normalisation-event batches (python/PDFs/physics/detail/NormEvents_4Body_*.cpp upstream)
    py::class_<NormEvents_4Body_Base>(m, "NormEvents_4Body_Base")
        .def("isWeighted", &NormEvents_4Body_Base::isWeighted);
    py::class_<NormEvents_4Body_DeviceCached, NormEvents_4Body_Base>(m, "NormEvents_4Body_DeviceCached")
        .def_static("buildBatches", &NormEvents_4Body_DeviceCached::buildBatches, "normSeeds"_a, "numNormEventsToGenPerBatch"_a,
                    "motherAndDaughterMasses"_a, py::return_value_policy::take_ownership);
    py::class_<NormEvents_4Body_HostCached, NormEvents_4Body_Base>(m, "NormEvents_4Body_HostCached")
        .def_static("buildBatches", &NormEvents_4Body_HostCached::buildBatches, "normSeeds"_a, "numNormEventsToGenPerBatch"_a,
                    "motherAndDaughterMasses"_a, py::return_value_policy::take_ownership);
    py::class_<NormEvents_4Body_Weighted, NormEvents_4Body_Base>(m, "NormEvents_4Body_Weighted")
        .def(py::init<std::vector<fptype>, std::vector<fptype>, std::vector<fptype>, std::vector<fptype>, std::vector<fptype>,
                      std::vector<fptype>, std::vector<fptype>, std::vector<fptype>>(),
             "m12"_a, "m34"_a, "cos12"_a, "cos34"_a, "phi"_a, "dtime"_a, "sigma"_a, "weights"_a);
    py::class_<NormEvents_4Body_WeightedDevice, NormEvents_4Body_Weighted>(m, "NormEvents_4Body_WeightedDevice")
        .def(py::init<std::vector<fptype>, std::vector<fptype>, std::vector<fptype>, std::vector<fptype>, std::vector<fptype>,
                      std::vector<fptype>, std::vector<fptype>, std::vector<fptype>>(),
             "m12"_a, "m34"_a, "cos12"_a, "cos34"_a, "phi"_a, "dtime"_a, "sigma"_a, "weights"_a);
    py::class_<Amp4Body_TD, Amp4BodyBase>(m, "Amp4Body_TD")
        .def(py::init([](std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *r, GooPdf *eff,
                         Observable *mistag, std::vector<NormEvents_4Body_Base *> normEvents) {
                 std::vector<NormEvents_4Body_Base *> copies; // the C++ constructor takes ownership; Python keeps its own objects
                 for(auto *b : normEvents) copies.push_back(new NormEvents_4Body_Base(*b));
                 return new Amp4Body_TD(n, observables, decay, r, eff, mistag, copies);
             }),
             "n"_a, "observables"_a, "decay"_a, "r"_a, "eff"_a, "mistag"_a, "normEvents"_a, py::keep_alive<1, 4>(), py::keep_alive<1, 5>(),
             py::keep_alive<1, 6>(), py::keep_alive<1, 7>())
        .def("hasWeightedNormEvents", &Amp4Body_TD::hasWeightedNormEvents)
        .def(py::init<std::string, std::vector<Observable>, DecayInfo4t, MixingTimeResolution *, GooPdf *, Observable *, long,
                      unsigned int>(),
             "n"_a, "observables"_a, "decay"_a, "r"_a, "eff"_a, "mistag"_a, "normSeed"_a, "MCeventsNorm"_a = 5000000,
             py::keep_alive<1, 4>(), py::keep_alive<1, 5>(), py::keep_alive<1, 6>(), py::keep_alive<1, 7>())
        .def("setDataSize", &Amp4Body_TD::setDataSize, "dataSize"_a, "evtSize"_a = 8)
        .def("getNumAccNormEvents", &Amp4Body_TD::getNumAccNormEvents)
        .def("getSumInitNormEventWeights", &Amp4Body_TD::getSumInitNormEventWeights)
        .def("setGenerationOffset", &Amp4Body_TD::setGenerationOffset, "off"_a)
        .def("getGenerationOffset", &Amp4Body_TD::getGenerationOffset)
        .def("setMaxWeight", &Amp4Body_TD::setMaxWeight, "wmax"_a)
        .def("getMaxWeight", &Amp4Body_TD::getMaxWeight)
        .def(
            "GenerateSig",
            [](Amp4Body_TD &self, size_t numEvents, int seed) { return generated_to_numpy(self.GenerateSigColumns(numEvents, seed), 4); },
            "numEvents"_a, "seed"_a = 0,
            "(particles[4, nAcc, 3 or 4], variables[6, nAcc] = m12, m34, cos12, cos34, phi, dtime, weights[nAcc] = PDF at t = 0, flags[nAcc])");
    m.def(
        "generatePhaseSpaceParticles",
        [](std::vector<fptype> masses, size_t n, unsigned long seed, size_t offset) {
            GeneratedSignal G;
            generatePhaseSpaceParticles(masses, n, seed, offset, G.particles, G.weights);
            G.flags.assign(n, false);
            py::tuple t = generated_to_numpy(G, 4);
            return py::make_tuple(t[0], t[2]);
        },
        "mother_and_daughter_masses"_a, "nEvents"_a, "seed"_a = 0, "offset"_a = 0,
        "GENBOD phase-space events for 3 or 4 daughters: (particles[nd, N, 4] as (E, px, py, pz), weights[N])");
    m.def(
        "devicePhaseSpace",
        [](std::vector<fptype> masses, size_t n, unsigned long seed, size_t offset) {
            GeneratedSignal G;
            std::vector<std::vector<fptype>> v5;
            devicePhaseSpace(masses, n, seed, offset, &G.particles, G.weights, masses.size() == 5 ? &v5 : nullptr);
            G.flags.assign(n, false);
            py::tuple t = generated_to_numpy(G, 4);
            py::array_t<fptype> r({(size_t)(v5.empty() ? 0 : 5), n});
            for(size_t k = 0; k < v5.size(); k++) std::copy(v5[k].begin(), v5[k].end(), r.mutable_data(k, 0));
            return py::make_tuple(t[0], t[2], r);
        },
        "mother_and_daughter_masses"_a, "nEvents"_a, "seed"_a = 0, "offset"_a = 0,
        "the stream of generatePhaseSpaceParticles generated on the GPU: (particles[nd, N, 4], weights[N], dim5[5, N] for four daughters)");
    m.def(
        "devicePhaseSpace5",
        [](std::vector<fptype> masses, size_t n, unsigned long seed, size_t offset) {
            std::vector<fptype> w;
            std::vector<std::vector<fptype>> v5;
            devicePhaseSpace(masses, n, seed, offset, nullptr, w, &v5);
            py::array_t<fptype> r({(size_t)5, n}), wa(n);
            for(int k = 0; k < 5; k++) std::copy(v5[k].begin(), v5[k].end(), r.mutable_data(k, 0));
            std::copy(w.begin(), w.end(), wa.mutable_data());
            return py::make_tuple(r, wa);
        },
        "particle_masses"_a, "n"_a, "seed"_a = 0, "offset"_a = 0,
        "generatePhaseSpace5's stream generated on the GPU: (m12, m34, cos12, cos34, phi) columns + weights");
    m.def(
        "counterUniform",
        [](unsigned long seed, py::array_t<long long, py::array::c_style | py::array::forcecast> events, unsigned long stream) {
            py::array_t<fptype> out(events.size());
            for(py::ssize_t i = 0; i < events.size(); i++) out.mutable_data()[i] = detail::counterUniform(seed, (size_t)events.data()[i], stream);
            return out;
        },
        "seed"_a, "events"_a, "stream"_a, "uniform numbers in (0, 1) of the counter-based generator (stream `stream` of each event index)");
    m.def("generatePhaseSpace5",
          [](std::vector<fptype> masses, size_t n, unsigned long seed) {
              std::vector<fptype> w;
              auto c = generatePhaseSpace5(masses, n, seed, &w);
              py::array_t<fptype> r({(size_t)5, n}), wa(n);
              for(int k = 0; k < 5; k++) std::copy(c[k].begin(), c[k].end(), r.mutable_data(k, 0));
              std::copy(w.begin(), w.end(), wa.mutable_data());
              return py::make_tuple(r, wa);
          },
          "particle_masses"_a, "n"_a, "seed"_a = 0, "four-body phase-space sample as (m12, m34, cos12, cos34, phi) columns + weights");
    m.def("fourVectorsToDim5",
          [](py::array_t<fptype, py::array::c_style | py::array::forcecast> p) {
              if(p.size() != 16) throw GeneralError("fourVectorsToDim5 expects 16 numbers: 4 x (E, px, py, pz)");
              py::array_t<fptype> out(5);
              fourVectorsToDim5(p.data(), out.mutable_data());
              return out;
          });
}
