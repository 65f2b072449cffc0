// module.cpp — pybind11 module `_goofit` exposing the host API with the reference's Python names
// (reference: python/goofit/{Variable,DataSet,UnbinnedDataSet,BinnedDataSet,PdfBase,FitManager,FitControl}.cpp,
//  python/PDFs/GooPdf.cpp and one file per PDF under python/PDFs/). Only the classes on the north-star
// path are bound.
#include <pybind11/complex.h>
#include <pybind11/functional.h>
#include <pybind11/iostream.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <Minuit2/FunctionMinimum.h>
#include <Minuit2/MnUserParameterState.h>
#include <goofit/Application.h>
#include <goofit/detail/Fit.h>
#include <goofit/detail/Physics3.h>

namespace py = pybind11;
using namespace pybind11::literals;
using namespace GooFit;

void init_physics4(py::module &m);
void init_minuit2(py::module &m);
py::tuple generated_to_numpy(const GooFit::GeneratedSignal &G, int ncomp);


template <class DS>
static DS *make_dataset(py::args args, py::kwargs kwargs) {
    std::string name;
    std::vector<Observable> vars;
    for(auto arg : args) vars.push_back(arg.cast<Observable>());
    if(kwargs.contains("name")) name = kwargs["name"].cast<std::string>();
    return new DS(vars, name);
}

PYBIND11_MODULE(_goofit, m) {
    m.doc() = "goofit_b200: B200-native GooFit likelihood engine (host API over hand-written sm_100a kernels)";
    py::register_exception<GeneralError>(m, "GeneralError");

    // module-level functions and constants of python/goofit/Application.cpp and python/goofit/Version.cpp of the reference: every
    // example there opens with `from goofit import *; print_goofit_info()`
    m.def("print_splash", &print_splash, "Print a splash screen", py::call_guard<py::scoped_ostream_redirect>());
    m.def(
        "goofit_info", [](int gpuDev) { return goofit_info_version() + "\n" + goofit_info_device(gpuDev); }, "Get GooFit information", "gpuDevice"_a = 0);
    m.def(
        "print_goofit_info", [](int gpuDev) { py::print(goofit_info_version() + "\n" + goofit_info_device(gpuDev)); },
        "Print GooFit information (same as goofit_info, kept for legacy code)", "gpuDevice"_a = 0);
    m.def("set_floating_exceptions", &Application::set_floating_exceptions, "Enable floating point exceptions as Exceptions (no effect on a CUDA build, as upstream)");
    m.def("__main__", []() { py::print(goofit_info_version() + "\n" + goofit_info_device(0)); });
    m.attr("__version__") = py::make_tuple(GOOFIT_VERSION_MAJOR, GOOFIT_VERSION_MINOR, GOOFIT_VERSION_PATCH);
    m.attr("GOOFIT_VERSION_TUPLE") = py::make_tuple(GOOFIT_VERSION_MAJOR, GOOFIT_VERSION_MINOR, GOOFIT_VERSION_PATCH);
    m.attr("GOOFIT_VERSION") = GOOFIT_VERSION;
    m.attr("GOOFIT_GIT_VERSION") = GOOFIT_VERSION;
    m.attr("GOOFIT_TAG") = "";
    m.attr("GOOFIT_SOURCE_DIR") = GOOFIT_SOURCE_DIR;
    m.attr("CMAKE_CXX_FLAGS") = CMAKE_CXX_FLAGS;
    m.attr("CMAKE_BUILD_TYPE") = CMAKE_BUILD_TYPE;
    m.attr("GOOFIT_DEVICE") = GOOFIT_DEVICE;
    m.attr("GOOFIT_HOST") = GOOFIT_HOST;
    m.attr("ARCH_FLAGS") = GOOFIT_ARCH_FLAGS;
    m.attr("ROOT_FOUND") = GOOFIT_ROOT_FOUND;
    m.attr("INT_MAX") = INT_MAX;

    m.def("set_device", [](int dev) { detail::Engine::get().setDevice(dev); }, "device"_a);
    m.def("set_shard", [](int rank, int size) { detail::Engine::get().setShard(rank, size); }, "rank"_a, "size"_a);
    m.def("attach_comm_from_env", [](int rank, int size) { detail::Engine::get().attachCommFromEnv(rank, size); }, "rank"_a, "size"_a,
          "file rendezvous of the NCCL id (keyed on MASTER_PORT), gf_comm_init and set_shard in one call");
    m.def("comm_attached", [] { return gf_comm_attached(detail::Engine::get().ctx()); });
    m.def("ctx_handle", [] { return reinterpret_cast<uintptr_t>(detail::Engine::get().ctx()); },
          "raw gf_ctx* of the engine (for attaching an NCCL communicator through the C ABI)");
    m.def("launch_count", [] { return gf_launch_count(detail::Engine::get().ctx()); });
    m.def("last_kernel_ms", [] { return gf_last_kernel_ms(detail::Engine::get().ctx()); });
    m.def("fast_enable", [](bool on) { return gf_set_option(detail::Engine::get().ctx(), "fast", on ? 1 : 0) != 0; }, "on"_a,
          "shape-specialised likelihood kernels on/off (off = interpreter for every program); returns the previous setting");
    m.def("fused_enable", [](bool on) { return gf_set_option(detail::Engine::get().ctx(), "fused", on ? 1 : 0) != 0; }, "on"_a,
          "one-pass lineshape -> amplitude -> Gram update on/off (off = separate kernels); returns the previous setting");
    m.def("last_kernel_name", [] { return std::string(gf_last_kernel_name(detail::Engine::get().ctx())); },
          "which copy of the likelihood kernel the last NLL evaluation launched");
    m.def("program_params", [] { return py::array(py::cast(detail::Engine::get().prog.params)); },
          "copy of the flattened parameter block as last pushed by updateParameters()");
    m.def("program_norms", [] { return py::array(py::cast(detail::Engine::get().prog.norms)); });
    m.def("local_events", [] { return detail::Engine::get().localEvents; });
    m.def("shard_range",
          [](size_t n, int rank, int size) { // the split PdfBase::setData applies (reference PdfBase.cu:364-378)
              size_t per = n / size, first = per * rank, count = rank == size - 1 ? n - first : per;
              return py::make_tuple(first, count);
          },
          "n"_a, "rank"_a, "size"_a);
    m.def("shutdown", [] { detail::Engine::get().shutdown(); });

    // ---- variables ----
    py::class_<Indexable>(m, "Indexable")
        .def_property_readonly("name", &Indexable::getName)
        .def_property("value", &Indexable::getValue, &Indexable::setValue)
        .def_property("index", &Indexable::getIndex, &Indexable::setIndex)
        .def_property("upperlimit", &Indexable::getUpperLimit, &Indexable::setUpperLimit)
        .def_property("lowerlimit", &Indexable::getLowerLimit, &Indexable::setLowerLimit)
        .def("getName", &Indexable::getName)
        .def("getValue", &Indexable::getValue)
        .def("setValue", &Indexable::setValue)
        .def("getIndex", &Indexable::getIndex)
        .def("getUpperLimit", &Indexable::getUpperLimit)
        .def("getLowerLimit", &Indexable::getLowerLimit)
        .def("__repr__", [](const Indexable &v) { return "<Indexable: " + v.getName() + ">"; })
        .def("__bool__", &Indexable::inRange);
    py::class_<Observable, Indexable>(m, "Observable")
        .def(py::init<std::string, fptype, fptype>(), "name"_a, "min"_a, "max"_a)
        .def_property("numbins", &Observable::getNumBins, &Observable::setNumBins)
        .def("getNumBins", &Observable::getNumBins)
        .def("setNumBins", &Observable::setNumBins)
        .def("getBinSize", &Observable::getBinSize)
        .def_property_readonly("bin_size", &Observable::getBinSize)
        .def_property_readonly("bin_centers",
                               [](const Observable &v) {
                                   py::array_t<fptype> r(v.getNumBins());
                                   for(size_t i = 0; i < v.getNumBins(); i++)
                                       r.mutable_at(i) = v.getLowerLimit() + (i + 0.5) * v.getBinSize();
                                   return r;
                               })
        .def("__repr__", [](const Observable &v) { return "<Observable: " + v.getName() + ">"; })
        .def("__str__", [](const Observable &v) {
            std::ostringstream o;
            o << v;
            return o.str();
        });
    py::class_<Variable, Indexable>(m, "Variable")
        .def(py::init<std::string, fptype>(), "A constant variable", "name"_a, "value"_a)
        .def(py::init<std::string, fptype, fptype>(), "name"_a, "value"_a, "error"_a)
        .def(py::init<std::string, fptype, fptype, fptype>(), "name"_a, "value"_a, "min"_a, "max"_a)
        .def(py::init<std::string, fptype, fptype, fptype, fptype, bool>(), "name"_a, "value"_a, "error"_a, "min"_a,
             "max"_a, "fixed"_a = false)
        .def_property("error", &Variable::getError, &Variable::setError)
        .def_property("fixed", &Variable::IsFixed, &Variable::setFixed)
        .def_property("fitter_index", &Variable::getFitterIndex, &Variable::setFitterIndex)
        .def_property_readonly("fitterIndex", &Variable::getFitterIndex)
        .def_property("blind", nullptr, &Variable::setBlind)
        .def_property("changed", &Variable::getChanged, &Variable::setChanged)
        .def("getError", &Variable::getError)
        .def("setFixed", &Variable::setFixed)
        .def("setBlind", &Variable::setBlind)
        .def("__repr__", [](const Variable &v) { return "<Variable: " + v.getName() + ">"; })
        .def("__str__", [](const Variable &v) {
            std::ostringstream o;
            o << v;
            return o.str();
        });
    py::class_<EventNumber, Observable>(m, "EventNumber")
        .def(py::init<std::string, fptype, fptype>(), "name"_a, "min"_a = 0, "max"_a = EventNumber::maxint)
        .def_property_readonly_static("maxint", [](py::object) { return EventNumber::maxint; })
        .def("__repr__", [](const EventNumber &v) { return "<EventNumber: " + v.getName() + ">"; });

    // ---- datasets ----
    py::class_<DataSet>(m, "DataSet")
        .def("addEvent", (void (DataSet::*)()) & DataSet::addEvent)
        .def("addWeightedEvent", &DataSet::addWeightedEvent)
        .def("getNumEvents", &DataSet::getNumEvents)
        .def("__len__", &DataSet::getNumEvents)
        .def_property_readonly("name", &DataSet::getName)
        .def_property_readonly("variables", &DataSet::getObservables)
        .def("getObservables", &DataSet::getObservables);
    auto to_numpy = [](const UnbinnedDataSet &d) {
        size_t rows = d.getObservables().size(), cols = d.getNumEvents();
        py::array_t<fptype> r({rows, cols});
        for(size_t i = 0; i < rows; i++) std::copy(d.column(i).begin(), d.column(i).end(), r.mutable_data(i, 0));
        return r;
    };
    auto from_numpy = [](UnbinnedDataSet &d, py::array_t<fptype, py::array::c_style | py::array::forcecast> a, bool filter) {
        if(a.ndim() != 2) throw GeneralError("Expected a 2-D array (variables x events)");
        size_t rows = a.shape(0), cols = a.shape(1), nobs = d.getObservables().size();
        if(rows != nobs && rows + 1 != nobs)
            throw GeneralError("The wrong number of rows, expected {}, but matrix had {}", nobs, rows);
        if(cols == 0) return; // no events: nothing to append (and no element to take the address of)
        if(!filter) { // bulk path: columns are appended without touching the Observables
            std::vector<const fptype *> cp;
            std::vector<fptype> idx;
            for(size_t i = 0; i < rows; i++) cp.push_back(a.data(i, 0));
            if(rows + 1 == nobs) {
                idx.resize(cols);
                for(size_t j = 0; j < cols; j++) idx[j] = (fptype)(d.getNumEvents() + j);
                cp.push_back(idx.data());
            }
            for(size_t i = 0; i < nobs; i++) { // same range check as addEvent()
                const Observable &o = d.getObservables()[i];
                if(o.getLowerLimit() == o.getUpperLimit()) continue;
                for(size_t j = 0; j < cols; j++)
                    if(cp[i][j] > o.getUpperLimit() || cp[i][j] < o.getLowerLimit())
                        throw OutOfRange(o.getName(), cp[i][j], o.getLowerLimit(), o.getUpperLimit());
            }
            d.appendColumns(cp, cols);
            return;
        }
        struct View {
            const py::array_t<fptype, py::array::c_style | py::array::forcecast> &a;
            size_t rows() const { return a.shape(0); }
            size_t cols() const { return a.shape(1); }
            fptype operator()(size_t i, size_t j) const { return a.at(i, j); }
        } v{a};
        d.from_matrix(v, true);
    };
    py::class_<UnbinnedDataSet, DataSet>(m, "UnbinnedDataSet")
        .def(py::init(&make_dataset<UnbinnedDataSet>),
             "Takes an unlimited number of Observerables, use 'name=' to set a name if desired")
        .def("__getitem__", [to_numpy](const UnbinnedDataSet &d, py::object idx) { return py::object(to_numpy(d)).attr("__getitem__")(idx); })
        .def("from_numpy", from_numpy, "array"_a, "filter"_a = false)
        .def("from_matrix", from_numpy, "matrix"_a, "filter"_a = false)
        .def("to_numpy", to_numpy)
        .def("to_matrix", to_numpy)
        .def("loadEvent", &UnbinnedDataSet::loadEvent, "event_number"_a)
        .def("clear", &UnbinnedDataSet::clear)
        .def("fillWithGrid", &UnbinnedDataSet::fillWithGrid)
        .def("setValueForAllEvents", &UnbinnedDataSet::setValueForAllEvents, "var"_a)
        .def("getValue", &UnbinnedDataSet::getValue, "var"_a, "idx"_a);
    py::class_<BinnedDataSet, DataSet>(m, "BinnedDataSet")
        .def(py::init(&make_dataset<BinnedDataSet>))
        .def("getBinCenter", (fptype(BinnedDataSet::*)(size_t, size_t) const) & BinnedDataSet::getBinCenter)
        .def("getBinContent", &BinnedDataSet::getBinContent, "bin"_a)
        .def("setBinContent", &BinnedDataSet::setBinContent, "bin"_a, "value"_a)
        .def("getBinError", &BinnedDataSet::getBinError, "bin"_a)
        .def("setBinError", &BinnedDataSet::setBinError, "bin"_a, "value"_a)
        .def("getBinSize", &BinnedDataSet::getBinSize, "ivar"_a)
        .def("getBinVolume", &BinnedDataSet::getBinVolume, "bin"_a)
        .def("getNumBins", &BinnedDataSet::getNumBins)
        .def("getNumWeightedEvents", &BinnedDataSet::getNumWeightedEvents)
        // python/goofit/BinnedDataSet.cpp:25-34
        .def_property_readonly("bin_center", [](const BinnedDataSet &d) {
            std::vector<std::vector<fptype>> c(d.numVariables());
            for(size_t v = 0; v < c.size(); v++)
                for(size_t b = 0; b < d.getNumBins(); b++) c[v].push_back(d.getBinCenter(v, b));
            return c;
        })
        .def_property_readonly("bin_number", &BinnedDataSet::getBinNumber)
        .def_property_readonly("bin_volume", [](const BinnedDataSet &d) { std::vector<fptype> v; for(size_t b = 0; b < d.getNumBins(); b++) v.push_back(d.getBinVolume(b)); return v; })
        .def_property_readonly("bin_error", [](const BinnedDataSet &d) { std::vector<fptype> v; for(size_t b = 0; b < d.getNumBins(); b++) v.push_back(d.getBinError(b)); return v; })
        .def_property_readonly("num_bins", &BinnedDataSet::getNumBins)
        .def_property_readonly("num_events", &BinnedDataSet::getNumEvents)
        .def_property_readonly("diminsions", &BinnedDataSet::getDiminsions);

    // ---- fit control ----
    py::enum_<EvalFunc>(m, "EvalFunc")
        .value("Eval", EvalFunc::Eval).value("NLL", EvalFunc::NLL).value("Prob", EvalFunc::Prob)
        .value("BinAvg", EvalFunc::BinAvg).value("BinWithError", EvalFunc::BinWithError).value("Chisq", EvalFunc::Chisq);
    py::class_<FitControl, std::shared_ptr<FitControl>>(m, "FitControl")
        .def("binnedFit", &FitControl::binnedFit)
        .def("binErrors", &FitControl::binErrors)
        .def("metricIsPdf", &FitControl::metricIsPdf)
        .def("getMetric", &FitControl::getMetric)
        .def("getName", &FitControl::getName);
    py::class_<UnbinnedNllFit, FitControl, std::shared_ptr<UnbinnedNllFit>>(m, "UnbinnedNllFit").def(py::init<>());
    py::class_<BinnedNllFit, FitControl, std::shared_ptr<BinnedNllFit>>(m, "BinnedNllFit").def(py::init<>());
    py::class_<BinnedErrorFit, FitControl, std::shared_ptr<BinnedErrorFit>>(m, "BinnedErrorFit").def(py::init<>());
    py::class_<BinnedChisqFit, FitControl, std::shared_ptr<BinnedChisqFit>>(m, "BinnedChisqFit").def(py::init<>());
    py::class_<ProbFit, FitControl, std::shared_ptr<ProbFit>>(m, "ProbFit").def(py::init<>());
    py::class_<EvalFit, FitControl, std::shared_ptr<EvalFit>>(m, "EvalFit").def(py::init<>());

    // ---- PDFs ----
    py::class_<PdfBase>(m, "PdfBase")
        .def("setData", &PdfBase::setData, "ptr"_a = nullptr, py::keep_alive<1, 2>())
        .def("getData", &PdfBase::getData, py::return_value_policy::reference)
        .def("getName", &PdfBase::getName)
        .def("getPdfName", &PdfBase::getPdfName)
        .def("getParameters", &PdfBase::getParameters)
        .def("getParameterByName", &PdfBase::getParameterByName, py::return_value_policy::reference_internal)
        .def("getObservables", &PdfBase::getObservables)
        .def("getFunctionIndex", &PdfBase::getFunctionIndex)
        .def("getParameterIndex", &PdfBase::getParameterIndex)
        .def("getNormalizationIndex", &PdfBase::getNormalizationIndex)
        .def("getNormalization", &PdfBase::getNormalization)
        .def("parametersChanged", &PdfBase::parametersChanged)
        .def("updateVariable", &PdfBase::updateVariable)
        .def("updateParameters", &PdfBase::updateParameters)
        .def("setIntegrationFineness", &PdfBase::setIntegrationFineness)
        .def("setCommonNorm", &PdfBase::setCommonNorm, "v"_a = true)
        .def("__str__", [](const PdfBase &p) {
            std::ostringstream o;
            o << p;
            return o.str();
        })
        .def("fillMCDataSimple", &PdfBase::fillMCDataSimple, "events"_a, "seed"_a = 0)
        .def("fitTo", &PdfBase::fitTo, "data"_a, "verbosity"_a = 3, py::call_guard<py::scoped_ostream_redirect>())
        .def("fit", &PdfBase::fit, "verbosity"_a = 3, py::call_guard<py::scoped_ostream_redirect>());
    py::class_<GooPdf, PdfBase>(m, "GooPdf")
        .def("get_bin_grid_size", &GooPdf::get_bin_grid_size)
        .def("get_bin_grid_volume", &GooPdf::get_bin_grid_volume)
        .def("reduce_with_bins", &GooPdf::reduce_with_bins)
        .def("reduce_with_metric", &GooPdf::reduce_with_metric)
        .def("evaluate_with_metric", [](GooPdf &p) { return py::array(py::cast(p.evaluate_with_metric())); })
        .def("normalize", &GooPdf::normalize)
        .def("normalise", &GooPdf::normalize)
        .def("integrate", &GooPdf::integrate, "lo"_a, "hi"_a)
        .def("copyParams", &GooPdf::copyParams)
        .def("makeGrid", &GooPdf::makeGrid)
        .def("evaluateAtPoints", &GooPdf::evaluateAtPoints)
        .def("evaluatePdf", &GooPdf::evaluateAtPoints)
        .def("getValue", &GooPdf::getValue, "evalfunc"_a = EvalFunc::Eval)
        .def("setParameterConstantness", &GooPdf::setParameterConstantness, "constant"_a = true)
        .def("hasAnalyticIntegral", &GooPdf::hasAnalyticIntegral)
        .def("calculateNLL", &GooPdf::calculateNLL)
        .def("scan",
             [](GooPdf &p, Observable v) {
                 std::vector<fptype> out;
                 p.scan(v, out);
                 return out;
             })
        .def("setFitControl", &GooPdf::setFitControl, "fit_control"_a)
        .def("getCompProbsAtDataPoints", &GooPdf::getCompProbsAtDataPoints);

    py::class_<ExpPdf, GooPdf>(m, "ExpPdf")
        .def(py::init<std::string, Observable, Variable>(), "name"_a, "x"_a, "alpha"_a, py::keep_alive<1, 3>(), py::keep_alive<1, 4>())
        .def(py::init<std::string, Observable, Variable, Variable>(), "name"_a, "x"_a, "alpha"_a, "offset"_a)
        .def(py::init([](std::string n, Observable x, std::vector<Variable> w) { return new ExpPdf(n, x, w); }), "name"_a, "x"_a, "weights"_a)
        .def(py::init([](std::string n, Observable x, std::vector<Variable> w, Variable off) { return new ExpPdf(n, x, w, off); }), "name"_a, "x"_a, "weights"_a, "offset"_a);
    py::class_<GaussianPdf, GooPdf>(m, "GaussianPdf")
        .def(py::init<std::string, Observable, Variable, Variable>(), "name"_a, "x"_a, "mean"_a, "sigma"_a);
    py::class_<PolynomialPdf, GooPdf>(m, "PolynomialPdf")
        .def(py::init<std::string, Observable, std::vector<Variable>, unsigned int>(), "name"_a, "x"_a, "weights"_a, "lowestDegree"_a = 0)
        .def(py::init<std::string, Observable, std::vector<Variable>, Variable, unsigned int>(), "name"_a, "x"_a, "weights"_a, "x0"_a, "lowestDegree"_a = 0)
        .def(py::init<std::string, std::vector<Observable>, std::vector<Variable>, std::vector<Variable>, unsigned int>(),
             "name"_a, "obses"_a, "coeffs"_a, "offsets"_a, "maxDegree"_a = 0)
        .def("getCoefficient", &PolynomialPdf::getCoefficient);
    py::class_<ArgusPdf, GooPdf>(m, "ArgusPdf")
        .def(py::init<std::string, Observable, Variable, Variable, bool>(), "name"_a, "x"_a, "m"_a, "s"_a, "upper"_a)
        .def(py::init<std::string, Observable, Variable, Variable, bool, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "upper"_a, "power"_a);
    py::class_<KinLimitBWPdf, GooPdf>(m, "KinLimitBWPdf")
        .def(py::init<std::string, Observable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a)
        .def("setMasses", &KinLimitBWPdf::setMasses, "bigM"_a, "smallM"_a);
    py::class_<ScaledGaussianPdf, GooPdf>(m, "ScaledGaussianPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "d"_a, "e"_a);
    py::class_<JohnsonSUPdf, GooPdf>(m, "JohnsonSUPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "g"_a, "d"_a);
    py::class_<BWPdf, GooPdf>(m, "BWPdf").def(py::init<std::string, Observable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a);
    py::class_<BifurGaussPdf, GooPdf>(m, "BifurGaussPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "sL"_a, "sR"_a);
    py::class_<CrystalBallPdf, GooPdf>(m, "CrystalBallPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "a"_a)
        .def(py::init<std::string, Observable, Variable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "a"_a, "power"_a);
    py::class_<ExpGausPdf, GooPdf>(m, "ExpGausPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "t"_a);
    py::class_<NovosibirskPdf, GooPdf>(m, "NovosibirskPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "t"_a);
    py::class_<StepPdf, GooPdf>(m, "StepPdf").def(py::init<std::string, Observable, Variable, int>(), "name"_a, "x"_a, "x0"_a, "Up"_a = 1);
    py::class_<TrigThresholdPdf, GooPdf>(m, "TrigThresholdPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable, bool>(), "name"_a, "x"_a, "thresh"_a, "trigConst"_a, "linConst"_a,
             "upper"_a = true)
        .def(py::init<std::string, Observable, Observable, Variable, Variable, Variable, Variable, bool>(), "name"_a, "x"_a, "y"_a, "thresh"_a,
             "trigConst"_a, "linConst"_a, "massConstant"_a, "upper"_a);
    py::class_<BernsteinPdf, GooPdf>(m, "BernsteinPdf")
        .def(py::init<std::string, Observable, std::vector<Variable>, unsigned int>(), "name"_a, "x"_a, "coeffs"_a, "mindeg"_a = 0);
    py::class_<CorrGaussianPdf, GooPdf>(m, "CorrGaussianPdf")
        .def(py::init<std::string, Observable, Observable, Variable, Variable, Variable, Variable, Variable>(), "name"_a, "x"_a, "y"_a, "mean1"_a,
             "sigma1"_a, "mean2"_a, "sigma2"_a, "correlation"_a);
    py::class_<BinTransformPdf, GooPdf>(m, "BinTransformPdf")
        .def(py::init<std::string, std::vector<Observable>, std::vector<fptype>, std::vector<fptype>, std::vector<int>>(), "name"_a, "obses"_a,
             "limits"_a, "binSizes"_a, "numBins"_a);
    py::class_<SmoothHistogramPdf, GooPdf>(m, "SmoothHistogramPdf")
        .def(py::init<std::string, BinnedDataSet *, Variable>(), "name"_a, "hist"_a, "smoothing"_a, py::keep_alive<1, 3>());
    py::class_<VoigtianPdf, GooPdf>(m, "VoigtianPdf")
        .def(py::init<std::string, Observable, Variable, Variable, Variable>(), "name"_a, "x"_a, "m"_a, "s"_a, "w"_a);
    py::class_<VariableBinTransform1DPdf, GooPdf>(m, "VariableBinTransform1DPdf")
        .def(py::init<std::string, Observable, std::vector<fptype>>(), "name"_a, "x"_a, "binlimits"_a);
    py::class_<LandauPdf, GooPdf>(m, "LandauPdf").def(py::init<std::string, Observable, Variable, Variable>(), "name"_a, "x"_a, "mpv"_a, "sigma"_a);
    py::class_<CombinePdf, GooPdf>(m, "CombinePdf");
    py::class_<AddPdf, CombinePdf>(m, "AddPdf")
        .def(py::init<std::string, std::vector<Variable>, std::vector<PdfBase *>>(), "name"_a, "weights"_a, "comps"_a, py::keep_alive<1, 4>())
        .def(py::init<std::string, Variable, PdfBase *, PdfBase *>(), "name"_a, "frac1"_a, "func1"_a, "func2"_a, py::keep_alive<1, 4>(), py::keep_alive<1, 5>());
    py::class_<ProdPdf, CombinePdf>(m, "ProdPdf")
        .def(py::init<std::string, std::vector<PdfBase *>>(), "name"_a, "comps"_a, py::keep_alive<1, 3>());
    py::class_<EventWeightedAddPdf, CombinePdf>(m, "EventWeightedAddPdf")
        .def(py::init<std::string, std::vector<Observable>, std::vector<PdfBase *>>(), "name"_a, "weights"_a, "comps"_a, py::keep_alive<1, 4>());
    py::class_<MappedPdf, CombinePdf>(m, "MappedPdf")
        .def(py::init([](std::string n, GooPdf *map, std::vector<GooPdf *> t) { return new MappedPdf(n, map, t); }), "name"_a, "m"_a, "t"_a,
             py::keep_alive<1, 3>(), py::keep_alive<1, 4>());
    py::class_<CompositePdf, CombinePdf>(m, "CompositePdf")
        .def(py::init<std::string, PdfBase *, PdfBase *>(), "name"_a, "core"_a, "shell"_a, py::keep_alive<1, 3>(), py::keep_alive<1, 4>());
    py::class_<ConvolutionPdf, CombinePdf>(m, "ConvolutionPdf")
        .def(py::init<std::string, Observable, GooPdf *, GooPdf *>(), "name"_a, "x"_a, "model"_a, "resolution"_a, py::keep_alive<1, 4>(), py::keep_alive<1, 5>())
        .def(py::init<std::string, Observable, GooPdf *, GooPdf *, unsigned int>(), "name"_a, "x"_a, "model"_a, "resolution"_a, "numOthers"_a,
             py::keep_alive<1, 4>(), py::keep_alive<1, 5>())
        .def("registerOthers", &ConvolutionPdf::registerOthers, "others"_a)
        .def("setIntegrationConstants", &ConvolutionPdf::setIntegrationConstants, "lo"_a, "hi"_a, "step"_a);

    // ---- three-body amplitude ----
    py::enum_<DaughterPair>(m, "DaughterPair")
        .value("PAIR_12", PAIR_12).value("PAIR_13", PAIR_13).value("PAIR_23", PAIR_23).export_values();
    m.def("inDalitz", &inDalitz, "m12"_a, "m13"_a, "bigM"_a, "dm1"_a, "dm2"_a, "dm3"_a);
    py::class_<VetoInfo>(m, "VetoInfo")
        .def(py::init<Variable, Variable, DaughterPair>(), "minimum"_a, "maximum"_a, "cyclic_index"_a)
        .def_readwrite("minimum", &VetoInfo::minimum)
        .def_readwrite("maximum", &VetoInfo::maximum)
        .def_readwrite("cyclic_index", &VetoInfo::cyclic_index);
    py::class_<DalitzVetoPdf, GooPdf>(m, "DalitzVetoPdf")
        .def(py::init<std::string, Observable, Observable, Variable, Variable, Variable, Variable, std::vector<VetoInfo>>(), "name"_a, "x"_a, "y"_a,
             "motherM"_a, "d1m"_a, "d2m"_a, "d3m"_a, "vetos"_a);
    py::class_<SquareDalitzEffPdf, GooPdf>(m, "SquareDalitzEffPdf")
        .def(py::init<std::string, Observable, Observable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable>(), "name"_a,
             "m12"_a, "m13"_a, "c0"_a, "c1"_a, "c2"_a, "c3"_a, "c4"_a, "c5"_a, "c6"_a, "c7"_a);
    py::class_<AmpComponent, GooPdf>(m, "AmpComponent");
    py::class_<ResonancePdf, AmpComponent>(m, "ResonancePdf")
        .def("get_amp_real", &ResonancePdf::get_amp_real)
        .def("get_amp_img", &ResonancePdf::get_amp_img);
    auto mres = m.def_submodule("Resonances");
    py::class_<Resonances::RBW, ResonancePdf>(mres, "RBW")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, unsigned int, unsigned int, bool, bool>(),
             "name"_a, "ar"_a, "ai"_a, "mass"_a, "width"_a, "sp"_a, "cyc"_a, "norm"_a = true, "sym"_a = false);
    py::class_<Resonances::GS, ResonancePdf>(mres, "GS")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, unsigned int, unsigned int, bool, bool>(),
             "name"_a, "ar"_a, "ai"_a, "mass"_a, "width"_a, "sp"_a, "cyc"_a, "norm"_a = true, "sym"_a = false);
    py::class_<Resonances::NonRes, ResonancePdf>(mres, "NonRes")
        .def(py::init<std::string, Variable, Variable>(), "name"_a, "ar"_a, "ai"_a);
    py::class_<Resonances::Gauss, ResonancePdf>(mres, "Gauss")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, unsigned int>(), "name"_a, "ar"_a, "ai"_a, "mean"_a, "sigma"_a, "cyc"_a);
    py::class_<Resonances::LASS, ResonancePdf>(mres, "LASS")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable,
                      unsigned int, unsigned int, bool>(),
             "name"_a, "ar"_a, "ai"_a, "mass"_a, "width"_a, "_a"_a, "_r"_a, "_R"_a, "_phiR"_a, "_B"_a, "_phiB"_a, "sp"_a, "cyc"_a,
             "norm"_a = true);
    py::class_<Resonances::FLATTE, ResonancePdf>(mres, "FLATTE")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, Variable, unsigned int, bool>(), "name"_a, "ar"_a, "ai"_a,
             "mean"_a, "g1"_a, "rg2og1"_a, "cyc"_a, "symmDP"_a);
    py::class_<Resonances::kMatrix, ResonancePdf>(mres, "kMatrix")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, Variable, Variable, std::vector<Variable>, std::vector<Variable>,
                      std::vector<Variable>, std::vector<Variable>, std::vector<Variable>, std::vector<Variable>, unsigned int, unsigned int>(),
             "name"_a, "a_r"_a, "a_i"_a, "sA0"_a, "sA"_a, "s0_prod"_a, "s0_scatt"_a, "beta_r"_a, "beta_i"_a, "f_prod_r"_a, "f_prod_i"_a, "fscat"_a,
             "poles"_a, "L"_a, "Mpair"_a);
    py::class_<Resonances::Spline, ResonancePdf>(mres, "Spline")
        .def(py::init([](std::string name, Variable ar, Variable ai, std::vector<fptype> limits, std::vector<Variable> re,
                         std::vector<Variable> im, unsigned int cyc, bool symmDP) {
                 return new Resonances::Spline(name, ar, ai, limits, re, im, cyc, symmDP);
             }),
             "name"_a, "ar"_a, "ai"_a, "HH_bin_limits"_a, "pwa_coefs_reals"_a, "pwa_coefs_imags"_a, "cyc"_a, "symmDP"_a = false)
        .def_static("secondDerivatives", &Resonances::Spline::secondDerivatives, "x"_a, "y"_a);
    py::class_<Resonances::RhoOmegaMix, ResonancePdf>(mres, "RhoOmegaMix")
        .def(py::init<std::string, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, Variable, unsigned int,
                      unsigned int, bool, bool>(),
             "name"_a, "ar"_a, "ai"_a, "omega_mass"_a, "omega_width"_a, "rho_mass"_a, "rho_width"_a, "real"_a, "imag"_a, "delta"_a,
             "sp"_a, "cyc"_a, "norm"_a, "sym"_a);
    py::class_<DecayInfo3>(m, "DecayInfo3")
        .def(py::init<>())
        .def_readwrite("motherMass", &DecayInfo3::motherMass)
        .def_readwrite("daug1Mass", &DecayInfo3::daug1Mass)
        .def_readwrite("daug2Mass", &DecayInfo3::daug2Mass)
        .def_readwrite("daug3Mass", &DecayInfo3::daug3Mass)
        .def_readwrite("meson_radius", &DecayInfo3::meson_radius)
        .def_readwrite("mother_meson_radius", &DecayInfo3::mother_meson_radius)
        .def_readwrite("resonances", &DecayInfo3::resonances)
        .def("add_resonance", [](DecayInfo3 &d, ResonancePdf *r) { d.resonances.push_back(r); }, py::keep_alive<1, 2>());
    py::class_<Amp3BodyBase, GooPdf>(m, "Amp3BodyBase");
    py::class_<Amp3Body, Amp3BodyBase>(m, "Amp3Body")
        .def(py::init<std::string, Observable, Observable, EventNumber, DecayInfo3, GooPdf *>(), "name"_a, "m12"_a, "m13"_a,
             "eventNumber"_a, "decay"_a, "eff"_a, py::keep_alive<1, 6>(), py::keep_alive<1, 7>())
        .def("setDataSize", &Amp3Body::setDataSize, "dataSize"_a, "evtSize"_a = 3, "offset"_a = 0)
        .def("setForceIntegrals", &Amp3Body::setForceIntegrals, "f"_a = true)
        .def("getCachedWave", [](const Amp3Body &a, size_t i) { return py::array(py::cast(a.getCachedWave(i))); }, "i"_a)
        .def("sumCachedWave", &Amp3Body::sumCachedWave, "i"_a)
        .def("getIntegrals", [](const Amp3Body &a) { return py::array(py::cast(a.getIntegrals())); })
        .def("fit_fractions", &Amp3Body::fit_fractions)
        .def("setGenerationOffset", &Amp3Body::setGenerationOffset, "off"_a)
        .def("getGenerationOffset", &Amp3Body::getGenerationOffset)
        // upstream's binding exports three of the four components of each particle (python/PDFs/physics/Amp3Body.cpp:49-62)
        .def(
            "GenerateSig",
            [](Amp3Body &self, size_t numEvents, int seed) { return generated_to_numpy(self.GenerateSigColumns(numEvents, seed), 3); },
            "numEvents"_a, "seed"_a = 0)
        .def("getDecayInfo", &Amp3Body::getDecayInfo, py::return_value_policy::reference_internal);
    py::class_<DalitzPlotter>(m, "DalitzPlotter")
        .def(py::init<GooPdf *, Amp3Body *>(), "overallSignal"_a, "signalDalitz"_a, py::keep_alive<1, 2>(), py::keep_alive<1, 3>())
        .def("fillDataSetMC", &DalitzPlotter::fillDataSetMC, "dataset"_a, "nTotal"_a, "seed"_a = 0)
        .def("getNumEvents", &DalitzPlotter::getNumEvents)
        .def("getX", &DalitzPlotter::getX)
        .def("getY", &DalitzPlotter::getY)
        .def("getXval", &DalitzPlotter::getXval)
        .def("getYval", &DalitzPlotter::getYval)
        .def("getZval", &DalitzPlotter::getZval)
        .def("getVal", &DalitzPlotter::getVal, "event"_a, "num"_a = 0)
        .def("getDataSet", &DalitzPlotter::getDataSet, py::return_value_policy::reference_internal)
        .def("getM12", &DalitzPlotter::getM12)
        .def("getM13", &DalitzPlotter::getM13)
        .def("make2D", [](const DalitzPlotter &self) { // python/PDFs/physics/DalitzPlotter.cpp:31-49: bins without an evaluation stay 0
                 const size_t n12 = self.getM12().getNumBins(), n13 = self.getM13().getNumBins();
                 py::array_t<fptype> result({n12, n13});
                 std::fill(result.mutable_data(), result.mutable_data() + n12 * n13, 0.0);
                 for(size_t j = 0; j < self.getNumEvents(); ++j) result.mutable_at(self.getX(j), self.getY(j)) = self.getVal(j);
                 return result;
             }, "Make a 2D array for plotting")
        .def("getExtent", [](const DalitzPlotter &self) {
                 return std::vector<double>{self.getM12().getLowerLimit(), self.getM12().getUpperLimit(), self.getM13().getLowerLimit(),
                                            self.getM13().getUpperLimit()};
             }, "Get the extents for plotting");

    init_physics4(m);
    // pre-2.2 class names the reference's bindings keep as module attributes (python/PDFs/physics/Amp3Body.cpp:79, Amp3Body_TD.cpp:129)
    m.attr("DalitzPlotPdf") = m.attr("Amp3Body");
    m.attr("TddpPdf") = m.attr("Amp3Body_TD");

    // ---- fitting ----
    init_minuit2(m); // goofit_b200.minuit2: FCNBase, MnUserParameters, MnMigrad, FunctionMinimum, ... (python/Minuit2/*.cpp upstream)
    m.attr("FunctionMinimum") = m.attr("minuit2").attr("FunctionMinimum");
    py::class_<Params>(m, "Params")
        .def(py::init<PdfBase &>(), "pdf"_a, py::keep_alive<1, 2>())
        .def("size", &Params::size)
        .def("GetGooFitParams", &Params::GetGooFitParams)
        .def("make_minuit_vector", &Params::make_minuit_vector)
        .def("from_minuit_vector", &Params::from_minuit_vector, "values"_a, "force_changed"_a = false)
        .def("set_record", &Params::set_record, "do_record"_a = true)
        .def("get_recorded", &Params::get_recorded)
        .def("SetGooFitParams", &Params::SetGooFitParams, "input"_a)
        .def_property("record", nullptr, [](Params &p, bool on) { p.set_record(on); })
        .def_property_readonly("recorded", &Params::get_recorded);
    py::class_<FCN, Minuit2::FCNBase>(m, "FCN")
        .def(py::init<Params &>(), "params"_a, py::keep_alive<1, 2>())
        .def("__call__", (double (FCN::*)(const std::vector<double> &) const) & FCN::operator())
        .def("__call__", (double (FCN::*)() const) & FCN::operator())
        .def("Up", &FCN::Up);
    py::class_<FitManagerMinuit2>(m, "FitManager")
        .def(py::init<PdfBase *>(), py::keep_alive<1, 2>())
        .def("fit", &FitManagerMinuit2::fit, py::call_guard<py::scoped_ostream_redirect>())
        .def("getParams", &FitManagerMinuit2::getParams, py::return_value_policy::reference_internal)
        .def("getFCN", &FitManagerMinuit2::getFCN, py::return_value_policy::reference_internal)
        .def("getMinosErrors", &FitManagerMinuit2::getMinosErrors)
        .def("setMaxCalls", &FitManagerMinuit2::setMaxCalls)
        .def("setVerbosity", &FitManagerMinuit2::setVerbosity)
        .def("getVerbosity", &FitManagerMinuit2::getVerbosity)
        .def("setMinos", &FitManagerMinuit2::setMinos, "minos_flag"_a = true)
        .def("setStrategy", &FitManagerMinuit2::setStrategy)
        .def("lastFitSeconds", &FitManagerMinuit2::lastFitSeconds)
        .def("lastFitCalls", &FitManagerMinuit2::lastFitCalls)
        .def("__int__", &FitManagerMinuit2::operator int)
        .def("__bool__", &FitManagerMinuit2::operator bool)
        .def("__nonzero__", &FitManagerMinuit2::operator bool)
        .def("getMnScan", &FitManagerMinuit2::getMnScan)
        .def_property("verbosity", &FitManagerMinuit2::getVerbosity, &FitManagerMinuit2::setVerbosity)
        .def_property("max_calls", nullptr, &FitManagerMinuit2::setMaxCalls)
        .def_property_readonly("params", &FitManagerMinuit2::getParams, py::return_value_policy::reference_internal)
        .def_property_readonly("fcn", &FitManagerMinuit2::getFCN, py::return_value_policy::reference_internal);
    m.attr("FitManagerMinuit2") = m.attr("FitManager"); // the C++ class name; `FitManager` is what upstream's module exposes
}
