// goofit.minuit2 of the reference (python/Minuit2/main.cpp:1-27 and the eleven class files next to it): a Python interface to
// the bundled Minuit2 - FCNBase (overridable from Python), MinuitParameter, MnUserParameters, MnUserParameterState,
// MnUserCovariance, FunctionMinimum, MnApplication, MnMigrad, MnMinos, MnScan, MnPrint - with the reference's names and
// signatures. Built into the _goofit extension as the submodule `goofit_b200.minuit2`, so that GooFit's FCN (a FCNBase) and the
// FunctionMinimum FitManager.fit() returns are the very types this module exposes.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <sstream>

#include <Minuit2/FCNBase.h>
#include <Minuit2/FunctionMinimum.h>
#include <Minuit2/MinuitParameter.h>
#include <Minuit2/MnApplication.h>
#include <Minuit2/MnMigrad.h>
#include <Minuit2/MinosError.h>
#include <Minuit2/MnMinos.h>
#include <Minuit2/MnPrint.h>
#include <Minuit2/MnScan.h>
#include <Minuit2/MnStrategy.h>
#include <Minuit2/MnUserCovariance.h>
#include <Minuit2/MnUserParameterState.h>
#include <Minuit2/MnUserParameters.h>

namespace py = pybind11;
using namespace pybind11::literals;
using namespace ROOT::Minuit2;

namespace {
class PyFCNBase : public FCNBase { // python/Minuit2/FCNBase.cpp:10-21: a likelihood written in Python
  public:
    using FCNBase::FCNBase;
    double operator()(const std::vector<double> &v) const override { PYBIND11_OVERLOAD_PURE_NAME(double, FCNBase, "__call__", operator(), v); }
    double Up() const override { PYBIND11_OVERLOAD_PURE(double, FCNBase, Up, ); }
    void SetErrorDef(double v) override { PYBIND11_OVERLOAD(void, FCNBase, SetErrorDef, v); }
    double ErrorDef() const override { PYBIND11_OVERLOAD(double, FCNBase, ErrorDef, ); }
};
} // namespace

void init_minuit2(py::module &parent) {
    py::module m = parent.def_submodule("minuit2", "Python interface for Minuit2");
    py::class_<FCNBase, PyFCNBase>(m, "FCNBase")
        .def(py::init<>())
        .def("__call__", &FCNBase::operator())
        .def("SetErrorDef", &FCNBase::SetErrorDef)
        .def("ErrorDef", &FCNBase::ErrorDef)
        .def("Up", &FCNBase::Up);
    py::class_<MinuitParameter>(m, "MinuitParameter")
        .def("Error", &MinuitParameter::Error)
        .def("Fix", &MinuitParameter::Fix)
        .def("GetName", &MinuitParameter::GetName)
        .def("HasLimits", &MinuitParameter::HasLimits)
        .def("HasLowerLimit", &MinuitParameter::HasLowerLimit)
        .def("HasUpperLimit", &MinuitParameter::HasUpperLimit)
        .def("IsConst", &MinuitParameter::IsConst)
        .def("IsFixed", &MinuitParameter::IsFixed)
        .def("LowerLimit", &MinuitParameter::LowerLimit)
        .def("Name", &MinuitParameter::Name)
        .def("Number", &MinuitParameter::Number)
        .def("Release", &MinuitParameter::Release)
        .def("RemoveLimits", &MinuitParameter::RemoveLimits)
        .def("SetError", &MinuitParameter::SetError)
        .def("SetLimits", &MinuitParameter::SetLimits)
        .def("SetLowerLimit", &MinuitParameter::SetLowerLimit)
        .def("SetName", &MinuitParameter::SetName)
        .def("SetUpperLimit", &MinuitParameter::SetUpperLimit)
        .def("SetValue", &MinuitParameter::SetValue)
        .def("UpperLimit", &MinuitParameter::UpperLimit)
        .def("Value", &MinuitParameter::Value);
    py::class_<MnUserParameterState>(m, "MnUserParameterState")
        .def("Value", (double (MnUserParameterState::*)(unsigned int) const) & MnUserParameterState::Value, "n"_a)
        .def("Error", (double (MnUserParameterState::*)(unsigned int) const) & MnUserParameterState::Error, "n"_a)
        .def("Params", &MnUserParameterState::Params)
        .def("Errors", &MnUserParameterState::Errors);
    py::class_<MnUserParameters>(m, "MnUserParameters")
        .def(py::init<>())
        .def("Add", (bool (MnUserParameters::*)(const std::string &, double)) & MnUserParameters::Add)
        .def("Add", (bool (MnUserParameters::*)(const std::string &, double, double)) & MnUserParameters::Add)
        .def("Add", (bool (MnUserParameters::*)(const std::string &, double, double, double, double)) & MnUserParameters::Add)
        .def("Parameter", &MnUserParameters::Parameter, "Get a parameter by number")
        .def("Parameters", &MnUserParameters::Parameters, "Get all the parameters");
    py::class_<MnUserCovariance>(m, "MnUserCovariance")
        .def("Nrow", &MnUserCovariance::Nrow)
        .def("size", &MnUserCovariance::size)
        .def("__call__", (double (MnUserCovariance::*)(unsigned int, unsigned int) const) & MnUserCovariance::operator())
        .def("to_matrix", [](const MnUserCovariance &self) {
            size_t n = self.Nrow();
            py::array_t<double> matrix({n, n});
            for(size_t i = 0; i < n; i++)
                for(size_t j = 0; j < n; j++) matrix.mutable_at(i, j) = self(i, j);
            return matrix;
        });
    py::class_<FunctionMinimum>(m, "FunctionMinimum")
        .def("UserState", &FunctionMinimum::UserState)
        .def("UserParameters", &FunctionMinimum::UserParameters)
        .def("UserCovariance", &FunctionMinimum::UserCovariance)
        .def("Fval", &FunctionMinimum::Fval)
        .def("Edm", &FunctionMinimum::Edm)
        .def("NFcn", &FunctionMinimum::NFcn, "Number of function calls")
        .def("Up", &FunctionMinimum::Up)
        .def("IsValid", &FunctionMinimum::IsValid)
        .def("HasValidParameters", &FunctionMinimum::HasValidParameters)
        .def("HasValidCovariance", &FunctionMinimum::HasValidCovariance)
        .def("HasAccurateCovar", &FunctionMinimum::HasAccurateCovar)
        .def("HasPosDefCovar", &FunctionMinimum::HasPosDefCovar)
        .def("HasMadePosDefCovar", &FunctionMinimum::HasMadePosDefCovar)
        .def("HesseFailed", &FunctionMinimum::HesseFailed)
        .def("HasCovariance", &FunctionMinimum::HasCovariance)
        .def("IsAboveMaxEdm", &FunctionMinimum::IsAboveMaxEdm)
        .def("HasReachedCallLimit", &FunctionMinimum::HasReachedCallLimit)
        .def("values", [](const FunctionMinimum &f) { return f.UserState().Params(); })
        .def("errors", [](const FunctionMinimum &f) { return f.UserState().Errors(); })
        .def("__str__", [](const FunctionMinimum &self) {
            std::stringstream os;
            os << self;
            return os.str();
        });
    py::class_<MnApplication>(m, "MnApplication")
        .def("__call__", &MnApplication::operator(), "Minimize the function, returns a function minimum", "maxfcn"_a = 0, "tolerance"_a = 0.1);
    py::class_<MnMigrad, MnApplication>(m, "MnMigrad")
        .def(py::init([](const FCNBase &fcn, const MnUserParameters &par, unsigned int stra) { return new MnMigrad(fcn, par, MnStrategy(stra)); }), "fcn"_a,
             "par"_a, "stra"_a = 1, py::keep_alive<1, 2>());
    py::class_<MnScan, MnApplication>(m, "MnScan")
        .def(py::init([](const FCNBase &fcn, const MnUserParameters &par, unsigned int stra) { return new MnScan(fcn, par, MnStrategy(stra)); }), "fcn"_a,
             "par"_a, "stra"_a = 1, py::keep_alive<1, 2>())
        .def("Scan", &MnScan::Scan, "Scan parameter.", "par"_a, "maxsteps"_a = 41, "low"_a = 0, "high"_a = 0);
    py::class_<MinosError>(m, "MinosError") // python/Minuit2/MnMinos.cpp:17-35
        .def_property_readonly("number", &MinosError::Parameter)
        .def_property_readonly("lower", &MinosError::Lower)
        .def_property_readonly("upper", &MinosError::Upper)
        .def_property_readonly("is_valid", &MinosError::IsValid)
        .def_property_readonly("lower_valid", &MinosError::LowerValid)
        .def_property_readonly("upper_valid", &MinosError::UpperValid)
        .def_property_readonly("at_lower_limit", &MinosError::AtLowerLimit)
        .def_property_readonly("at_upper_limit", &MinosError::AtUpperLimit)
        .def_property_readonly("at_lower_max_fcn", &MinosError::AtLowerMaxFcn)
        .def_property_readonly("at_upper_max_fcn", &MinosError::AtUpperMaxFcn)
        .def_property_readonly("lower_new_min", &MinosError::LowerNewMin)
        .def_property_readonly("upper_new_min", &MinosError::UpperNewMin)
        .def_property_readonly("nfcn", &MinosError::NFcn)
        .def_property_readonly("min", &MinosError::Min);
    py::class_<MnMinos>(m, "MnMinos")
        .def(py::init([](const FCNBase &fcn, const FunctionMinimum &min, unsigned int stra) { return new MnMinos(fcn, min, MnStrategy(stra)); }), "fcn"_a,
             "min"_a, "stra"_a = 1, py::keep_alive<1, 2>(), py::keep_alive<1, 3>())
        .def("__call__", [](const MnMinos &self, unsigned int n, unsigned int maxcalls, double toler) { return self(n, maxcalls, toler); }, "n"_a,
             "maxcalls"_a = 0, "toler"_a = 0.1, "the (lower, upper) MINOS errors of parameter n")
        .def("Minos", [](const MnMinos &self, unsigned int n, unsigned int maxcalls, double toler) { return self.Minos(n, maxcalls, toler); }, "n"_a,
             "maxcalls"_a = 0, "toler"_a = 0.1, "the full MinosError of parameter n");
    py::class_<MnPrint> m_ls(m, "m_ls");
    m_ls.def_static("GlobalLevel", &MnPrint::GlobalLevel);
    m_ls.def_static("SetGlobalLevel", &MnPrint::SetGlobalLevel);
    m.attr("MnPrint") = m_ls;
}
