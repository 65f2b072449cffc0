// Core.h — host side of the B200 drop-in: GooFit's public model API (Variable/Observable, datasets,
// FitControl, PdfBase/GooPdf) re-implemented over the C ABI in include/goofit_b200.h.
// The class and method names follow the reference so user code compiles unchanged
// (/root/reference/include/goofit/{Variable.h,DataSet.h,UnbinnedDataSet.h,BinnedDataSet.h,FitControl.h,PdfBase.h},
//  include/goofit/PDFs/GooPdf.h); everything below the API is new: the PDF tree is flattened to a
// post-order node program (no function-pointer tables) and all evaluation goes to the CUDA library.
#pragma once
#include <climits>
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <initializer_list>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "goofit_b200.h"
#include <goofit/Version.h>
#ifdef ROOT_FOUND
#include <TH1D.h>
#endif

#ifndef __host__
#define __host__
#endif
#ifndef __device__
#define __device__
#endif

// upstream's headers pull in {fmt} (goofit/Log.h); user code that calls fmt::print after including a goofit header keeps compiling where {fmt} is on
// the include path (this library itself formats its messages without it)
#if defined(__has_include)
#if __has_include(<fmt/format.h>)
#include <fmt/format.h>
#endif
#endif

namespace GooFit {

using fptype    = double;
using fpcomplex = std::complex<double>;

// ---- errors (reference: include/goofit/Error.h:10-19, src/goofit/Abort.cpp:10-42) ---------------
namespace detail {
inline void fmt_into(std::ostringstream &o, const char *s) { o << s; }
template <typename T, typename... R>
void fmt_into(std::ostringstream &o, const char *s, const T &v, const R &...rest) {
    for(; *s; ++s) {
        if(s[0] == '{' && s[1] == '}') {
            o << v;
            return fmt_into(o, s + 2, rest...);
        }
        o << *s;
    }
}
} // namespace detail

struct GeneralError : public std::runtime_error {
    template <typename... Args>
    explicit GeneralError(const std::string &name, const Args &...args)
        : std::runtime_error(format(name, args...)) {}
    template <typename... Args>
    static std::string format(const std::string &f, const Args &...args) {
        std::ostringstream o;
        detail::fmt_into(o, f.c_str(), args...);
        return o.str();
    }
    int get_exit_code() const { return 2; }
};
struct OutOfRange : public GeneralError {
    OutOfRange(const std::string &name, double value, double min, double max)
        : GeneralError("{} value {} is not in range {} to {}", name, value, min, max) {}
};

class PdfBase;
[[noreturn]] void abort(std::string file, int line, std::string reason, const PdfBase *pdf = nullptr);

#define GOOFIT_INFO(...) (std::cout << ::GooFit::GeneralError::format(__VA_ARGS__) << std::endl)
#define GOOFIT_WARN(...) (std::cout << ::GooFit::GeneralError::format(__VA_ARGS__) << std::endl)
#define GOOFIT_ERROR(...) (std::cerr << ::GooFit::GeneralError::format(__VA_ARGS__) << std::endl)
#define GOOFIT_STATUS(...) ((void)0)
#define GOOFIT_DEBUG(...) ((void)0)
#define GOOFIT_TRACE(...) ((void)0)

// ---- Variable / Observable / EventNumber (reference Variable.h:22-263) ---------------------------
class Params;
class Minuit1;
class Indexable {
  protected:
    struct State {
        fptype value = 0, lo = 0, hi = 0;
        int index = -1;
    };
    std::string name;
    std::shared_ptr<State> st;

  public:
    Indexable(std::string n, fptype val = 0)
        : name(std::move(n))
        , st(std::make_shared<State>()) {
        st->value = val;
    }
    Indexable(std::string n, fptype val, fptype lowerlimit, fptype upperlimit)
        : Indexable(std::move(n), val) {
        st->lo = lowerlimit;
        st->hi = upperlimit;
    }
    virtual ~Indexable() = default;
    int getIndex() const { return st->index; }
    void setIndex(int v) { st->index = v; }
    const std::string &getName() const { return name; }
    fptype getValue() const { return st->value; }
    void setValue(fptype v) { st->value = v; }
    fptype getUpperLimit() const { return st->hi; }
    void setUpperLimit(fptype v) { st->hi = v; }
    fptype getLowerLimit() const { return st->lo; }
    void setLowerLimit(fptype v) { st->lo = v; }
    const Indexable &operator=(const fptype &v) {
        setValue(v);
        return *this;
    }
    operator fptype() const { return getValue(); }
    // identity = shared state, exactly like the reference's pointer comparison
    bool operator<(const Indexable &o) const { return st.get() < o.st.get(); }
    bool operator==(const Indexable &o) const { return st.get() == o.st.get(); }
    bool operator!=(const Indexable &o) const { return st.get() != o.st.get(); }
    bool inRange() const { return st->hi == st->lo || (st->value <= st->hi && st->value >= st->lo); }
    explicit operator bool() const { return inRange(); }
    bool isEqualNameValLimits(const Indexable &o) const {
        return name == o.name && st->value == o.st->value && st->lo == o.st->lo && st->hi == o.st->hi;
    }
};

class Observable : public Indexable {
  protected:
    std::shared_ptr<size_t> numbins{std::make_shared<size_t>(100)};
    bool is_event_number{false};

  public:
    Observable(std::string name, fptype lowerlimit, fptype upperlimit)
        : Indexable(std::move(name), 0, lowerlimit, upperlimit) {}
    void setNumBins(size_t n) { *numbins = n; }
    size_t getNumBins() const { return *numbins; }
    fptype getBinSize() const { return (getUpperLimit() - getLowerLimit()) / getNumBins(); }
    bool isEventNumber() const { return is_event_number; }
    const Observable &operator=(const fptype &v) {
        setValue(v);
        return *this;
    }
};

class Variable : public Indexable {
    struct Fit {
        fptype error = 0, blind = 0;
        bool changed = true, fixed = false;
        int fitter_index = -1;
    };
    std::shared_ptr<Fit> fit{std::make_shared<Fit>()};

  public:
    class Key {
        friend Params;
        friend Minuit1; // goofit/fitting/FitManagerMinuit1.h (builds with ROOT)
        Key() = default;
    };
    Variable(const Variable &) = default;
    /// constant: fixed, limits straddle the value (reference Variable.h:148-152)
    Variable(std::string n, fptype v)
        : Indexable(std::move(n), v, v + 0.01, v - 0.01) {
        fit->error = 0.002;
        fit->fixed = true;
    }
    /// value and error, no limits; the reference marks this form fixed as well (Variable.h:155-159)
    Variable(std::string n, fptype v, fptype err)
        : Indexable(std::move(n), v, 0, 0) {
        fit->error = err;
        fit->fixed = true;
    }
    Variable(std::string n, fptype v, fptype dn, fptype up)
        : Indexable(std::move(n), v, dn, up) {
        fit->error = 0.1 * (up - dn);
    }
    Variable(std::string n, fptype v, fptype e, fptype dn, fptype up, bool fix = false)
        : Indexable(std::move(n), v, dn, up) {
        fit->error = e;
        fit->fixed = fix;
    }
    fptype getError() const { return fit->error; }
    void setError(fptype v) { fit->error = v; }
    int getFitterIndex() const { return fit->fitter_index; }
    void setFitterIndex(int v) { fit->fitter_index = v; }
    bool getChanged() const { return fit->changed; }
    bool IsFixed() const { return fit->fixed; }
    void setFixed(bool f) { fit->fixed = f; }
    void setChanged(bool v = true) { fit->changed = v; }
    void setBlind(fptype v) { fit->blind = v; }
    fptype getBlind(const Key &) { return fit->blind; }
    const Variable &operator=(const fptype &v) {
        setValue(v);
        return *this;
    }
    bool isEqualNameValLimitsErrBlindFixed(const Variable &o) const {
        return isEqualNameValLimits(o) && fit->error == o.fit->error && fit->blind == o.fit->blind
               && fit->fixed == o.fit->fixed;
    }
};
std::ostream &operator<<(std::ostream &o, const Variable &v);
std::ostream &operator<<(std::ostream &o, const Observable &v);
/// reads the next number of the stream into the observable's value (Variable.cpp:53)
std::istream &operator>>(std::istream &i, Observable &v);

class EventNumber : public Observable {
  public:
    static constexpr fptype maxint{static_cast<fptype>(1LL << std::numeric_limits<fptype>::digits)};
    EventNumber(std::string name, fptype min = 0, fptype max = EventNumber::maxint)
        : Observable(std::move(name), min, max) {
        is_event_number = true;
        st->value       = 0;
    }
    const EventNumber &operator=(const fptype &v) {
        setValue(v);
        return *this;
    }
};

int max_index(const std::vector<Variable> &vars);
int max_index(const std::vector<Observable> &vars);
int max_fitter_index(const std::vector<Variable> &vars);

// ---- datasets (reference DataSet.h, UnbinnedDataSet.h, BinnedDataSet.h) ---------------------------
class DataSet {
  public:
    DataSet(const Observable &var, std::string n = "");
    DataSet(const std::vector<Observable> &vars, std::string n = "");
    DataSet(std::initializer_list<Observable> vars, std::string n = "");
    virtual ~DataSet() = default;
    virtual void addEvent() = 0;
    virtual void addWeightedEvent(fptype weight);
    template <typename... Args>
    void addEvent(fptype value, Args... args) {
        std::vector<fptype> values{value, static_cast<fptype>(args)...};
        if(values.size() != observables.size())
            throw GeneralError("You must pass the correct number of values ({}) to addEvent", observables.size());
        for(size_t i = 0; i < values.size(); i++) observables[i].setValue(values[i]);
        addEvent();
    }
    const std::vector<Observable> &getObservables() const { return observables; }
    size_t numVariables() const { return observables.size(); }
    size_t getNumEvents() const { return numEventsAdded; }
    std::string getName() const { return name; }
    size_t indexOfVariable(const Observable &var) const;

  protected:
    std::vector<fptype> getCurrentValues() const;
    void checkAllVars() const;
    size_t numEventsAdded{0};
    std::string name;
    std::vector<Observable> observables;
};

class UnbinnedDataSet : public DataSet {
  public:
    using DataSet::addEvent;
    UnbinnedDataSet(const Observable &var, std::string n = "");
    UnbinnedDataSet(const std::vector<Observable> &vars, std::string n = "");
    UnbinnedDataSet(std::initializer_list<Observable> vars, std::string n = "");
    void addEvent() override;
    void clear();
    void fillWithGrid();
    fptype getValue(const Observable &var, size_t idx) const;
    void loadEvent(size_t idx);
    void setValueForAllEvents(const Observable &var);
    /// matrix-like input: input(j, i) = variable j of event i; an omitted trailing column is the running index
    template <typename M>
    void from_matrix(const M &input, bool filter = false) {
        size_t next = getNumEvents();
        size_t rows = static_cast<size_t>(input.rows());
        if(observables.size() != rows && observables.size() != rows + 1)
            throw GeneralError("The wrong number of rows, expected {}, but matrix had {}", observables.size(), rows);
        for(long i = 0; i < static_cast<long>(input.cols()); i++) {
            for(size_t j = 0; j < rows; j++) observables[j].setValue(input(j, i));
            if(observables.size() == rows + 1) observables[rows].setValue(static_cast<fptype>(next++));
            bool ok = true;
            if(filter)
                for(auto &o : observables) ok = ok && o.inRange();
            if(ok) addEvent();
        }
    }
    template <typename M>
    M to_matrix() const {
        M mat(data.size(), data.at(0).size());
        for(size_t i = 0; i < data.size(); i++)
            for(size_t j = 0; j < data[i].size(); j++) mat(i, j) = data[i][j];
        return mat;
    }
    /// bulk column access for the uploader and the Python bindings (columnar already, UnbinnedDataSet.h:85)
    const std::vector<fptype> &column(size_t ivar) const { return data.at(ivar); }
    void appendColumns(const std::vector<const fptype *> &cols, size_t n);

  private:
    std::vector<std::vector<fptype>> data;
};

class BinnedDataSet : public DataSet {
  public:
    using DataSet::addEvent;
    BinnedDataSet(const Observable &var, std::string n = "");
    BinnedDataSet(const std::vector<Observable> &vars, std::string n = "");
    BinnedDataSet(std::initializer_list<Observable> vars, std::string n = "");
    void addEvent() override;
    void addWeightedEvent(double weight) override;
    fptype getBinContent(size_t bin) const { return binvalues.at(bin); }
    fptype getBinCenter(size_t ivar, size_t bin) const;
    fptype getBinCenter(const Observable &var, size_t bin) const;
    fptype getBinSize(size_t ivar) const;
    size_t getBinNumber() const;
    fptype getBinVolume(size_t bin) const;
    fptype getBinError(size_t bin) const;
    std::vector<size_t> getDiminsions() const { return binsizes; }
    size_t getNumBins() const;
    fptype getNumWeightedEvents() const;
    void setBinContent(unsigned int bin, fptype value) { binvalues.at(bin) = value; }
    void setBinError(unsigned int bin, fptype value);

  private:
    void collectBins();
    std::vector<size_t> globalToLocal(size_t global) const;
    std::vector<size_t> binsizes;
    std::vector<fptype> binvalues, binerrors;
};

// ---- fit control (reference FitControl.h:11-76, PDFs/MetricPointer.h) ------------------------------
enum class EvalFunc : size_t { Eval = 0, NLL, Prob, BinAvg, BinWithError, Chisq };
/// FitControl.h:13-16 of the reference: the name upstream registers each metric's device function under
inline const char *evalfunc_to_string(EvalFunc val) {
    static const char *names[] = {"ptr_to_Eval", "ptr_to_NLL", "ptr_to_Prob", "ptr_to_BinAvg", "ptr_to_BinWithError", "ptr_to_Chisq"};
    return names[static_cast<size_t>(val)];
}

class FitControl {
  public:
    FitControl(bool bin, EvalFunc mn)
        : binned(bin)
        , metricFunc(mn) {}
    virtual ~FitControl() = default;
    bool binnedFit() const { return binned; }
    bool binErrors() const { return errorsOnBins; }
    bool metricIsPdf() const { return !errorsOnBins; }
    EvalFunc getMetric() const { return metricFunc; }
    std::string getName() const;

  protected:
    bool errorsOnBins{false};

  private:
    bool binned;
    EvalFunc metricFunc;
};
struct UnbinnedNllFit : FitControl {
    UnbinnedNllFit()
        : FitControl(false, EvalFunc::NLL) {}
};
struct BinnedNllFit : FitControl {
    BinnedNllFit()
        : FitControl(true, EvalFunc::BinAvg) {}
};
struct BinnedErrorFit : FitControl {
    BinnedErrorFit()
        : FitControl(true, EvalFunc::BinWithError) {
        errorsOnBins = true;
    }
};
struct BinnedChisqFit : FitControl {
    BinnedChisqFit()
        : FitControl(true, EvalFunc::Chisq) {}
};
struct ProbFit : FitControl {
    ProbFit()
        : FitControl(false, EvalFunc::Prob) {}
};
struct EvalFit : FitControl {
    EvalFit()
        : FitControl(false, EvalFunc::Eval) {}
};

// ---- the engine: one CUDA context + the currently flattened program -----------------------------
// Replaces the reference's five global flat tables (PDFs/detail/Globals.h:14-49). As there, one
// flattened tree and one dataset live at a time per process.
namespace detail {
struct ProgramBuilder {
    std::vector<gf_node> nodes;
    std::vector<double> params, consts, norms;
    std::vector<int32_t> ints;
    int next_aux = 0;
    void clear() {
        nodes.clear(); params.clear(); consts.clear(); norms.clear(); ints.clear();
        next_aux = 0;
    }
    int allocAux();
};

/// Event sharding for multi-GPU runs (replaces the GOOFIT_MPI split, PdfBase.cu:361-378): rank r of n
/// takes events [r*(N/n), ...) with the remainder on the last rank; partial sums are all-reduced
/// inside the C-ABI call (NCCL) when a communicator is attached to the context.
struct Shard {
    int rank = 0, size = 1;
};

class Engine {
  public:
    static Engine &get();
    gf_ctx *ctx();
    void setDevice(int dev);
    int device() const { return device_; }
    void setShard(int rank, int size); // throws unless 0 <= rank < size
    /// throws when the dataset is sharded but no communicator of the same size is attached to the context
    void requireComm(const char *what);
    /// rendezvous over a file (rank 0 writes the NCCL id, the others read it) keyed on MASTER_PORT / TORCHELASTIC_RUN_ID,
    /// then gf_comm_init + setShard: what a C++ program launched by torchrun / SLURM needs to run sharded
    void attachCommFromEnv(int rank, int size);
    Shard shard() const { return shard_; }
    void build(PdfBase *root);
    void sync();
    void release(const PdfBase *pdf);
    void shutdown();
    static void check(int rc, const char *what);

    ProgramBuilder prog;
    PdfBase *root = nullptr;
    std::vector<PdfBase *> members; // every node object in the current program
    bool paramsDirty = true, normsDirty = true;
    size_t localEvents = 0, firstEvent = 0, totalEvents = 0; // this rank's shard [firstEvent, firstEvent + localEvents) of totalEvents
    uint64_t generation = 0; // bumped by every build(); PDFs use it to notice stale device caches

  private:
    gf_ctx *ctx_ = nullptr;
    int device_  = -1;
    Shard shard_;
    int commEpoch_ = 0;
};
} // namespace detail

// ---- PdfBase (reference PdfBase.h:97-260) --------------------------------------------------------
namespace {
inline void filter_arguments(std::vector<Observable> &, std::vector<Variable> &) {}
template <typename... Args>
void filter_arguments(std::vector<Observable> &o, std::vector<Variable> &v, const Variable &var, Args... args);
template <typename... Args>
void filter_arguments(std::vector<Observable> &o, std::vector<Variable> &v, const Observable &obs, Args... args) {
    o.push_back(obs);
    filter_arguments(o, v, args...);
}
template <typename... Args>
void filter_arguments(std::vector<Observable> &o, std::vector<Variable> &v, const EventNumber &obs, Args... args) {
    o.push_back(obs);
    filter_arguments(o, v, args...);
}
template <typename... Args>
void filter_arguments(std::vector<Observable> &o, std::vector<Variable> &v, const Variable &var, Args... args) {
    v.push_back(var);
    filter_arguments(o, v, args...);
}
} // namespace

} // namespace GooFit
namespace ROOT { namespace Minuit2 { class FunctionMinimum; } }
namespace GooFit {

class PdfBase {
    friend std::ostream &operator<<(std::ostream &, const PdfBase &);
    friend class detail::Engine;

  public:
    template <typename... Args>
    explicit PdfBase(std::string pdf_name, std::string n, Args... args)
        : pdf_name_(std::move(pdf_name))
        , name(std::move(n)) {
        std::vector<Observable> obs;
        std::vector<Variable> vars;
        filter_arguments(obs, vars, args...);
        for(auto &ob : obs) registerObservable(ob);
        for(auto &var : vars) registerParameter(var);
    }
    virtual ~PdfBase();

    virtual double calculateNLL() = 0;
    virtual fptype normalize()    = 0;
    virtual void setFitControl(std::shared_ptr<FitControl>) = 0;
    virtual bool hasAnalyticIntegral() const { return false; }

    std::string getName() const { return name; }
    std::string getPdfName() const { return pdf_name_; }
    std::vector<Observable> getObservables() const;
    std::vector<Variable> getParameters() const;
    Variable *getParameterByName(std::string n);
    std::vector<PdfBase *> getComponents() { return components; }

    void setData(DataSet *data);
    DataSet *getData() { return data_; }
    void fillMCDataSimple(size_t events, unsigned int seed = 0);
    ROOT::Minuit2::FunctionMinimum fitTo(DataSet *data, int verbosity = 3);
    ROOT::Minuit2::FunctionMinimum fit(int verbosity = 3);

    void setNormalization(const fptype &v);
    fptype getNormalization() const { return cachedNormalization; }
    void setIntegrationFineness(int i) { integrationBins = i; }
    bool parametersChanged() const;
    void updateVariable(Variable v, fptype newValue);
    void updateParameters();
    void setupObservables();
    virtual void setIndices();
    void SigGenSetIndices() { setIndices(); }
    void initializeIndices() {}
    void copyParams() { updateParameters(); }
    void setCommonNorm(bool v = true) { commonNorm = v; }
    void setSeparateNorm(bool v = true) { separateNorm = v; }
    bool getCommonNorm() const { return commonNorm; }
    bool getSeparateNorm() const { return separateNorm; }
    /// index of this PDF's record in the flattened program (-1 when it is not part of it)
    int getFunctionIndex() const { return nodeIdx; }
    int getParameterIndex() const { return parOffs.empty() ? -1 : parOffs.back(); }
    int getNormalizationIndex() const { return normSlot; }
    bool areParamsandConstantsEqualByVal(const PdfBase &other) const;

    /// bumped whenever a parameter list of any PDF changes: holders of pointers into the lists (Params) re-resolve
    static uint64_t &structureEpoch() {
        static uint64_t e = 0;
        return e;
    }
    /// every parameter entry of this tree whose NAME is `n`, one per distinct shared state (what updateVariable touches)
    void collectParametersByName(const std::string &n, std::vector<Variable *> &out);

    // ---- flattening hooks (new; called by the Engine and by parent nodes, not by user code) ----
    /// emit records that are NOT evaluated per event (sub-PDFs of a convolution, resonances, ...)
    virtual void emitAux(detail::ProgramBuilder &B);
    /// emit the per-event subtree in post-order, return the index of this node's record
    virtual int emit(detail::ProgramBuilder &B);
    /// called after the program is on the device
    virtual void afterProgram() {}

  protected:
    void registerParameter(Variable var) {
        parametersList.push_back(var);
        ++structureEpoch();
    }
    void unregisterParameter(Variable var);
    void registerConstant(fptype value) { constantsList.push_back(value); }
    void registerObservable(Observable obs);
    /// the reference registers a device function pointer here; the B200 build records an opcode instead
    void registerFunction(std::string reflex_name, int op) {
        reflex_name_ = std::move(reflex_name);
        opcode_      = op;
    }
    void recursiveSetNormalization(fptype norm = 1.0, bool subpdf = false);
    void initialize() {}
    /// value of parameter i as last pushed by updateParameters() — the reference reads host_parameters[parametersIdx+i+1]
    /// (not the Variable) in integrate()/normalize(), which matters after a fit has rewritten the Variables
    fptype par(size_t i) const;

    int emitRecord(detail::ProgramBuilder &B, int op, int nchild, int size, const std::vector<int32_t> &extraInts);

    DataSet *data_ = nullptr;
    std::string reflex_name_;
    int opcode_ = 0;
    fptype numEvents{0};        // sum of weights / entries (global, all shards)
    unsigned int numEntries{0}; // rows (global)
    fptype cachedNormalization{1.0};
    std::shared_ptr<FitControl> fitControl;
    std::vector<Observable> observablesList;
    std::vector<Variable> parametersList;
    std::vector<fptype> constantsList;
    std::vector<PdfBase *> components;
    int integrationBins{-1};
    bool properlyInitialised{true};
    bool commonNorm{false};
    bool separateNorm{false};

    // position in the current program
    int nodeIdx = -1, normSlot = -1;
    std::vector<int> parOffs; // one per occurrence of this object in the tree

  private:
    std::string pdf_name_;
    std::string name;
};
std::ostream &operator<<(std::ostream &, const PdfBase &);

// ---- GooPdf (reference PDFs/GooPdf.h:27-103, src/PDFs/GooPdf.cu) -----------------------------------
class GooPdf : public PdfBase {
  public:
    using PdfBase::PdfBase;
    ~GooPdf() override = default;

    double calculateNLL() override;
    fptype normalize() override;
    fptype normalise() { return normalize(); }
    virtual fptype integrate(fptype lo, fptype hi) const;
    bool hasAnalyticIntegral() const override { return false; }
    std::vector<fptype> evaluateAtPoints(Observable var);
    fptype getValue(EvalFunc evalfunc = EvalFunc::Eval);
    std::vector<std::vector<fptype>> getCompProbsAtDataPoints();
    UnbinnedDataSet makeGrid();
    void scan(Observable var, std::vector<fptype> &values);
    void setFitControl(std::shared_ptr<FitControl> fc) override;
    void setParameterConstantness(bool constant = true);
    /// GooPdf.cu:237-252 upstream switches printf debugging inside its device functions on; the kernels here carry none: accepted, no effect
    void setDebugMask(int, bool = true) const {}
    int get_bin_grid_size() const;
    fptype get_bin_grid_volume() const;
    double reduce_with_bins() const;
    void setIndices() override;
    /// per-event metric values of this (sub)tree on the uploaded data (GooPdf.cu:112-173)
    std::vector<fptype> evaluate_with_metric() const;
    /// sum over events of the metric (GooPdf.cu:45-78), all-reduced over shards
    double reduce_with_metric() const;
#ifdef ROOT_FOUND
    /// GooPdf.cpp:126-144 of the reference: the PDF over `var` as a ROOT histogram, normalised to normFactor (header-only here: the
    /// library itself is built without ROOT). The caller owns the histogram.
    TH1D *plotToROOT(Observable var, double normFactor = 1, std::string name = "") {
        if(name.empty()) name = getName() + "_hist";
        auto *ret = new TH1D(name.c_str(), "", (int)var.getNumBins(), var.getLowerLimit(), var.getUpperLimit());
        std::vector<fptype> binValues = evaluateAtPoints(var);
        double pdf_int = 0;
        for(size_t i = 0; i < var.getNumBins(); ++i) pdf_int += binValues[i];
        for(size_t i = 0; i < var.getNumBins(); ++i) ret->SetBinContent((int)i + 1, binValues[i] * normFactor / pdf_int / var.getBinSize());
        return ret;
    }
#endif

  protected:
    int metric() const;
    // memo of the last leaf grid integral (normalize()): inputs it depended on, and the sum
    std::vector<double> gridKey;
    fptype gridSum{0};
};

using device_function_ptr = int; // source compatibility for user code that names the type
} // namespace GooFit
