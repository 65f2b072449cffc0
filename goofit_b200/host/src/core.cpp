// core.cpp — host model API over the C ABI: Variable helpers, datasets, the Engine (program flattening
// and upload), PdfBase and GooPdf. Reference behaviour cited per function (paths under /root/reference).
#include <goofit/detail/Core.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <numeric>
#include <random>
#include <thread>

namespace GooFit {

// ---- Variable helpers (src/goofit/Variable.cpp) ---------------------------------------------------
int max_index(const std::vector<Variable> &vars) {
    int m = -1;
    for(auto &v : vars) m = std::max(m, v.getIndex());
    return m;
}
int max_index(const std::vector<Observable> &vars) {
    int m = -1;
    for(auto &v : vars) m = std::max(m, v.getIndex());
    return m;
}
int max_fitter_index(const std::vector<Variable> &vars) {
    int m = -1;
    for(auto &v : vars) m = std::max(m, v.getFitterIndex());
    return m;
}
std::ostream &operator<<(std::ostream &o, const Variable &v) {
    o << v.getName() << ": " << v.getValue() << " +/- " << v.getError();
    if(!v.IsFixed()) o << " [" << v.getLowerLimit() << ", " << v.getUpperLimit() << "]";
    if(v.getIndex() >= 0) o << " GooFit index: " << v.getIndex();
    if(v.getFitterIndex() >= 0) o << " Fitter index: " << v.getFitterIndex();
    if(v.IsFixed()) o << " Fixed";
    return o;
}
std::ostream &operator<<(std::ostream &o, const Observable &v) {
    o << v.getName() << ": " << v.getValue() << " (" << v.getNumBins() << " bins)";
    o << " [" << v.getLowerLimit() << ", " << v.getUpperLimit() << "]";
    if(v.getIndex() >= 0) o << " GooFit index: " << v.getIndex();
    return o;
}

std::istream &operator>>(std::istream &i, Observable &v) {
    fptype x;
    if(i >> x) v.setValue(x);
    return i;
}

void abort(std::string file, int line, std::string reason, const PdfBase *pdf) {
    // same message shape as src/goofit/Abort.cpp:10-42 (parameter dump, then throw)
    std::ostringstream o;
    o << "Abort called from " << file << " line " << line << " due to " << reason;
    if(pdf) {
        o << "\nParameters of " << pdf->getName() << ":";
        for(const Variable &v : pdf->getParameters()) o << "\n  " << v.getName() << " = " << v.getValue();
    }
    std::cerr << o.str() << std::endl;
    throw GeneralError(reason);
}

// ---- DataSet (src/goofit/DataSet.cpp) -------------------------------------------------------------
static std::string join_names(const std::vector<Observable> &obs) {
    std::string s;
    for(size_t i = 0; i < obs.size(); i++) s += (i ? ", " : "") + obs[i].getName();
    return s;
}
DataSet::DataSet(const Observable &var, std::string n)
    : name(std::move(n))
    , observables({var}) {
    if(name.empty()) name = join_names(observables);
}
DataSet::DataSet(const std::vector<Observable> &vars, std::string n)
    : name(std::move(n))
    , observables(vars) {
    if(name.empty()) name = join_names(observables);
}
DataSet::DataSet(std::initializer_list<Observable> vars, std::string n)
    : name(std::move(n))
    , observables(vars) {
    if(name.empty()) name = join_names(observables);
}
void DataSet::addWeightedEvent(fptype) { throw GeneralError("AddWeightedEvent not implemented for this type of DataSet"); }
std::vector<fptype> DataSet::getCurrentValues() const {
    std::vector<fptype> v;
    for(auto &o : observables) v.push_back(o.getValue());
    return v;
}
size_t DataSet::indexOfVariable(const Observable &var) const {
    for(size_t i = 0; i < observables.size(); ++i)
        if(var == observables[i]) return i;
    throw GeneralError("Invalid variable access into dataset!");
}
void DataSet::checkAllVars() const {
    for(const Observable &v : observables)
        if(!v.inRange()) throw OutOfRange(v.getName(), v.getValue(), v.getLowerLimit(), v.getUpperLimit());
}

// ---- UnbinnedDataSet (src/goofit/UnbinnedDataSet.cpp) ----------------------------------------------
UnbinnedDataSet::UnbinnedDataSet(const Observable &var, std::string n)
    : DataSet(var, std::move(n))
    , data(1) {}
UnbinnedDataSet::UnbinnedDataSet(const std::vector<Observable> &vars, std::string n)
    : DataSet(vars, std::move(n))
    , data(vars.size()) {}
UnbinnedDataSet::UnbinnedDataSet(std::initializer_list<Observable> vars, std::string n)
    : DataSet(vars, std::move(n))
    , data(vars.size()) {}
void UnbinnedDataSet::addEvent() {
    checkAllVars();
    for(size_t i = 0; i < observables.size(); i++) data[i].push_back(observables[i].getValue());
    numEventsAdded++;
}
void UnbinnedDataSet::clear() {
    for(auto &c : data) c.clear();
    numEventsAdded = 0;
}
fptype UnbinnedDataSet::getValue(const Observable &var, size_t idx) const {
    if(idx >= getNumEvents())
        throw GeneralError("UnbinnedDataSet: Attempted to find {} in event {} when only {} events exits",
                           var.getName(), idx, getNumEvents());
    return data[indexOfVariable(var)][idx];
}
void UnbinnedDataSet::loadEvent(size_t idx) {
    for(size_t i = 0; i < observables.size(); i++) observables[i].setValue(data.at(i).at(idx));
}
void UnbinnedDataSet::setValueForAllEvents(const Observable &var) {
    auto &c = data[indexOfVariable(var)];
    std::fill(c.begin(), c.end(), var.getValue());
}
void UnbinnedDataSet::appendColumns(const std::vector<const fptype *> &cols, size_t n) {
    if(cols.size() != data.size()) throw GeneralError("appendColumns: expected {} columns", data.size());
    for(size_t k = 0; k < data.size(); k++) data[k].insert(data[k].end(), cols[k], cols[k] + n);
    numEventsAdded += n;
}
// grid of bin centres, last observable slowest... the reference recurses from the back of the list
// (UnbinnedDataSet.cpp:60-89), which makes the FIRST observable the fastest; event numbers count up.
void UnbinnedDataSet::fillWithGrid() {
    for(auto &c : data)
        if(!c.empty()) throw GeneralError("The dataset must be empty to fill with grid!");
    std::vector<size_t> dims;
    std::vector<size_t> which;
    int evt = -1;
    for(size_t i = 0; i < observables.size(); i++) {
        if(observables[i].isEventNumber()) { evt = (int)i; continue; }
        which.push_back(i);
        dims.push_back(observables[i].getNumBins());
    }
    size_t total = 1;
    for(size_t d : dims) total *= d;
    fptype evtval = evt >= 0 ? observables[evt].getValue() : 0;
    for(size_t g = 0; g < total; g++) {
        size_t r = g;
        for(size_t k = 0; k < which.size(); k++) {
            Observable &o = observables[which[k]];
            size_t b = r % dims[k];
            r /= dims[k];
            o.setValue(o.getLowerLimit() + (o.getUpperLimit() - o.getLowerLimit()) * (b + 0.5) / o.getNumBins());
        }
        if(evt >= 0) observables[evt].setValue(evtval);
        addEvent();
        evtval += 1;
    }
    if(evt >= 0) observables[evt].setValue(evtval);
}

// ---- BinnedDataSet (src/goofit/BinnedDataSet.cpp) ---------------------------------------------------
BinnedDataSet::BinnedDataSet(const Observable &var, std::string n)
    : DataSet(var, std::move(n)) {
    collectBins();
    binvalues.resize(getNumBins());
}
BinnedDataSet::BinnedDataSet(const std::vector<Observable> &vars, std::string n)
    : DataSet(vars, std::move(n)) {
    collectBins();
    binvalues.resize(getNumBins());
}
BinnedDataSet::BinnedDataSet(std::initializer_list<Observable> vars, std::string n)
    : DataSet(vars, std::move(n)) {
    collectBins();
    binvalues.resize(getNumBins());
}
void BinnedDataSet::collectBins() {
    binsizes.clear();
    for(auto &o : observables) binsizes.push_back(o.getNumBins());
}
void BinnedDataSet::addEvent() { addWeightedEvent(1.0); }
void BinnedDataSet::addWeightedEvent(double w) {
    checkAllVars();
    binvalues.at(getBinNumber()) += w;
    numEventsAdded++;
}
size_t BinnedDataSet::getBinNumber() const {
    size_t ret = 0, prior = 1;
    for(size_t i = 0; i < observables.size(); i++) {
        const Observable &o = observables[i];
        fptype v = std::min(std::max(o.getValue(), o.getLowerLimit()), o.getUpperLimit());
        if(v != o.getValue())
            GOOFIT_INFO("Warning: Value {} outside {} range [{},{}] - clamping to {}", o.getValue(), o.getName(),
                        o.getLowerLimit(), o.getUpperLimit(), v);
        auto local = static_cast<size_t>(std::floor((v - o.getLowerLimit()) / getBinSize(i)));
        ret += local * prior;
        prior *= binsizes[i];
    }
    return ret;
}
std::vector<size_t> BinnedDataSet::globalToLocal(size_t global) const {
    std::vector<size_t> l;
    for(size_t i = 0; i < observables.size(); i++) {
        l.push_back(global % binsizes[i]);
        global /= binsizes[i];
    }
    return l;
}
fptype BinnedDataSet::getBinCenter(size_t ivar, size_t bin) const {
    size_t local = globalToLocal(bin).at(ivar);
    fptype ret   = getBinSize(ivar);
    ret *= (local + 0.5);
    ret += observables[ivar].getLowerLimit();
    return ret;
}
fptype BinnedDataSet::getBinCenter(const Observable &var, size_t bin) const { return getBinCenter(indexOfVariable(var), bin); }
fptype BinnedDataSet::getBinSize(size_t ivar) const {
    return (observables.at(ivar).getUpperLimit() - observables[ivar].getLowerLimit()) / binsizes[ivar];
}
fptype BinnedDataSet::getBinVolume(size_t) const {
    fptype r = 1;
    for(size_t i = 0; i < observables.size(); i++) r *= getBinSize(i);
    return r;
}
fptype BinnedDataSet::getBinError(size_t bin) const {
    return binerrors.empty() ? std::sqrt(binvalues.at(bin)) : binerrors.at(bin);
}
void BinnedDataSet::setBinError(unsigned int bin, fptype e) {
    if(binerrors.empty()) binerrors.resize(binvalues.size());
    binerrors.at(bin) = e;
}
size_t BinnedDataSet::getNumBins() const {
    size_t n = 1;
    for(size_t b : binsizes) n *= b;
    return n;
}
fptype BinnedDataSet::getNumWeightedEvents() const { return std::accumulate(binvalues.begin(), binvalues.end(), 0.0); }

std::string FitControl::getName() const {
    return evalfunc_to_string(metricFunc);
}

// ---- Engine -----------------------------------------------------------------------------------------
namespace detail {
Engine &Engine::get() {
    static Engine e;
    return e;
}
void Engine::check(int rc, const char *what) {
    if(rc != 0) throw GeneralError("{}: {}", what, gf_last_error());
}
void Engine::setDevice(int dev) {
    if(ctx_ && dev != device_) throw GeneralError("The CUDA device must be chosen before the first PDF is evaluated");
    device_ = dev;
}
gf_ctx *Engine::ctx() {
    if(ctx_) return ctx_;
    if(device_ < 0) {
        const char *e = std::getenv("GOOFIT_B200_DEVICE");
        if(!e) e = std::getenv("LOCAL_RANK");
        device_ = e ? std::atoi(e) : 0;
    }
    ctx_ = gf_create(device_);
    if(!ctx_) throw GeneralError("goofit_b200: cannot create a CUDA context ({}). There is no CPU path.", gf_last_error());
    return ctx_;
}
void Engine::shutdown() {
    if(ctx_) gf_destroy(ctx_);
    ctx_ = nullptr;
    root = nullptr;
    members.clear();
}
int ProgramBuilder::allocAux() {
    if(next_aux >= 16) throw GeneralError("Too many cached PDF components (max 16 convolution/amplitude caches)");
    return next_aux++;
}
void Engine::build(PdfBase *r) {
    for(PdfBase *m : members) {
        m->nodeIdx = -1;
        m->normSlot = -1;
        m->parOffs.clear();
    }
    members.clear();
    prog.clear();
    root = r;
    r->emitAux(prog);
    r->emit(prog);
    gf_ctx *c = ctx();
    check(gf_reset_aux(c), "gf_reset_aux");
    check(gf_set_program(c, prog.nodes.data(), (int)prog.nodes.size(), prog.consts.data(), (int)prog.consts.size(),
                         prog.ints.data(), (int)prog.ints.size(), (int)prog.params.size(), (int)prog.norms.size()),
          "gf_set_program");
    paramsDirty = normsDirty = true;
    generation++;
    std::vector<PdfBase *> m = members; // hooks may not modify the list
    for(PdfBase *p : m) p->afterProgram();
}
void Engine::sync() {
    gf_ctx *c = ctx();
    // one stream-ordered copy of the parameter + normalisation block through the pinned ring: no host-side wait
    if(paramsDirty || normsDirty)
        check(gf_set_params_async(c, prog.params.data(), (int)prog.params.size(), prog.norms.data(), (int)prog.norms.size()),
              "gf_set_params_async");
    paramsDirty = normsDirty = false;
}
void Engine::setShard(int rank, int size) {
    if(size < 1 || rank < 0 || rank >= size) throw GeneralError("setShard: need 0 <= rank < size, got rank {} of {}", rank, size);
    shard_ = {rank, size};
}
// Sharding and the reduction are one operation (upstream pairs its MPI split with MPI_Allreduce, GooPdf.cu:72-75): a
// sharded dataset without a communicator of the same size would hand each rank a different, partial likelihood.
void Engine::requireComm(const char *what) {
    if(shard_.size <= 1) return;
    int n = gf_comm_attached(ctx());
    if(n != shard_.size)
        throw GeneralError("{}: the dataset is sharded over {} ranks but {} — attach one with gf_comm_init / "
                           "Engine::attachCommFromEnv() before evaluating, or do not shard",
                           what, shard_.size, n == 0 ? std::string("no communicator is attached") : "the communicator has " + std::to_string(n) + " ranks");
}
void Engine::attachCommFromEnv(int rank, int size) {
    if(size < 1 || rank < 0 || rank >= size) throw GeneralError("attachCommFromEnv: need 0 <= rank < size, got rank {} of {}", rank, size);
    if(size == 1) return;
    gf_ctx *c = ctx();
    if(gf_comm_attached(c) == size) {
        setShard(rank, size);
        return;
    }
    // the 128-byte NCCL id travels through a file: rank 0 writes it under a temporary name and renames it into place
    // (atomic on one filesystem), the other ranks poll for it. One node only, which is what this library targets.
    const char *dir = std::getenv("GOOFIT_B200_RENDEZVOUS_DIR");
    const char *port = std::getenv("MASTER_PORT"), *run = std::getenv("TORCHELASTIC_RUN_ID");
    std::string path = std::string(dir ? dir : "/tmp") + "/goofit_b200_nccl_" + (port ? port : "0") + "_" + (run ? run : "none") + "_"
                       + std::to_string(commEpoch_++) + ".id";
    char id[128];
    if(rank == 0) {
        check(gf_comm_unique_id(id), "gf_comm_unique_id");
        std::string tmp = path + ".tmp";
        {
            std::ofstream f(tmp, std::ios::binary | std::ios::trunc);
            f.write(id, sizeof id);
            if(!f) throw GeneralError("attachCommFromEnv: cannot write {}", tmp);
        }
        if(std::rename(tmp.c_str(), path.c_str()) != 0) throw GeneralError("attachCommFromEnv: cannot publish {}", path);
    } else {
        auto t0 = std::chrono::steady_clock::now();
        for(;;) {
            std::ifstream f(path, std::ios::binary);
            if(f && f.read(id, sizeof id) && f.gcount() == (std::streamsize)sizeof id) break;
            if(std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120)
                throw GeneralError("attachCommFromEnv: rank {} waited 120 s for {} (is rank 0 running?)", rank, path);
            std::this_thread::sleep_for(std::chrono::milliseconds(20));
        }
    }
    check(gf_comm_init(c, id, rank, size), "gf_comm_init"); // collective: returns once every rank has joined
    if(rank == 0) std::remove(path.c_str());
    setShard(rank, size);
}
void Engine::release(const PdfBase *pdf) {
    auto it = std::find(members.begin(), members.end(), pdf);
    if(it != members.end()) members.erase(it);
    if(root == pdf) root = nullptr;
}
} // namespace detail
using detail::Engine;

// ---- PdfBase ------------------------------------------------------------------------------------------
PdfBase::~PdfBase() {
    ++structureEpoch();
    Engine::get().release(this);
}

template <typename T>
static bool find_in(const std::vector<T> &list, const T &item) {
    return std::find(list.begin(), list.end(), item) != list.end();
}
std::vector<Variable> PdfBase::getParameters() const { // src/goofit/PdfBase.cpp:56-70
    std::vector<Variable> ret = parametersList;
    for(const PdfBase *comp : components)
        for(const Variable &v : comp->getParameters())
            if(!find_in(ret, v)) ret.push_back(v);
    return ret;
}
Variable *PdfBase::getParameterByName(std::string n) {
    for(Variable &p : parametersList)
        if(p.getName() == n) return &p;
    for(auto *c : components)
        if(Variable *v = c->getParameterByName(n)) return v;
    return nullptr;
}
std::vector<Observable> PdfBase::getObservables() const { // PdfBase.cpp:88-100
    std::vector<Observable> ret = observablesList;
    for(const PdfBase *comp : components)
        for(const Observable &o : comp->getObservables())
            if(!find_in(ret, o)) ret.push_back(o);
    return ret;
}
void PdfBase::registerObservable(Observable obs) {
    if(!find_in(observablesList, obs)) observablesList.push_back(obs);
}
void PdfBase::collectParametersByName(const std::string &n, std::vector<Variable *> &out) {
    for(auto &p : parametersList)
        if(p.getName() == n && std::none_of(out.begin(), out.end(), [&](const Variable *q) { return *q == p; })) out.push_back(&p);
    for(auto *c : components) c->collectParametersByName(n, out);
}
void PdfBase::unregisterParameter(Variable var) {
    ++structureEpoch();
    auto it = std::find(parametersList.begin(), parametersList.end(), var);
    if(it != parametersList.end()) parametersList.erase(it);
    for(auto *c : components) c->unregisterParameter(var);
}
void PdfBase::setupObservables() { // PdfBase.cpp:109-117
    int counter = 0;
    for(auto &o : observablesList) o.setIndex(counter++);
}
bool PdfBase::parametersChanged() const { // PdfBase.cpp:124-127: own parameters only
    return std::any_of(parametersList.begin(), parametersList.end(), [](const Variable &v) { return v.getChanged(); });
}
void PdfBase::updateVariable(Variable var, fptype newValue) { // PdfBase.cu:121-129 (match by NAME)
    for(auto &p : parametersList)
        if(p.getName() == var.getName()) p.setValue(newValue);
    for(auto *c : components) c->updateVariable(var, newValue);
}
void PdfBase::updateParameters() { // PdfBase.cu:131-140; one upload per call instead of one write per element
    Engine &E = Engine::get();
    for(int off : parOffs)
        for(size_t i = 0; i < parametersList.size(); i++) E.prog.params[off + i] = parametersList[i].getValue();
    for(auto *c : components) c->updateParameters();
    E.paramsDirty = true;
}
fptype PdfBase::par(size_t i) const {
    if(nodeIdx >= 0 && !parOffs.empty()) return Engine::get().prog.params[parOffs.back() + i];
    return parametersList.at(i).getValue();
}
void PdfBase::setNormalization(const fptype &v) {
    cachedNormalization = v;
    if(normSlot >= 0) {
        Engine &E = Engine::get();
        E.prog.norms[normSlot] = v;
        E.normsDirty           = true;
    }
}
void PdfBase::recursiveSetNormalization(fptype norm, bool) { // src/goofit/PdfBase.cpp:30-39
    setNormalization(norm);
    for(auto *c : components) c->recursiveSetNormalization(norm, true);
}
bool PdfBase::areParamsandConstantsEqualByVal(const PdfBase &o) const {
    if(parametersList.size() != o.parametersList.size() || constantsList != o.constantsList) return false;
    for(size_t i = 0; i < parametersList.size(); i++)
        if(!parametersList[i].isEqualNameValLimitsErrBlindFixed(o.parametersList[i])) return false;
    return true;
}

int PdfBase::emitRecord(detail::ProgramBuilder &B, int op, int nchild, int size, const std::vector<int32_t> &extra) {
    if(op == 0) throw GeneralError("PDF {} has no B200 opcode (registerFunction was not called)", getName());
    gf_node nd{};
    nd.op = op; nd.nchild = nchild; nd.size = size;
    nd.par_off = (int)B.params.size(); nd.npar = (int)parametersList.size();
    for(auto &p : parametersList) B.params.push_back(p.getValue());
    nd.con_off = (int)B.consts.size(); nd.ncon = (int)constantsList.size();
    for(fptype c : constantsList) B.consts.push_back(c);
    nd.obs_off = (int)B.ints.size(); nd.nobs = (int)observablesList.size();
    for(auto &o : observablesList) B.ints.push_back(o.getIndex());
    nd.norm_slot = (int)B.norms.size();
    B.norms.push_back(cachedNormalization);
    nd.int_off = (int)B.ints.size(); nd.nint = (int)extra.size();
    for(int32_t v : extra) B.ints.push_back(v);
    B.nodes.push_back(nd);
    nodeIdx  = (int)B.nodes.size() - 1;
    normSlot = nd.norm_slot;
    parOffs.push_back(nd.par_off);
    Engine &E = Engine::get();
    if(!find_in(E.members, this)) E.members.push_back(this);
    return nodeIdx;
}
void PdfBase::emitAux(detail::ProgramBuilder &B) {
    for(auto *c : components) c->emitAux(B);
}
int PdfBase::emit(detail::ProgramBuilder &B) {
    int first = (int)B.nodes.size();
    std::vector<int32_t> slots;
    for(auto *c : components) slots.push_back(B.nodes[c->emit(B)].norm_slot);
    return emitRecord(B, opcode_, (int)components.size(), (int)B.nodes.size() - first + 1, slots);
}
void PdfBase::setIndices() { Engine::get().build(this); }

// Dataset upload (PdfBase.cu:329-530). Columns stay columnar: no AoS transpose, no per-element
// name lookups; the library stages them through pinned memory into 256-B aligned SoA columns.
void PdfBase::setData(DataSet *data) {
    setupObservables();
    setIndices();
    data_ = data;
    if(data == nullptr) return;
    Engine &E      = Engine::get();
    detail::Shard sh = E.shard();
    std::vector<std::vector<fptype>> owned;
    std::vector<const fptype *> cols;
    size_t first = 0, count = 0;
    auto split = [&](size_t n) { // PdfBase.cu:364-378: N/P each, remainder on the last rank
        size_t per = n / sh.size;
        first = per * sh.rank;
        count = sh.rank == sh.size - 1 ? n - first : per;
    };
    if(auto *ub = dynamic_cast<UnbinnedDataSet *>(data)) {
        numEntries = (unsigned int)data->getNumEvents();
        numEvents  = numEntries;
        split(numEntries);
        for(const Observable &o : observablesList) {
            const std::vector<fptype> &c = ub->column(ub->indexOfVariable(o));
            if(o.isEventNumber() && sh.rank > 0) { // re-based to the local row (PdfBase.cu:408-415)
                owned.emplace_back(count);
                std::iota(owned.back().begin(), owned.back().end(), 0.0);
                cols.push_back(nullptr);
            } else {
                // a null pointer marks an owned column below: an empty dataset's vector may legitimately have none, so point at a
                // dummy instead (no element of it is ever read: count is 0)
                static const fptype none = 0;
                cols.push_back(c.empty() ? &none : c.data() + first);
            }
        }
    } else if(auto *bd = dynamic_cast<BinnedDataSet *>(data)) {
        numEvents  = 0;
        numEntries = (unsigned int)bd->getNumBins();
        if(!fitControl || !fitControl->binnedFit()) setFitControl(std::make_shared<BinnedNllFit>());
        split(numEntries);
        size_t nobs = observablesList.size();
        owned.assign(nobs + 2, std::vector<fptype>(count));
        for(size_t i = 0; i < numEntries; i++) numEvents += bd->getBinContent(i);
        for(size_t j = 0; j < count; j++) {
            size_t i = first + j;
            for(size_t k = 0; k < nobs; k++) owned[k][j] = bd->getBinCenter(observablesList[k], i);
            owned[nobs][j]     = bd->getBinContent(i);
            owned[nobs + 1][j] = fitControl->binErrors() ? bd->getBinError(i) : bd->getBinVolume(i);
        }
        cols.assign(nobs + 2, nullptr);
    } else
        throw GeneralError("Dataset must be binned or unbinned!");
    size_t k = 0;
    for(auto &c : cols)
        if(c == nullptr) c = owned[k++].data();
    E.localEvents = count;
    E.firstEvent  = first;
    E.totalEvents = numEntries;
    Engine::check(gf_upload_columns(E.ctx(), cols.data(), (int)cols.size(), count, (double)numEntries), "gf_upload_columns");
}

std::ostream &operator<<(std::ostream &out, const PdfBase &pdf) {
    out << "GooPdf::" << pdf.getPdfName() << "(\"" << pdf.getName() << "\") :\n";
    out << "  Device function: " << pdf.reflex_name_ << "\n";
    out << "  Observable" << (pdf.observablesList.size() > 1 ? "s" : "") << ": ";
    for(size_t i = 0; i < pdf.observablesList.size(); i++) out << (i ? ", " : "") << pdf.observablesList[i].getName();
    out << "\n  Parameter" << (pdf.parametersList.size() > 1 ? "s" : "") << ": ";
    for(size_t i = 0; i < pdf.parametersList.size(); i++) out << (i ? ", " : "") << pdf.parametersList[i].getName();
    out << "\n";
    for(auto *c : pdf.components) out << "  Component: " << c->getName() << "\n";
    return out;
}

// toy generation by inverse CDF on the observable's grid (src/goofit/PdfBase.cpp:154-203)
void PdfBase::fillMCDataSimple(size_t events, unsigned int seed) {
    if(observablesList.size() != 1) throw GeneralError("You can only fill MC on a 1D dataset with a simple fill");
    Observable var = observablesList.at(0);
    UnbinnedDataSet grid{var};
    grid.fillWithGrid();
    auto *orig = getData();
    if(orig == nullptr) throw GeneralError("Can't run on a PDF with no DataSet to fill!");
    setData(&grid);
    std::vector<double> pdfValues = dynamic_cast<GooPdf *>(this)->getCompProbsAtDataPoints()[0];
    if(seed == 0) seed = std::random_device{}();
    std::mt19937 gen(seed);
    std::uniform_real_distribution<> unihalf(-.5, .5), uniwhole(0.0, 1.0);
    std::vector<double> integral(pdfValues.size());
    std::partial_sum(pdfValues.begin(), pdfValues.end(), integral.begin());
    for(double &v : integral) v /= integral.back();
    for(size_t i = 0; i < events; i++) {
        double r = uniwhole(gen);
        size_t j = std::lower_bound(integral.begin(), integral.end(), r) - integral.begin();
        var.setValue(grid.getValue(var, j) + var.getBinSize() * unihalf(gen));
        orig->addEvent();
    }
    setData(orig);
}

// ---- GooPdf (src/PDFs/GooPdf.cu) ------------------------------------------------------------------------
int GooPdf::metric() const {
    return static_cast<int>(fitControl ? fitControl->getMetric() : EvalFunc::NLL); // gf_metric == EvalFunc order
}
void GooPdf::setIndices() {
    if(!fitControl) setFitControl(std::make_shared<UnbinnedNllFit>()); // GooPdf.cu:183-186
    PdfBase::setIndices();
}
void GooPdf::setFitControl(std::shared_ptr<FitControl> fc) {
    fitControl = fc;
    for(auto *c : components) c->setFitControl(fc);
}
void GooPdf::setParameterConstantness(bool constant) {
    for(Variable &v : getParameters()) v.setFixed(constant);
}
fptype GooPdf::integrate(fptype, fptype) const { return 0; }

int GooPdf::get_bin_grid_size() const {
    size_t total = 1;
    for(const Observable &v : observablesList) total *= integrationBins > 0 ? (size_t)integrationBins : v.getNumBins();
    return (int)total;
}
fptype GooPdf::get_bin_grid_volume() const {
    fptype ret = 1.0;
    for(const Observable &v : observablesList) ret *= v.getUpperLimit() - v.getLowerLimit();
    return ret;
}
static void require_in_program(const PdfBase *p, int nodeIdx) {
    if(nodeIdx < 0)
        throw GeneralError("PDF {} is not part of the flattened program; call setData() on the top-level PDF first", p->getName());
}
double GooPdf::reduce_with_bins() const { // GooPdf.cu:83-108 -> k_grid_sum
    require_in_program(this, nodeIdx);
    std::vector<double> spec;
    for(const Observable &v : observablesList) { // PdfBase.cu:532-557 generateNormRange
        spec.push_back(v.getLowerLimit());
        spec.push_back(v.getUpperLimit());
        spec.push_back(integrationBins > 0 ? integrationBins : (double)v.getNumBins());
    }
    Engine &E = Engine::get();
    E.sync();
    double out = 0;
    Engine::check(gf_integrate_grid(E.ctx(), nodeIdx, spec.data(), (int)observablesList.size(), &out), "gf_integrate_grid");
    return out;
}
double GooPdf::reduce_with_metric() const { // GooPdf.cu:45-78 -> k_eval_sum (+ NCCL all-reduce when sharded)
    require_in_program(this, nodeIdx);
    Engine &E = Engine::get();
    E.requireComm("reduce_with_metric");
    E.sync();
    double out = 0;
    Engine::check(gf_eval_sum(E.ctx(), nodeIdx, metric(), &out), "gf_eval_sum");
    return out;
}
std::vector<fptype> GooPdf::evaluate_with_metric() const { // GooPdf.cu:112-173 -> k_eval_events
    require_in_program(this, nodeIdx);
    Engine &E = Engine::get();
    E.sync();
    std::vector<fptype> out(E.localEvents);
    Engine::check(gf_eval_events(E.ctx(), nodeIdx, metric(), out.data(), out.size()), "gf_eval_events");
    detail::Shard sh = E.shard();
    if(sh.size > 1) { // every rank returns the values of ALL events, in event order (the reference's MPI_Allgatherv, GooPdf.cu:139-172)
        E.requireComm("evaluate_with_metric");
        std::vector<size_t> counts(sh.size);
        size_t per = E.totalEvents / sh.size;
        for(int r = 0; r < sh.size; r++) counts[r] = r == sh.size - 1 ? E.totalEvents - per * r : per;
        std::vector<fptype> all(E.totalEvents);
        Engine::check(gf_allgather_events(E.ctx(), out.data(), counts.data(), all.data()), "gf_allgather_events");
        return all;
    }
    return out;
}

fptype GooPdf::normalize() { // GooPdf.cu:340-396
    if(!fitControl) setFitControl(std::make_shared<UnbinnedNllFit>());
    if(!fitControl->metricIsPdf()) {
        setNormalization(1.0);
        return 1.0;
    }
    fptype ret = 1;
    if(hasAnalyticIntegral()) {
        for(const Observable &v : observablesList) ret *= integrate(v.getLowerLimit(), v.getUpperLimit());
        setNormalization(1.0 / ret);
        return ret;
    }
    size_t totalBins = get_bin_grid_size();
    ret *= get_bin_grid_volume();
    ret /= totalBins;
    // A leaf's grid sum is a pure function of its pushed parameter block and of the grid; the reference re-sums it on
    // every call (GooPdf.cu:366-388). Re-running the kernel only when one of those inputs moved returns the same
    // number and saves a launch + readback per likelihood call for every fixed-shape component (e.g. a background).
    fptype sum;
    std::vector<double> key;
    if(components.empty() && nodeIdx >= 0) {
        key.push_back(static_cast<double>(Engine::get().generation));
        key.push_back(static_cast<double>(integrationBins));
        for(fptype cst : constantsList) key.push_back(cst);
        for(size_t i = 0; i < parametersList.size(); ++i) key.push_back(par(i));
        for(const Observable &v : observablesList) {
            key.push_back(v.getLowerLimit());
            key.push_back(v.getUpperLimit());
            key.push_back(static_cast<double>(v.getNumBins()));
        }
    }
    if(!key.empty() && key == gridKey)
        sum = gridSum;
    else {
        sum     = reduce_with_bins();
        gridKey = key;
        gridSum = sum;
    }
    if(std::isnan(sum))
        GooFit::abort(__FILE__, __LINE__, getName() + " NaN in normalization", this);
    else if(0 >= sum)
        GooFit::abort(__FILE__, __LINE__, "Non-positive normalization", this);
    ret *= sum;
    if(0 == ret) GooFit::abort(__FILE__, __LINE__, "Zero integral");
    setNormalization(1.0 / ret);
    return ret;
}

double GooPdf::calculateNLL() { // GooPdf.cu:254-273
    normalize();
    if(cachedNormalization <= 0) GooFit::abort(__FILE__, __LINE__, getName() + " non-positive normalization", this);
    double ret = reduce_with_metric();
    if(0.0 == ret) GooFit::abort(__FILE__, __LINE__, getName() + " zero NLL", this);
    return ret;
}

std::vector<fptype> GooPdf::evaluateAtPoints(Observable var) { // GooPdf.cu:275-312
    setFitControl(std::make_shared<EvalFit>());
    setIndices();
    normalize();
    UnbinnedDataSet tempdata(observablesList);
    double step = var.getBinSize();
    for(size_t i = 0; i < var.getNumBins(); ++i) {
        var.setValue(var.getLowerLimit() + (i + 0.5) * step);
        tempdata.addEvent();
    }
    auto *old = getData();
    setData(&tempdata);
    normalize();
    std::vector<fptype> res = evaluate_with_metric();
    for(auto &v : res) v *= cachedNormalization;
    setData(old);
    return res;
}

fptype GooPdf::getValue(EvalFunc evalfunc) { // GooPdf.cu:314-338
    if(evalfunc == EvalFunc::Prob)
        setFitControl(std::make_shared<ProbFit>());
    else if(evalfunc == EvalFunc::Eval)
        setFitControl(std::make_shared<EvalFit>());
    else if(evalfunc == EvalFunc::NLL)
        setFitControl(std::make_shared<UnbinnedNllFit>());
    else
        throw GeneralError("That EvalFunc is not supported");
    setIndices();
    normalize();
    UnbinnedDataSet point(observablesList);
    point.addEvent();
    auto *old = getData();
    setData(&point);
    normalize(); // the re-flattened program starts from the cached norms; keep them current
    auto results = evaluate_with_metric();
    setData(old);
    return results[0];
}

// Per-event probabilities of the root and of each direct component (GooPdf.cu:402-435). The reference
// re-flattens every component as its own root; here a component is evaluated in place as a subtree.
std::vector<std::vector<fptype>> GooPdf::getCompProbsAtDataPoints() {
    auto fc = fitControl;
    setFitControl(std::make_shared<ProbFit>());
    normalize();
    std::vector<std::vector<fptype>> values(components.size() + 1);
    values[0] = evaluate_with_metric();
    for(size_t i = 0; i < components.size(); ++i) {
        auto *g = dynamic_cast<GooPdf *>(components[i]);
        if(!g) throw GeneralError("Component {} is not a GooPdf", components[i]->getName());
        g->normalize();
        values[1 + i] = g->evaluate_with_metric();
    }
    setFitControl(fc);
    return values;
}

UnbinnedDataSet GooPdf::makeGrid() {
    UnbinnedDataSet ret{observablesList};
    ret.fillWithGrid();
    return ret;
}

void GooPdf::scan(Observable var, std::vector<fptype> &values) {
    fptype step = var.getUpperLimit();
    step -= var.getLowerLimit();
    step /= var.getNumBins();
    values.clear();
    for(fptype v = var.getLowerLimit() + 0.5 * step; v < var.getUpperLimit(); v += step) {
        var.setValue(v);
        copyParams();
        values.push_back(calculateNLL());
    }
}

} // namespace GooFit
