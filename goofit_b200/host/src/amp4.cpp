// amp4.cpp — host half of the four-body amplitude PDF (reference: src/PDFs/physics/Amp4Body.cu:94-558,
// Amplitude.cu, SpinFactors.cu:570-640, lineshapes/{Lineshape,BW,Bugg}.cu). Device work lives in csrc/gf_amp4.cu.
#include <goofit/detail/Physics4.h>

#include <functional>
#include <numeric>

namespace GooFit {
using detail::Engine;

// ---- lineshapes ------------------------------------------------------------------------------------------
Lineshape::Lineshape(std::string pdf_name, std::string name, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius, int kind)
    : AmpComponent("Lineshapes::" + pdf_name, name)
    , _L(L)
    , _Mpair(Mpair)
    , _FormFac(FormFac)
    , _radius(radius)
    , _kind(kind) {
    registerConstant(Mpair); // Lineshape.cu:21-35
    registerFunction("ptr_to_" + pdf_name, GF_OP_LS);
}
bool Lineshape::isEqualByValue(const Lineshape &o) const { // Lineshape.cu:37-41
    return getName() == o.getName() && _L == o._L && _Mpair == o._Mpair && _FormFac == o._FormFac && _radius == o._radius
           && areParamsandConstantsEqualByVal(o);
}
namespace Lineshapes {
// kinds: gf_ls_kind in csrc/gf_amp4.cuh {RBW, SBW, LASS, BUGG, BUGG3, ONE, NONRES, FLATTE, GLASS, FOCUS}
#define GOOFIT_B200_LS_BW(NAME, KIND)                                                                                  \
    NAME::NAME(std::string name, Variable mass, Variable width, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius) \
        : Lineshape(#NAME, name, L, Mpair, FormFac, radius, KIND) {                                                    \
        registerParameter(mass);                                                                                       \
        registerParameter(width);                                                                                      \
        registerConstant(L);                                                                                           \
        registerConstant(static_cast<unsigned int>(FormFac));                                                          \
        registerConstant(radius);                                                                                      \
    }
GOOFIT_B200_LS_BW(RBW, 0)
GOOFIT_B200_LS_BW(SBW, 1)
GOOFIT_B200_LS_BW(LASS, 2)
#undef GOOFIT_B200_LS_BW
Bugg::Bugg(std::string name, Variable, Variable, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("Bugg", name, L, Mpair, FormFac, radius, 3) {}
Bugg3::Bugg3(std::string name, Variable, Variable, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("Bugg3", name, L, Mpair, FormFac, radius, 4) {}
// The shapes below keep the common constant block {Mpair, L, FormFac, radius} so the device dispatch reads one layout.
#define GOOFIT_B200_LS_CONSTS                                                                                          \
    registerConstant(L);                                                                                               \
    registerConstant(static_cast<unsigned int>(FormFac));                                                              \
    registerConstant(radius);
One::One(std::string name, Variable, Variable, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("One", name, L, Mpair, FormFac, radius, 5) {}
NonRes::NonRes(std::string name, Variable, Variable, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("NonRes", name, L, Mpair, FormFac, radius, 6) {
    GOOFIT_B200_LS_CONSTS
}
Flatte::Flatte(std::string name, Variable mass, Variable, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("Flatte", name, L, Mpair, FormFac, radius, 7) {
    registerParameter(mass);
    GOOFIT_B200_LS_CONSTS
}
GLASS::GLASS(std::string name, Variable mass, Variable width, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius,
             std::vector<Variable> AdditionalVars)
    : Lineshape("GLASS", name, L, Mpair, FormFac, radius, 8) {
    if(5 != AdditionalVars.size())
        throw GeneralError("It seems you forgot to provide the vector with the five necessary variables for GLASS, a, "
                           "r, phiF, phiR and F (in that order)");
    registerParameter(mass);
    registerParameter(width);
    for(auto &v : AdditionalVars) registerParameter(v);
    GOOFIT_B200_LS_CONSTS
}
FOCUS::FOCUS(std::string name, Mod mod, Variable, Variable, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("FOCUS", name, L, Mpair, FormFac, radius, 9) {
    GOOFIT_B200_LS_CONSTS
    registerConstant(static_cast<unsigned int>(mod));
}
static void registerKMatrix(Lineshape *self, std::vector<Variable> heads, std::vector<Variable> &beta_r, std::vector<Variable> &beta_i,
                           std::vector<Variable> &f_prod_r, std::vector<Variable> &f_prod_i, std::vector<Variable> &fscat,
                           std::vector<Variable> &poles, const std::function<void(Variable)> &reg) {
    if(fscat.size() != 5) throw GeneralError("You must have {} channels in fscat, not {}", 5, fscat.size());
    if(poles.size() != 5 * 6) throw GeneralError("You must have {}x{} channels in poles, not {}", 5, 6, poles.size());
    if(beta_r.size() < 5 || beta_i.size() < 5 || f_prod_r.size() < 5 || f_prod_i.size() < 5)
        throw GeneralError("kMatrix {}: beta and f_prod need five (real, imaginary) pairs", self->getName());
    for(auto &v : heads) reg(v);
    for(int i = 0; i < 5; i++) reg(fscat.at(i));
    for(int i = 0; i < 5; i++) {
        reg(beta_r.at(i));
        reg(beta_i.at(i));
    }
    for(int i = 0; i < 5; i++) {
        reg(f_prod_r.at(i));
        reg(f_prod_i.at(i));
    }
    for(int i = 0; i < 30; i++) reg(poles.at(i));
}
kMatrix::kMatrix(std::string name, Variable, Variable, Variable sA0, Variable sA, Variable s0_prod, Variable s0_scatt, std::vector<Variable> beta_r,
                 std::vector<Variable> beta_i, std::vector<Variable> f_prod_r, std::vector<Variable> f_prod_i, std::vector<Variable> fscat,
                 std::vector<Variable> poles, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius)
    : Lineshape("kMatrix", name, L, Mpair, FormFac, radius, 10) {
    registerKMatrix(this, {sA0, sA, s0_prod, s0_scatt}, beta_r, beta_i, f_prod_r, f_prod_i, fscat, poles, [this](Variable v) { registerParameter(v); });
    GOOFIT_B200_LS_CONSTS
}
// GSpline.cu:97-124 make_spline_curvatures, as written upstream: the (n-2) x (n-2) system has 4 on the diagonal, 1 on the
// super-diagonal and 1 in COLUMN 1 of every row below the first (`m(i + 1, 1) = 1`), it is inverted, applied to the second
// differences of the knot values, and scaled by 6 / spacing^2 with spacing = (first - second entry of SplineInfo) / bins.
// Upstream leaves the other matrix elements uninitialised (Eigen::Matrix is not zero-filled); they are zero here.
std::vector<fptype> GSpline::curvatures(const std::vector<Variable> &vars, const spline_t &info) {
    const size_t size = std::get<2>(info) - 2;
    std::vector<std::vector<fptype>> m(size, std::vector<fptype>(size, 0.0)), inv(size, std::vector<fptype>(size, 0.0));
    for(size_t i = 0; i < size; i++) {
        m[i][i] = 4;
        if(i != size - 1) {
            m[i][i + 1] = 1;
            m[i + 1][1] = 1;
        }
        inv[i][i] = 1;
    }
    for(size_t col = 0; col < size; col++) { // Gauss-Jordan with partial pivoting
        size_t piv = col;
        for(size_t r = col + 1; r < size; r++)
            if(std::fabs(m[r][col]) > std::fabs(m[piv][col])) piv = r;
        if(m[piv][col] == 0) throw GeneralError("GSpline: singular curvature system");
        std::swap(m[piv], m[col]);
        std::swap(inv[piv], inv[col]);
        const fptype d = m[col][col];
        for(size_t k = 0; k < size; k++) {
            m[col][k] /= d;
            inv[col][k] /= d;
        }
        for(size_t r = 0; r < size; r++) {
            if(r == col || m[r][col] == 0) continue;
            const fptype f = m[r][col];
            for(size_t k = 0; k < size; k++) {
                m[r][k] -= f * m[col][k];
                inv[r][k] -= f * inv[col][k];
            }
        }
    }
    std::vector<fptype> Lv(size);
    for(size_t i = 0; i < size; ++i) Lv[i] = vars[i + 2].getValue() - 2 * vars[i + 1].getValue() + vars[i].getValue();
    const fptype spacing = (std::get<0>(info) - std::get<1>(info)) / std::get<2>(info);
    std::vector<fptype> ret(vars.size(), 0);
    for(size_t i = 0; i < size; ++i) {
        fptype mtv = 0;
        for(size_t k = 0; k < size; k++) mtv += inv[i][k] * Lv[k];
        ret.at(i + 1) = 6 * mtv / (spacing * spacing);
    }
    return ret;
}
GSpline::GSpline(std::string name, Variable mass, Variable width, unsigned int L, unsigned int Mpair, FF FormFac, fptype radius,
                 std::vector<Variable> AdditionalVars, spline_t SplineInfo)
    : Lineshape("GSpline", name, L, Mpair, FormFac, radius, 11) {
    registerParameter(mass);
    registerParameter(width);
    GOOFIT_B200_LS_CONSTS
    if(std::get<2>(SplineInfo) != AdditionalVars.size()) throw GeneralError("bins {} != vars {}", std::get<2>(SplineInfo), AdditionalVars.size());
    if(AdditionalVars.size() < 3) throw GeneralError("GSpline {} needs at least three knots", name);
    registerConstant(std::get<0>(SplineInfo));
    registerConstant(std::get<1>(SplineInfo));
    registerConstant(std::get<2>(SplineInfo));
    std::vector<fptype> terms = curvatures(AdditionalVars, SplineInfo);
    registerConstant(static_cast<fptype>(terms.size()));
    for(fptype t : terms) registerConstant(t);
}
#undef GOOFIT_B200_LS_CONSTS
} // namespace Lineshapes

// ---- spin factors / amplitudes -------------------------------------------------------------------------------
SpinFactor::SpinFactor(std::string name, SF_4Body SF, fptype mD0, unsigned int P0, unsigned int P1, unsigned int P2, unsigned int P3)
    : AmpComponent("SpinFactor", name)
    , _SF(SF)
    , _P0(P0)
    , _P1(P1)
    , _P2(P2)
    , _P3(P3) {
    registerConstant(mD0);
    registerConstant(P0);
    registerConstant(P1);
    registerConstant(P2);
    registerConstant(P3);
    registerFunction("ptr_to_SpinFactor", GF_OP_SF);
}

Amplitude::Amplitude(std::string uniqueDecayStr, Variable ar, Variable ai, std::vector<Lineshape *> LS, std::vector<SpinFactor *> SF,
                     unsigned int nPerm)
    : AmpComponent("Amplitude", uniqueDecayStr)
    , _uniqueDecayStr(uniqueDecayStr)
    , _ar(ar)
    , _ai(ai)
    , _SF(std::move(SF))
    , _LS(std::move(LS))
    , _nPerm(nPerm) {
    if(_nPerm < 1) throw GeneralError("Amplitude {}: nPerm must be at least 1", uniqueDecayStr);
    for(auto *l : _LS) components.push_back(l);
    for(auto *s : _SF) components.push_back(s);
}
bool Amplitude::operator==(const Amplitude &A) const {
    return _uniqueDecayStr == A._uniqueDecayStr && _ar == A._ar && _ai == A._ai && _LS == A._LS && _SF == A._SF && _nPerm == A._nPerm;
}
bool Amplitude::lineshapeParametersChanged() const {
    return std::any_of(_LS.cbegin(), _LS.cend(), [](const Lineshape *p) { return p->parametersChanged(); });
}

// ---- phase space -----------------------------------------------------------------------------------------------
namespace {
struct V4 { // (E, px, py, pz)
    double e, x, y, z;
};
inline V4 operator+(V4 a, V4 b) { return {a.e + b.e, a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V4 operator-(V4 a, V4 b) { return {a.e - b.e, a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V4 operator*(double c, V4 a) { return {a.e * c, a.x * c, a.y * c, a.z * c}; }
inline V4 operator/(V4 a, double c) {
    double ci = 1.0 / c;
    return {a.e * ci, a.x * ci, a.y * ci, a.z * ci};
}
inline double mink(V4 a, V4 b) { return a.e * b.e - a.x * b.x - a.y * b.y - a.z * b.z; }
inline double mass2(V4 a) { return mink(a, a); }
inline double mass(V4 a) {
    double m2 = mass2(a);
    return m2 > 0 ? std::sqrt(m2) : 0.0;
}
inline double dot3(V4 a, V4 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline double mag3(V4 a) { return std::sqrt(dot3(a, a)); }
inline V4 cross3(V4 a, V4 b) { return {0.0, a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// helicity angle (detail/Dim5.h:22-34)
double cosHel(V4 p, V4 q, V4 d) {
    double pd = mink(p, d), pq = mink(p, q), qd = mink(q, d);
    double mp2 = mass2(p), mq2 = mass2(q), md2 = mass2(d);
    return (pd * mq2 - pq * qd) / std::sqrt((pq * pq - mq2 * mp2) * (qd * qd - mq2 * md2));
}
// angle between the two decay planes (detail/Dim5.h:36-92; only the second estimate is returned upstream)
double planeAngle(V4 d1, V4 d2, V4 h1, V4 h2) {
    V4 d1n = d1 / mag3(d1), d2n = d2 / mag3(d2), h1n = h1 / mag3(h1), h2n = h2 / mag3(h2);
    V4 h12 = h1 + h2;
    h12 = (1.0 / mag3(h12)) * h12;
    V4 n1 = cross3(d1n, d2n), n2 = cross3(h1n, h2n);
    n1 = (1.0 / mag3(n1)) * n1;
    n2 = (1.0 / mag3(n2)) * n2;
    V4 n3 = cross3(n1, n2);
    double cp = dot3(n1, n2), sp = dot3(n3, h12);
    double phi2 = std::acos(cp);
    if(sp < 0) phi2 *= -1;
    return phi2;
}
inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline double uniform01(uint64_t seed, uint64_t event, uint64_t k) {
    uint64_t h = mix64(mix64(seed ^ 0xD1B54A32D192ED03ull) + event * 0x2545F4914F6CDD1Dull + k);
    return ((h >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}
} // namespace
namespace detail {
double counterUniform(unsigned long seed, size_t event, unsigned long stream) { return uniform01(seed, event, stream); }
} // namespace detail
namespace {
inline double pdk(double a, double b, double c) {
    double x = (a - b - c) * (a + b + c) * (a + b - c) * (a - b + c);
    return std::sqrt(x > 0 ? x : 0) / (2 * a);
}
void boost(V4 &p, double bx, double by, double bz) {
    double b2 = bx * bx + by * by + bz * bz;
    if(b2 <= 0) return;
    double gamma = 1.0 / std::sqrt(1.0 - b2), bp = bx * p.x + by * p.y + bz * p.z, g2 = (gamma - 1.0) / b2;
    p.x += g2 * bp * bx + gamma * bx * p.e;
    p.y += g2 * bp * by + gamma * by * p.e;
    p.z += g2 * bp * bz + gamma * bz * p.e;
    p.e = gamma * (p.e + bp);
}
} // namespace

void fourVectorsToDim5(const fptype *p, fptype *out) { // detail/Dim5.h:94-110
    V4 a{p[0], p[1], p[2], p[3]}, b{p[4], p[5], p[6], p[7]}, c{p[8], p[9], p[10], p[11]}, d{p[12], p[13], p[14], p[15]};
    V4 pM = a + b + c + d, p12 = a + b, p34 = c + d;
    out[0] = mass(p12);
    out[1] = mass(p34);
    out[2] = cosHel(pM, p12, a);
    out[3] = cosHel(pM, p34, c);
    out[4] = planeAngle(a, b, c, d);
}

std::vector<std::vector<fptype>> generatePhaseSpace5(const std::vector<fptype> &pm, size_t n, unsigned long seed, std::vector<fptype> *weights) {
    if(pm.size() != 5) throw GeneralError("generatePhaseSpace5 needs the mother and four daughter masses");
    const double M = pm[0], m[4] = {pm[1], pm[2], pm[3], pm[4]};
    double tecm = M - m[0] - m[1] - m[2] - m[3];
    if(tecm <= 0) throw GeneralError("generatePhaseSpace5: not enough energy");
    double emmax = tecm + m[0], emmin = 0, wtmax = 1;
    for(int k = 1; k < 4; k++) {
        emmin += m[k - 1];
        emmax += m[k];
        wtmax *= pdk(emmax, emmin, m[k]);
    }
    std::vector<std::vector<fptype>> cols(5, std::vector<fptype>(n));
    if(weights) weights->assign(n, 0.0);
    for(size_t ev = 0; ev < n; ev++) { // GENBOD: sorted uniforms -> intermediate masses -> sequential two-body decays
        uint64_t ctr = 0;
        double r1 = uniform01(seed, ev, ctr++), r2 = uniform01(seed, ev, ctr++);
        if(r1 > r2) std::swap(r1, r2);
        double rno[4] = {0, r1, r2, 1}, inv[4], pd[3], sum = 0;
        for(int k = 0; k < 4; k++) {
            sum += m[k];
            inv[k] = rno[k] * tecm + sum;
        }
        double wt = 1.0 / wtmax;
        for(int k = 0; k < 3; k++) {
            pd[k] = pdk(inv[k + 1], inv[k], m[k + 1]);
            wt *= pd[k];
        }
        V4 p[4];
        p[0] = {std::sqrt(pd[0] * pd[0] + m[0] * m[0]), 0, pd[0], 0};
        for(int i = 1;; i++) {
            p[i] = {std::sqrt(pd[i - 1] * pd[i - 1] + m[i] * m[i]), 0, -pd[i - 1], 0};
            double cZ = 2 * uniform01(seed, ev, ctr++) - 1, sZ = std::sqrt(1 - cZ * cZ);
            double ang = 2 * M_PI * uniform01(seed, ev, ctr++), cY = std::cos(ang), sY = std::sin(ang);
            for(int j = 0; j <= i; j++) {
                double x = p[j].x, y = p[j].y;
                p[j].x = cZ * x - sZ * y;
                p[j].y = sZ * x + cZ * y;
                x = p[j].x;
                double z = p[j].z;
                p[j].x = cY * x - sY * z;
                p[j].z = sY * x + cY * z;
            }
            if(i == 3) break;
            double beta = pd[i] / std::sqrt(pd[i] * pd[i] + inv[i] * inv[i]);
            for(int j = 0; j <= i; j++) boost(p[j], 0, beta, 0);
        }
        double in[16], out[5];
        for(int k = 0; k < 4; k++) {
            in[4 * k] = p[k].e; in[4 * k + 1] = p[k].x; in[4 * k + 2] = p[k].y; in[4 * k + 3] = p[k].z;
        }
        fourVectorsToDim5(in, out);
        for(int k = 0; k < 5; k++) cols[k][ev] = out[k];
        if(weights) (*weights)[ev] = wt;
    }
    return cols;
}

void generatePhaseSpaceParticles(const std::vector<fptype> &pm, size_t n, unsigned long seed, size_t offset,
                                 std::vector<std::vector<fptype>> &particles, std::vector<fptype> &weights) {
    const int nd = static_cast<int>(pm.size()) - 1;
    if(nd != 3 && nd != 4) throw GeneralError("generatePhaseSpaceParticles: 3 or 4 daughters, got {}", nd);
    const double M = pm[0];
    const double *m = pm.data() + 1;
    double tecm = M;
    for(int k = 0; k < nd; k++) tecm -= m[k];
    if(tecm <= 0) throw GeneralError("generatePhaseSpaceParticles: not enough energy");
    double emmax = tecm + m[0], emmin = 0, wtmax = 1;
    for(int k = 1; k < nd; k++) {
        emmin += m[k - 1];
        emmax += m[k];
        wtmax *= pdk(emmax, emmin, m[k]);
    }
    particles.assign(nd, std::vector<fptype>(4 * n));
    weights.assign(n, 0.0);
    for(size_t i = 0; i < n; i++) {
        const uint64_t ev = i + offset;
        uint64_t ctr = 0;
        double rno[4] = {0, 0, 0, 0}, inv[4], pd[3], sum = 0;
        for(int k = 1; k < nd - 1; k++) rno[k] = uniform01(seed, ev, ctr++);
        if(nd == 4 && rno[1] > rno[2]) std::swap(rno[1], rno[2]);
        rno[nd - 1] = 1;
        for(int k = 0; k < nd; k++) {
            sum += m[k];
            inv[k] = rno[k] * tecm + sum;
        }
        double wt = 1.0 / wtmax;
        for(int k = 0; k < nd - 1; k++) {
            pd[k] = pdk(inv[k + 1], inv[k], m[k + 1]);
            wt *= pd[k];
        }
        V4 p[4];
        p[0] = {std::sqrt(pd[0] * pd[0] + m[0] * m[0]), 0, pd[0], 0};
        for(int d = 1;; d++) { // sequential two-body decays, each rotated at random and boosted into its parent's frame
            p[d] = {std::sqrt(pd[d - 1] * pd[d - 1] + m[d] * m[d]), 0, -pd[d - 1], 0};
            double cZ = 2 * uniform01(seed, ev, ctr++) - 1, sZ = std::sqrt(1 - cZ * cZ);
            double ang = 2 * M_PI * uniform01(seed, ev, ctr++), cY = std::cos(ang), sY = std::sin(ang);
            for(int j = 0; j <= d; j++) {
                double x = p[j].x, y = p[j].y;
                p[j].x = cZ * x - sZ * y;
                p[j].y = sZ * x + cZ * y;
                x = p[j].x;
                double z = p[j].z;
                p[j].x = cY * x - sY * z;
                p[j].z = sY * x + cY * z;
            }
            if(d == nd - 1) break;
            double beta = pd[d] / std::sqrt(pd[d] * pd[d] + inv[d] * inv[d]);
            for(int j = 0; j <= d; j++) boost(p[j], 0, beta, 0);
        }
        for(int d = 0; d < nd; d++) {
            fptype *o = particles[d].data() + 4 * i;
            o[0] = p[d].e; o[1] = p[d].x; o[2] = p[d].y; o[3] = p[d].z;
        }
        weights[i] = wt;
    }
}
void devicePhaseSpace(const std::vector<fptype> &pm, size_t n, unsigned long seed, size_t offset, std::vector<std::vector<fptype>> *particles,
                      std::vector<fptype> &weights, std::vector<std::vector<fptype>> *dim5) {
    const int nd = static_cast<int>(pm.size()) - 1;
    if(nd != 3 && nd != 4) throw GeneralError("devicePhaseSpace: 3 or 4 daughters, got {}", nd);
    if(dim5 && nd != 4) throw GeneralError("devicePhaseSpace: the five-variable form needs four daughters");
    weights.assign(n, 0.0);
    std::vector<fptype> flatP(particles ? 4 * n * nd : 0), flatV(dim5 ? 5 * n : 0);
    Engine &E = Engine::get();
    Engine::check(gf_genbod(E.ctx(), pm.data(), nd, n, seed, offset, particles ? flatP.data() : nullptr, weights.data(), dim5 ? flatV.data() : nullptr),
                  "gf_genbod");
    if(particles) {
        particles->assign(nd, std::vector<fptype>());
        for(int d = 0; d < nd; d++) (*particles)[d].assign(flatP.begin() + (size_t)d * 4 * n, flatP.begin() + (size_t)(d + 1) * 4 * n);
    }
    if(dim5) {
        dim5->assign(5, std::vector<fptype>());
        for(int k = 0; k < 5; k++) (*dim5)[k].assign(flatV.begin() + (size_t)k * n, flatV.begin() + (size_t)(k + 1) * n);
    }
}
std::vector<bool> acceptRejectFlags(const std::vector<fptype> &w, unsigned long seed, size_t offset) {
    fptype wmax = 0;
    for(fptype v : w) wmax = std::max(wmax, v);
    std::vector<bool> flags(w.size());
    for(size_t i = 0; i < w.size(); i++) flags[i] = wmax > 0 && w[i] / wmax > uniform01(seed, i + offset, 2000);
    return flags;
}

std::vector<bool> unweightPhaseSpace(const std::vector<fptype> &w, unsigned long seed) {
    std::vector<bool> flags(w.size());
    for(size_t i = 0; i < w.size(); i++) flags[i] = w[i] > uniform01(seed, i, 1000);
    return flags;
}

GeneratedSignal Amp4Body::GenerateSigColumns(unsigned int numEvents, int seed) {
    GeneratedSignal G;
    devicePhaseSpace(decayInfo.particle_masses, numEvents, static_cast<unsigned long>(seed), generation_offset, &G.particles, G.weights, &G.variables);
    std::vector<fptype> evt(numEvents);
    std::iota(evt.begin(), evt.end(), 0.0);
    std::vector<Observable> obs(observablesList.begin(), observablesList.begin() + 6);
    generated = std::make_unique<UnbinnedDataSet>(obs);
    generated->appendColumns({G.variables[0].data(), G.variables[1].data(), G.variables[2].data(), G.variables[3].data(),
                              G.variables[4].data(), evt.data()},
                             numEvents);
    setData(generated.get());
    setDataSize(numEvents, 6);
    auto fc = fitControl;
    setFitControl(std::make_shared<EvalFit>()); // the un-normalised value, upstream's ProbFit with the normalisation forced to 1
    normalize();
    std::vector<fptype> values = evaluate_with_metric();
    setFitControl(fc);
    for(size_t i = 0; i < numEvents; i++) G.weights[i] *= values[i];
    G.flags = acceptRejectFlags(G.weights, static_cast<unsigned long>(seed), generation_offset);
    return G;
}

// ---- what Amp4Body and Amp4Body_TD share ------------------------------------------------------------------------
void Amp4BodyBase::addAmplitudeRecord(Amplitude *amp) { // Amp4Body.cu:120-160 / Amp4Body_TD.cu:366-420: unique by VALUE
    if(amp->_LS.size() % amp->_nPerm || amp->_SF.size() % amp->_nPerm)
        throw GeneralError("Amplitude {}: lineshapes and spin factors must split evenly over the permutations", amp->getName());
    ampStarts.push_back(static_cast<int32_t>(ampRecords.size()));
    ampRecords.push_back(static_cast<int32_t>(amp->_LS.size()));
    ampRecords.push_back(static_cast<int32_t>(amp->_SF.size()));
    ampRecords.push_back(static_cast<int32_t>(amp->_nPerm));
    for(auto *ls : amp->_LS) {
        auto it = std::find_if(LineShapes.begin(), LineShapes.end(), [ls](const Lineshape *L) { return *ls == *L; });
        if(it == LineShapes.end()) {
            LineShapes.push_back(ls);
            it = LineShapes.end() - 1;
        }
        ampRecords.push_back(static_cast<int32_t>(it - LineShapes.begin()));
    }
    for(auto *sf : amp->_SF) {
        auto it = std::find_if(SpinFactors.begin(), SpinFactors.end(), [sf](const SpinFactor *S) { return *sf == *S; });
        if(it == SpinFactors.end()) {
            SpinFactors.push_back(sf);
            it = SpinFactors.end() - 1;
        }
        ampRecords.push_back(static_cast<int32_t>(it - SpinFactors.begin()));
    }
    ++nAmps;
    if(LineShapes.size() > 64 || SpinFactors.size() > 64 || nAmps > 64)
        throw GeneralError("{} supports at most 64 unique lineshapes, spin factors and amplitudes", getName());
}
std::vector<int32_t> Amp4BodyBase::headerInts(detail::ProgramBuilder &B, int tdOffset) const {
    std::vector<int32_t> ints{B.allocAux(), (int32_t)LineShapes.size(), (int32_t)SpinFactors.size(), (int32_t)nAmps, tdOffset};
    for(auto *ls : LineShapes) ints.push_back(ls->getFunctionIndex());
    for(auto *sf : SpinFactors) ints.push_back(sf->getFunctionIndex());
    ints.insert(ints.end(), ampStarts.begin(), ampStarts.end());
    ints.insert(ints.end(), ampRecords.begin(), ampRecords.end());
    return ints;
}
// The generated normalisation sample: per seed, GENBOD candidates and their five variables on the GPU (k_genbod), the
// accept-reject "Unweight" step against the counter-based uniform stream, then either the COMPACTED accepted events
// (Amp4Body_TD: detail/NormEvents_4Body_Base.cu:24-78 does run the remove_if) or the FIRST nAcc events of the un-compacted sample
// (Amp4Body.cu:243-289: the remove_if is commented out there). Generation is deferred to the first use so that building a model
// needs no device and a 10^8-event sample costs no host loop over GENBOD.
void Amp4BodyBase::materialiseNormEvents() {
    if(normSpec.perBatch == 0) return;
    NormSpec spec = normSpec;
    normSpec.perBatch = 0;
    std::vector<std::vector<fptype>> all(5);
    std::vector<fptype> firstCandidate;
    for(long seed : spec.seeds) {
        std::vector<fptype> w;
        std::vector<std::vector<fptype>> v5;
        devicePhaseSpace(spec.masses, spec.perBatch, static_cast<unsigned long>(seed), 0, nullptr, w, &v5);
        auto acc = unweightPhaseSpace(w, static_cast<unsigned long>(seed));
        if(firstCandidate.empty() && !w.empty())
            for(int k = 0; k < 5; k++) firstCandidate.push_back(v5[k][0]);
        if(spec.compact) {
            for(size_t i = 0; i < w.size(); i++)
                if(acc[i])
                    for(int k = 0; k < 5; k++) all[k].push_back(v5[k][i]);
        } else {
            size_t nAcc = std::count(acc.begin(), acc.end(), true);
            if(nAcc == 0) nAcc = 1;
            for(int k = 0; k < 5; k++) all[k].insert(all[k].end(), v5[k].begin(), v5[k].begin() + nAcc);
        }
    }
    // Nothing accepted (python/examples/generate_sig_example.py asks for ONE candidate because it only generates): upstream is left
    // with an empty sample, which its GenerateSig never touches and whose integral would be 0 / 0. The caches here need a sample
    // to be allocated against, so the first candidate stands in, and normalize() refuses to integrate over it.
    const bool none = all[0].empty();
    if(none) {
        if(firstCandidate.empty()) throw GeneralError("{}: no normalisation event was generated", getName());
        for(int k = 0; k < 5; k++) all[k].push_back(firstCandidate[k]);
    }
    setNormEvents(all);
    normSampleEmpty = none;
}
void Amp4BodyBase::setNormEvents(const std::vector<std::vector<fptype>> &cols5) {
    normSpec.perBatch = 0; // an explicit sample replaces a pending generated one
    normSampleEmpty   = false;
    normWeights.clear();
    normDtime.clear();
    normSigma.clear();
    if(cols5.size() != 5) throw GeneralError("setNormEvents needs five columns (m12, m34, cos12, cos34, phi)");
    for(auto &c : cols5)
        if(c.size() != cols5[0].size() || c.empty()) throw GeneralError("setNormEvents: columns must be non-empty and of equal length");
    normCols       = cols5;
    MCevents       = cols5[0].size();
    cacheAllocated = false;
    forceRedoIntegrals = true;
}
void Amp4BodyBase::emitAux(detail::ProgramBuilder &B) {
    for(auto *ls : LineShapes) ls->emit(B);
    for(auto *sf : SpinFactors) sf->emit(B);
    efficiency->emitAux(B);
}
void Amp4BodyBase::afterProgram() {
    cacheAllocated     = false;
    forceRedoIntegrals = true;
}
void Amp4BodyBase::setDataSizeCommon(unsigned int dataSize, unsigned int evtSize) { // Amp4Body.cu:348-395, Amp4Body_TD.cu:575-604
    totalEventSize = evtSize;
    if(totalEventSize < 3) throw GeneralError("totalEventSize {} should be 3 or more", totalEventSize);
    numEntries     = dataSize;
    cacheAllocated = false;
    setForceIntegrals();
}
void Amp4BodyBase::ensureCache() {
    if(cacheAllocated) return;
    if(nodeIdx < 0) throw GeneralError("{}: call setData() on the top-level PDF before evaluating", getName());
    materialiseNormEvents();
    if(MCevents == 0) throw GeneralError("{} has no normalisation events", getName());
    Engine &E = Engine::get();
    E.requireComm("Amp4Body normalisation");
    Engine::check(gf_amp4_alloc(E.ctx(), nodeIdx, E.localEvents), "gf_amp4_alloc");
    // the MC normalisation sample is sharded over ranks like the data (the reference replicates it)
    detail::Shard sh = E.shard();
    size_t per = MCevents / sh.size, first = per * sh.rank, count = sh.rank == sh.size - 1 ? MCevents - first : per;
    const double *cols[5];
    for(int k = 0; k < 5; k++) cols[k] = normCols[k].data() + first;
    const bool wtd = !normWeights.empty();
    Engine::check(gf_amp4_set_mc(E.ctx(), nodeIdx, cols, count, wtd ? normWeights.data() + first : nullptr, wtd ? normDtime.data() + first : nullptr,
                                 wtd ? normSigma.data() + first : nullptr),
                  "gf_amp4_set_mc");
    cacheAllocated     = true;
    forceRedoIntegrals = true;
    gramValid          = false;
}
// Rebuild what a moved mass or width invalidated. In Gram mode the lineshapes, the amplitudes that contain them and G_ab are
// refreshed in one pass per sample (gf_amp4_cache_gram); in PerEvent mode only the caches are (the MC integral follows per call).
void Amp4BodyBase::refreshCaches(bool any, int spins, const std::vector<uint8_t> &lsRedo, const std::vector<uint8_t> &ampRedo) {
    if(!any) return;
    Engine &E = Engine::get();
    if(normMode == NormMode::Gram) {
        G.assign(nAmps * nAmps, {0.0, 0.0});
        Engine::check(gf_amp4_cache_gram(E.ctx(), nodeIdx, spins, lsRedo.data(), ampRedo.data(), reinterpret_cast<double *>(G.data())),
                      "gf_amp4_cache_gram");
        gramValid = true;
    } else
        Engine::check(gf_amp4_cache(E.ctx(), nodeIdx, spins, lsRedo.data(), ampRedo.data()), "gf_amp4_cache");
}
const std::vector<std::complex<fptype>> &Amp4BodyBase::gram(bool mcAmpsChanged) {
    if(mcAmpsChanged || !gramValid || G.size() != nAmps * nAmps) {
        G.assign(nAmps * nAmps, {0.0, 0.0});
        Engine &E = Engine::get();
        Engine::check(gf_amp4_gram(E.ctx(), nodeIdx, reinterpret_cast<double *>(G.data())), "gf_amp4_gram");
        gramValid = true;
    }
    return G;
}
std::vector<std::complex<fptype>> Amp4BodyBase::getCachedAmp(size_t a) const {
    if(!cacheAllocated) throw GeneralError("{}: no cached amplitudes yet", getName());
    Engine &E = Engine::get();
    std::vector<std::complex<fptype>> ret(E.localEvents);
    Engine::check(gf_amp4_get_amp(E.ctx(), nodeIdx, (int)a, reinterpret_cast<double *>(ret.data()), ret.size()), "gf_amp4_get_amp");
    return ret;
}

// ---- Amp4Body ------------------------------------------------------------------------------------------------------
Amp4Body::Amp4Body(std::string n, std::vector<Observable> observables, DecayInfo4 decay, GooPdf *eff, unsigned int MCeventsNorm)
    : Amp4BodyBase("Amp4Body", n)
    , decayInfo(decay) {
    efficiency     = eff;
    totalEventSize = static_cast<int>(observables.size());
    if(observables.size() < 6) throw GeneralError("Amp4Body {} needs m12, m34, cos12, cos34, phi and an event number", n);
    if(decayInfo.particle_masses.size() != 5) throw GeneralError("Amp4Body {}: particle_masses must hold the mother and four daughters", n);
    if(decayInfo.amplitudes.empty()) throw GeneralError("Amp4Body {} needs at least one amplitude", n);
    if(!efficiency) throw GeneralError("Amp4Body {} needs an efficiency function", n);
    for(auto &o : observables) registerObservable(o);
    for(double pm : decayInfo.particle_masses) registerConstant(pm);
    for(auto *amp : decayInfo.amplitudes) {
        components.push_back(amp);
        registerParameter(amp->_ar);
        registerParameter(amp->_ai);
        addAmplitudeRecord(amp);
    }
    components.push_back(efficiency);
    registerFunction("ptr_to_DP", GF_OP_AMP4);
    // Normalisation sample. The reference generates MCeventsNorm phase-space events, runs the accept-reject
    // "Unweight" step and then keeps the FIRST nAcc events of the un-compacted sample (Amp4Body.cu:243-289: the
    // remove_if is commented out); the same selection is made here.
    // The sample itself is generated on the device when it is first needed (materialiseNormEvents).
    if(MCeventsNorm > 0) normSpec = NormSpec{{0}, MCeventsNorm, false, decayInfo.particle_masses};
    setSeparateNorm();
}
int Amp4Body::emit(detail::ProgramBuilder &B) {
    int first = (int)B.nodes.size();
    efficiency->emit(B);
    return emitRecord(B, GF_OP_AMP4, 1, (int)B.nodes.size() - first + 1, headerInts(B, 0));
}
fptype Amp4Body::normalize() { // Amp4Body.cu:398-558
    recursiveSetNormalization(1.0);
    ensureCache();
    Engine &E   = Engine::get();
    size_t nLS  = LineShapes.size(), nAmp = nAmps;
    std::vector<uint8_t> lsRedo(nLS), ampRedo(nAmp, 0);
    bool any = forceRedoIntegrals;
    for(size_t i = 0; i < nLS; ++i) {
        lsRedo[i] = forceRedoIntegrals || LineShapes[i]->parametersChanged();
        any       = any || lsRedo[i];
    }
    // Upstream flags the cached data amplitude a for recomputation with redoIntegral[a], an array that is filled per
    // LINESHAPE index (Amp4Body.cu:407-414, 490-492); kept as is, so amplitude a follows lineshape a.
    // Entries beyond the number of lineshapes keep their constructor value `true` upstream (Amp4Body.cu:196-199).
    for(size_t a = 0; a < nAmp; ++a) {
        ampRedo[a] = a < nLS ? lsRedo[a] : 1;
        any        = any || ampRedo[a];
    }
    int spins          = forceRedoIntegrals ? 1 : 0;
    forceRedoIntegrals = false;
    E.sync();
    refreshCaches(any, spins, lsRedo, ampRedo);
    double sum = 0;
    if(normMode == NormMode::Gram) {
        // sum_m |sum_a c_a A_a(m)|^2 = sum_a |c_a|^2 G_aa + 2 sum_{a<b} Re(c_a conj(c_b) G_ab), with the coefficients of the
        // parameter block the device would have read
        const auto &g = gram(false); // refreshCaches left G current
        for(size_t a = 0; a < nAmp; ++a) {
            std::complex<fptype> ca(par(2 * a), par(2 * a + 1));
            sum += std::norm(ca) * g[a * nAmp + a].real();
            for(size_t b = a + 1; b < nAmp; ++b) {
                std::complex<fptype> cb(par(2 * b), par(2 * b + 1));
                sum += 2 * (ca * std::conj(cb) * g[a * nAmp + b]).real();
            }
        }
    } else
        Engine::check(gf_amp4_norm(E.ctx(), nodeIdx, &sum), "gf_amp4_norm");
    fptype ret = sum / static_cast<fptype>(MCevents);
    if(std::isnan(ret)) GooFit::abort(__FILE__, __LINE__, getName() + " NAN normalization in Amp4Body", this);
    setNormalization(1.0 / ret);
    return ret;
}

// ---- decay-time resolutions ------------------------------------------------------------------------------------------
TruthResolution::TruthResolution()
    : MixingTimeResolution("TruthResolution", "mixing_resolution") {
    registerFunction("ptr_to_truth", GF_OP_TRES_TRUTH);
}
fptype TruthResolution::normalization(fptype di1, fptype di2, fptype di3, fptype di4, fptype tau, fptype xmixing, fptype ymixing) const {
    fptype timeIntegralOne = tau / (1 - ymixing * ymixing); // TruthResolution.cu:56-70
    fptype timeIntegralTwo = tau / (1 + xmixing * xmixing);
    fptype timeIntegralThr = ymixing * timeIntegralOne;
    fptype timeIntegralFou = xmixing * timeIntegralTwo;
    fptype ret             = timeIntegralOne * (di1 + di2);
    ret += timeIntegralTwo * (di1 - di2);
    ret -= 2 * timeIntegralThr * di3;
    ret -= 2 * timeIntegralFou * di4;
    return ret;
}
ThreeGaussResolution::ThreeGaussResolution(Variable cf, Variable tf, Variable cb, Variable cs, Variable tb, Variable ts, Variable ob,
                                           Variable os, Variable sb, Variable common_bias_offset, Variable common_width_scale,
                                           Variable common_bias_offset2, Variable common_width_scale2)
    : MixingTimeResolution("ThreeGaussResolution", "mixing_resolution", cf, tf, cb, cs, tb, ts, ob, os, sb)
    , selectionBias(sb) {
    registerParameter(common_bias_offset);
    registerParameter(common_width_scale);
    registerParameter(common_bias_offset2);
    registerParameter(common_width_scale2);
    registerFunction("ptr_to_threegauss", GF_OP_TRES_3GAUSS);
}
fptype ThreeGaussResolution::normalization(fptype di1, fptype di2, fptype di3, fptype di4, fptype tau, fptype xmixing,
                                           fptype ymixing) const { // ThreeGaussResolution.cu:155-182
    fptype selBias = selectionBias.getValue();
    fptype timeIntegralOne = (selBias + 1 / tau) / (selBias * selBias + 2 * selBias / tau + (1 - ymixing * ymixing) / (tau * tau));
    fptype timeIntegralTwo = (selBias + 1 / tau) / (selBias * selBias + 2 * selBias / tau + (1 + xmixing * xmixing) / (tau * tau));
    fptype timeIntegralThr = (ymixing / tau) / (selBias * selBias + 2 * selBias / tau + (1 - ymixing * ymixing) / (tau * tau));
    fptype timeIntegralFou = (xmixing / tau) / (selBias * selBias + 2 * selBias / tau + (1 + xmixing * xmixing) / (tau * tau));
    fptype ret             = timeIntegralOne * (di1 + di2);
    ret += timeIntegralTwo * (di1 - di2);
    ret -= 2 * timeIntegralThr * di3;
    ret -= 2 * timeIntegralFou * di4;
    return ret;
}

// ---- Amp4Body_TD -------------------------------------------------------------------------------------------------
Amp4Body_TD::Amp4Body_TD(std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *Tres,
                         GooPdf *eff, Observable *mistag, long normSeed, unsigned int numNormEventsToGen)
    : Amp4BodyBase("Amp4Body_TD", n)
    , decayInfo(decay)
    , resolution(Tres) {
    efficiency = eff;
    init(observables, mistag, {normSeed}, numNormEventsToGen);
}
Amp4Body_TD::Amp4Body_TD(std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *Tres,
                         GooPdf *eff, Observable *mistag, const std::vector<long> &normSeeds, unsigned int perBatch)
    : Amp4BodyBase("Amp4Body_TD", n)
    , decayInfo(decay)
    , resolution(Tres) {
    efficiency = eff;
    init(observables, mistag, normSeeds, perBatch);
}
// ---- normalisation-event containers ----------------------------------------------------------------------------------------
static std::vector<NormEvents_4Body_Base *> pendingBatches(const std::vector<long> &seeds, unsigned int perBatch, const std::vector<fptype> &masses,
                                                            bool device) {
    std::vector<NormEvents_4Body_Base *> ret;
    for(long seed : seeds) {
        NormEvents_4Body_Base *b = device ? static_cast<NormEvents_4Body_Base *>(new NormEvents_4Body_DeviceCached())
                                          : static_cast<NormEvents_4Body_Base *>(new NormEvents_4Body_HostCached());
        b->seed       = seed;
        b->toGenerate = perBatch;
        b->masses     = masses;
        ret.push_back(b);
    }
    return ret;
}
std::vector<NormEvents_4Body_Base *> NormEvents_4Body_DeviceCached::buildBatches(const std::vector<long> &seeds, unsigned int perBatch,
                                                                                 const std::vector<fptype> &masses) {
    return pendingBatches(seeds, perBatch, masses, true);
}
std::vector<NormEvents_4Body_Base *> NormEvents_4Body_HostCached::buildBatches(const std::vector<long> &seeds, unsigned int perBatch,
                                                                               const std::vector<fptype> &masses) {
    return pendingBatches(seeds, perBatch, masses, false);
}
NormEvents_4Body_Weighted::NormEvents_4Body_Weighted(std::vector<fptype> m12, std::vector<fptype> m34, std::vector<fptype> cos12,
                                                     std::vector<fptype> cos34, std::vector<fptype> phi, std::vector<fptype> t,
                                                     std::vector<fptype> s, std::vector<fptype> w) {
    size_t n = w.size();
    if(n == 0 || m12.size() != n || m34.size() != n || cos12.size() != n || cos34.size() != n || phi.size() != n || t.size() != n || s.size() != n)
        throw GeneralError("NormEvents_4Body_Weighted: eight non-empty columns of equal length are needed");
    cols    = {std::move(m12), std::move(m34), std::move(cos12), std::move(cos34), std::move(phi)};
    dtime   = std::move(t);
    sigma   = std::move(s);
    weights = std::move(w);
}
Amp4Body_TD::Amp4Body_TD(std::string n, std::vector<Observable> observables, DecayInfo4t decay, MixingTimeResolution *Tres, GooPdf *eff,
                         Observable *mistag, const std::vector<NormEvents_4Body_Base *> &normEvents)
    : Amp4BodyBase("Amp4Body_TD", n)
    , decayInfo(decay)
    , resolution(Tres) {
    efficiency = eff;
    std::vector<std::unique_ptr<NormEvents_4Body_Base>> own;
    for(auto *b : normEvents) own.emplace_back(b);
    if(own.empty()) throw GeneralError("Amp4Body_TD {}: no normalisation events", n);
    bool anyW = false, allW = true;
    for(auto &b : own) {
        if(!b) throw GeneralError("Amp4Body_TD {}: null normalisation-event batch", n);
        anyW = anyW || b->isWeighted();
        allW = allW && b->isWeighted();
    }
    if(anyW && !allW) throw GeneralError("Amp4Body_TD {}: weighted and unweighted normalisation batches cannot be mixed", n);
    std::vector<long> seeds;
    unsigned int perBatch = 0;
    if(!anyW) {
        for(auto &b : own) {
            if(b->toGenerate == 0) throw GeneralError("Amp4Body_TD {}: an unweighted batch must come from buildBatches()", n);
            if(perBatch && b->toGenerate != perBatch) throw GeneralError("Amp4Body_TD {}: batches of different sizes", n);
            perBatch = b->toGenerate;
            seeds.push_back(b->seed);
        }
    }
    init(observables, mistag, seeds, perBatch);
    if(anyW) {
        std::vector<std::vector<fptype>> all(5);
        sumInitWeights = 0;
        for(auto &b : own) {
            for(int k = 0; k < 5; k++) all[k].insert(all[k].end(), b->cols[k].begin(), b->cols[k].end());
            normDtime.insert(normDtime.end(), b->dtime.begin(), b->dtime.end());
            normSigma.insert(normSigma.end(), b->sigma.begin(), b->sigma.end());
            normWeights.insert(normWeights.end(), b->weights.begin(), b->weights.end());
            int acc = 0; // thrust::reduce(thrust::host, weights.begin(), weights.end(), 0): the accumulator has the type of the init value
            for(fptype w : b->weights) acc = acc + static_cast<int>(w);
            sumInitWeights += acc;
        }
        auto wkeep = normWeights; // setNormEvents describes an unweighted sample: keep the weighted columns across it
        auto tkeep = normDtime, skeep = normSigma;
        setNormEvents(all);
        normWeights  = wkeep;
        normDtime    = tkeep;
        normSigma    = skeep;
        weightedNorm = true;
        normMode     = NormMode::PerEvent; // the integrand depends on tau, x, y event by event: there is no Gram form
    }
}

void Amp4Body_TD::init(std::vector<Observable> &observables, Observable *mistag, const std::vector<long> &seeds, unsigned int perBatch) {
    const std::string n = getName();
    // m12, m34, cos12, cos34, phi, eventnumber, dtime, sigmat — in this order (Amp4Body_TD.cu:327-330)
    if(observables.size() < 8) throw GeneralError("Amp4Body_TD {} needs m12, m34, cos12, cos34, phi, event number, dtime and sigma_t", n);
    if(decayInfo.particle_masses.size() != 5) throw GeneralError("Amp4Body_TD {}: particle_masses must hold the mother and four daughters", n);
    if(decayInfo.amplitudes.empty() && decayInfo.amplitudes_B.empty()) throw GeneralError("Amp4Body_TD {} needs at least one amplitude", n);
    if(!efficiency) throw GeneralError("Amp4Body_TD {} needs an efficiency function", n);
    if(!resolution) throw GeneralError("Amp4Body_TD {} needs a decay-time resolution", n);
    totalEventSize = static_cast<int>(observables.size()) + 2;
    for(auto &o : observables) registerObservable(o);
    for(double pm : decayInfo.particle_masses) registerConstant(pm);
    registerParameter(decayInfo._tau);
    registerParameter(decayInfo._xmixing);
    registerParameter(decayInfo._ymixing);
    registerParameter(decayInfo._SqWStoRSrate);
    // Amp4Body_TD.cu:349-476: A amplitudes first (flag 0, or 2 when the same amplitude is also in the B list, which adds a
    // second coefficient pair right behind the first), then the B amplitudes not seen yet (flag 1)
    std::vector<Amplitude *> seen;
    for(auto *amp : decayInfo.amplitudes) {
        seen.push_back(amp);
        components.push_back(amp);
        registerParameter(amp->_ar);
        registerParameter(amp->_ai);
        addAmplitudeRecord(amp);
        auto inB = std::find_if(decayInfo.amplitudes_B.begin(), decayInfo.amplitudes_B.end(), [amp](const Amplitude *A) { return *amp == *A; });
        if(inB != decayInfo.amplitudes_B.end()) {
            registerParameter((*inB)->_ar);
            registerParameter((*inB)->_ai);
            flags.push_back(2);
        } else
            flags.push_back(0);
    }
    for(auto *amp : decayInfo.amplitudes_B) {
        if(std::find_if(seen.begin(), seen.end(), [amp](const Amplitude *A) { return *amp == *A; }) != seen.end()) continue;
        components.push_back(amp);
        registerParameter(amp->_ar);
        registerParameter(amp->_ai);
        addAmplitudeRecord(amp);
        flags.push_back(1);
    }
    components.push_back(resolution);
    components.push_back(efficiency);
    if(mistag) {
        registerObservable(*mistag);
        totalEventSize = 9;
    }
    registerFunction("ptr_to_Amp4Body_TD", GF_OP_AMP4_TD);
    // Normalisation sample: per seed, phase space -> accept-reject -> COMPACTED accepted events
    // (detail/NormEvents_4Body_Base.cu:24-78; unlike Amp4Body, upstream does run the remove_if here)
    // The sample itself is generated on the device when it is first needed (materialiseNormEvents).
    if(perBatch > 0 && !seeds.empty()) normSpec = NormSpec{seeds, perBatch, true, decayInfo.particle_masses};
    setSeparateNorm();
}
void Amp4Body_TD::emitAux(detail::ProgramBuilder &B) {
    Amp4BodyBase::emitAux(B);
    resolution->emit(B);
}
int Amp4Body_TD::emit(detail::ProgramBuilder &B) {
    int first = (int)B.nodes.size();
    efficiency->emit(B);
    int tdOff = 5 + (int)LineShapes.size() + (int)SpinFactors.size() + (int)ampStarts.size() + (int)ampRecords.size();
    std::vector<int32_t> ints = headerInts(B, tdOff);
    ints.push_back(resolution->getFunctionIndex());
    ints.insert(ints.end(), flags.begin(), flags.end());
    return emitRecord(B, GF_OP_AMP4_TD, 1, (int)B.nodes.size() - first + 1, ints);
}
fptype Amp4Body_TD::normalize() { // Amp4Body_TD.cu:744-799
    recursiveSetNormalization(1.0);
    ensureCache();
    Engine &E  = Engine::get();
    size_t nLS = LineShapes.size();
    std::vector<uint8_t> lsRedo(nLS), ampRedo(nAmps, 0);
    bool any = forceRedoIntegrals;
    for(size_t i = 0; i < nLS; ++i) { // areLineshapesChanged, Amp4Body_TD.cu:690-704
        lsRedo[i] = forceRedoIntegrals || LineShapes[i]->parametersChanged();
        any       = any || lsRedo[i];
    }
    for(size_t a = 0; a < nAmps; ++a) { // areAmplitudeComponentsChanged, Amp4Body_TD.cu:706-725: per amplitude, its own lineshapes
        auto *amp  = dynamic_cast<Amplitude *>(components[a]);
        ampRedo[a] = forceRedoIntegrals || (amp && amp->lineshapeParametersChanged());
        any        = any || ampRedo[a];
    }
    int spins          = forceRedoIntegrals ? 1 : 0;
    forceRedoIntegrals = false;
    E.sync();
    if(generation_no_norm) { // GenerateSig: data-side caches only, normalisation 1 (Amp4Body_TD.cu:764-765,946-950)
        if(any) Engine::check(gf_amp4_cache_data(E.ctx(), nodeIdx, spins, lsRedo.data(), ampRedo.data()), "gf_amp4_cache_data");
        setNormalization(1.0);
        return 1.0;
    }
    if(normSampleEmpty) throw GeneralError("{}: no normalisation event was accepted", getName());
    if(weightedNorm) { // NormEvents_4Body_Weighted::computeNorm_TD: sum_m w_m pdf(x_m, t_m, sigma_m) / sum of the initial weights
        if(any) Engine::check(gf_amp4_cache(E.ctx(), nodeIdx, spins, lsRedo.data(), ampRedo.data()), "gf_amp4_cache");
        double sum = 0;
        Engine::check(gf_amp4_norm(E.ctx(), nodeIdx, &sum), "gf_amp4_norm");
        fptype ret = sum / getSumInitNormEventWeights();
        setNormalization(1.0 / ret);
        return ret;
    }
    refreshCaches(any, spins, lsRedo, ampRedo);
    double sums[4] = {0, 0, 0, 0};
    if(normMode == NormMode::Gram) {
        // effective coefficient vectors of A (scaled by sqrt(WS/RS)) and B with the reference's coefficient walk
        // (NormIntegrator_TD.cu:69-92); then |A|^2, |B|^2 and A conj(B) are bilinear forms in G
        std::vector<std::complex<fptype>> cA(nAmps, {0.0, 0.0}), cB(nAmps, {0.0, 0.0});
        const fptype sq = par(3);
        int k = 0;
        for(size_t a = 0; a < nAmps; ++a) {
            std::complex<fptype> c(par(4 + 2 * (a + k)), par(5 + 2 * (a + k)));
            if(flags[a] == 1)
                cB[a] = c;
            else {
                cA[a] = c * sq;
                if(flags[a] == 2) {
                    ++k;
                    cB[a] = {par(4 + 2 * (a + k)), par(5 + 2 * (a + k))};
                }
            }
        }
        const auto &g = gram(false); // refreshCaches left G current
        std::complex<fptype> AA(0, 0), BB(0, 0), AB(0, 0);
        for(size_t a = 0; a < nAmps; ++a)
            for(size_t b = 0; b < nAmps; ++b) {
                const std::complex<fptype> gab = g[a * nAmps + b];
                AA += cA[a] * std::conj(cA[b]) * gab;
                BB += cB[a] * std::conj(cB[b]) * gab;
                AB += cA[a] * std::conj(cB[b]) * gab;
            }
        sums[0] = AA.real(); sums[1] = BB.real(); sums[2] = AB.real(); sums[3] = AB.imag();
    } else
        Engine::check(gf_amp4_norm(E.ctx(), nodeIdx, sums), "gf_amp4_norm");
    // the reference reads tau, x, y from the Variables here (parametersList[0..2], Amp4Body_TD.cu:770-772), while the
    // device sums above used the parameter block last pushed by updateParameters()
    fptype tau = parametersList[0].getValue(), x = parametersList[1].getValue(), y = parametersList[2].getValue();
    fptype ret = resolution->normalization(sums[0], sums[1], sums[2], sums[3], tau, x, y) / getSumInitNormEventWeights();
    setNormalization(1.0 / ret);
    return ret;
}


// Amp4Body_TD.cu:825-1018. Phase-space candidates are unweighted and COMPACTED first (the remove_if over the accept flags), decay
// times are drawn from exp(-gammamin t) with gammamin = (1 - |y|) / tau, the PDF (normalisation forced to 1) is evaluated once at
// t = 0 to find the envelope's height (1.1 x the largest value) and once at the drawn times; event i is accepted when
// u_i x wmax x exp(-gammamin t_i) < pdf_i. Returned: the accepted candidates' four-vectors, (m12, m34, cos12, cos34, phi, t), the
// t = 0 values ("weights") and the flags. The uniform numbers come from this build's counter-based generator (MCBooster's
// stream and thrust's minstd are not reproduced), keyed on (seed, generation_offset + index) like the phase space.
GeneratedSignal Amp4Body_TD::GenerateSigColumns(unsigned int numEvents, int seed) {
    GeneratedSignal G;
    std::vector<std::vector<fptype>> P, V5;
    std::vector<fptype> w;
    devicePhaseSpace(decayInfo.particle_masses, numEvents, static_cast<unsigned long>(seed), generation_offset, &P, w, &V5);
    std::vector<bool> acc(numEvents);
    size_t nAcc = 0;
    for(size_t i = 0; i < numEvents; i++) { // mcbooster::PhaseSpace::Unweight: weight / max weight against a uniform number
        acc[i] = w[i] > detail::counterUniform(static_cast<unsigned long>(seed), i + generation_offset, 1000);
        nAcc += acc[i];
    }
    if(nAcc == 0) throw GeneralError("{}: no phase-space candidate was accepted", getName());
    G.particles.assign(4, std::vector<fptype>(4 * nAcc));
    G.variables.assign(6, std::vector<fptype>(nAcc));
    const fptype tau = parametersList[0].getValue(), ymixing = parametersList[2].getValue();
    const fptype gammamin = 1.0 / tau - std::fabs(ymixing) / tau;
    for(size_t i = 0, k = 0; i < numEvents; i++) {
        if(!acc[i]) continue;
        for(int d = 0; d < 4; d++) std::copy(P[d].begin() + 4 * i, P[d].begin() + 4 * i + 4, G.particles[d].begin() + 4 * k);
        for(int v = 0; v < 5; v++) G.variables[v][k] = V5[v][i];
        G.variables[5][k] = -std::log(detail::counterUniform(static_cast<unsigned long>(seed), k + generation_offset, 3000)) / gammamin; // genExp
        k++;
    }
    std::vector<fptype> evt(nAcc), zeros(nAcc, 0.0);
    std::iota(evt.begin(), evt.end(), 0.0);
    std::vector<Observable> obs(observablesList.begin(), observablesList.begin() + 8);
    auto upload = [&](const std::vector<fptype> &times) {
        generated = std::make_unique<UnbinnedDataSet>(obs);
        generated->appendColumns({G.variables[0].data(), G.variables[1].data(), G.variables[2].data(), G.variables[3].data(), G.variables[4].data(),
                                  evt.data(), times.data(), zeros.data()},
                                 nAcc);
        setData(generated.get());
        setDataSize(nAcc, 8);
    };
    auto fc = fitControl;
    const bool hadNorm = generation_no_norm;
    generation_no_norm = true; // no MC integral while generating: the value is used un-normalised (Amp4Body_TD.cu:946-950)
    setFitControl(std::make_shared<EvalFit>());
    upload(zeros); // the envelope's height: the PDF at t = 0
    normalize();
    G.weights = evaluate_with_metric();
    fptype wmax = 1.1 * *std::max_element(G.weights.begin(), G.weights.end());
    if(wmax > maxWeight && maxWeight != 0)
        throw GeneralError("WARNING: you just encountered a higher maximum weight than observed in previous iterations.\n"
                           "WARNING: Consider recalculating your AccRej flags and accepting based upon these.\n"
                           "WARNING: previous weight: {}, new weight: {}\n",
                           maxWeight, wmax);
    maxWeight = wmax > maxWeight ? wmax : maxWeight;
    upload(G.variables[5]);
    normalize();
    std::vector<fptype> results = evaluate_with_metric();
    setFitControl(fc);
    generation_no_norm = hadNorm;
    setForceIntegrals();
    G.flags.resize(nAcc);
    for(size_t k = 0; k < nAcc; k++) { // exp_functor, Amp4Body_TD.cu:76-117
        const fptype envelope = std::exp(-G.variables[5][k] * gammamin) * maxWeight;
        G.flags[k] = detail::counterUniform(static_cast<unsigned long>(seed), k + generation_offset, 4000) * envelope < results[k];
    }
    return G;
}

} // namespace GooFit
