// gf_ops.cuh — per-node device math and the post-order stack machine.
// Each case cites the reference device function it reproduces (paths under /root/reference/src/PDFs).
#pragma once
#include "gf_math.cuh"

// std::pow(double, int) of the reference promotes to pow(double, double). For k = 0, 1, 2 the results 1, x and the
// correctly rounded x*x are what a correctly rounded pow returns, without ~200 instructions of generic pow; higher
// powers keep calling pow so that no extra rounding steps are introduced.
__device__ inline double gf_pow_i(double x, int k) {
    if(k == 0) return 1.0;
    if(k == 1) return x;
    if(k == 2) return x * x;
    return pow(x, (double)k);
}

// t^power exp(slope t) of both Argus forms (ArgusPdf.cu:7-47) for t > 0. The generic double pow is ~200 instructions out of line
// and was 40 % of the zachFit likelihood kernel (two Argus shapes per event): the default power 1/2 is a square root (exactly
// rounded, as the host libm's pow returns for 1/2), and any other power goes through exp(power log t + slope t): relative error
// ~|power log t| 2^-53, i.e. < 1e-14 for the t > 1e-30 a double-precision event can produce - against the 1e-12 of the parity tests.
__device__ inline double gf_argus_shape(double t, double power, double slope) {
    if(power == 0.5) return sqrt(t) * exp(slope * t);
    return exp(power * log(t) + slope * t);
}

// KinLimitBWPdf.cu:7-24
__device__ inline double gf_klbw_momentum(double mass, double pimass, double d0mass) {
    if(mass <= 0) return 0;
    double lambda = mass * mass - pimass * pimass - d0mass * d0mass;
    lambda *= lambda;
    lambda -= 4 * pimass * pimass * d0mass * d0mass;
    if(lambda <= 0) return 0;
    return sqrt(0.5 * lambda / mass);
}
__device__ inline double gf_klbw_factor(double momentum) { return 1 / sqrt(1.0 + 2.56 * momentum * momentum); }

// Landau density by the CERNLIB G110 DENLAN rational approximations (Koelbig & Schorr, Comput. Phys. Commun. 31 (1984) 97), the
// algorithm basic/LandauPdf.cu:8-86 evaluates: seven ranges of v = (x - mpv) / sigma, each a ratio of two quartic polynomials in v
// or 1/v (Horner form, innermost coefficient first as upstream), with asymptotic forms below -5.5 and above 300.
__device__ inline double gf_quartic(const double (&k)[5], double t) { return k[0] + (k[1] + (k[2] + (k[3] + k[4] * t) * t) * t) * t; }
__device__ inline double gf_landau(double x, double mpv, double sigma) {
    if(sigma <= 0) return 0;
    // rows: -5.5 <= v < -1, -1 <= v < 1, 1 <= v < 5 (in v); 5 <= v < 12, 12 <= v < 50, 50 <= v < 300 (in 1/v)
    const double num[6][5] = {{0.4259894875, -0.1249762550, 0.03984243700, -0.006298287635, 0.001511162253},
                              {0.1788541609, 0.1173957403, 0.01488850518, -0.001394989411, 0.0001283617211},
                              {0.1788544503, 0.09359161662, 0.006325387654, 0.00006611667319, -0.000002031049101},
                              {0.9874054407, 118.6723273, 849.2794360, -743.7792444, 427.0262186},
                              {1.003675074, 167.5702434, 4789.711289, 21217.86767, -22324.94910},
                              {1.000827619, 664.9143136, 62972.92665, 475554.6998, -5743609.109}};
    const double den[6][5] = {{1.0, -0.3388260629, 0.09594393323, -0.01608042283, 0.003778942063},
                              {1.0, 0.7428795082, 0.3153932961, 0.06694219548, 0.008790609714},
                              {1.0, 0.6097809921, 0.2560616665, 0.04746722384, 0.006957301675},
                              {1.0, 106.8615961, 337.6496214, 2016.712389, 1597.063511},
                              {1.0, 156.9424537, 3745.310488, 9834.698876, 66924.28357},
                              {1.0, 651.4101098, 56974.73333, 165917.4725, -2815759.939}};
    const double v = (x - mpv) / sigma;
    double d;
    if(v < -5.5) {
        const double u = exp(v + 1.0);
        if(u < 1e-10) return 0.0;
        d = 0.3989422803 * (exp(-1 / u) / sqrt(u)) * (1 + (0.04166666667 + (-0.01996527778 + 0.02709538966 * u) * u) * u);
    } else if(v < -1) {
        const double u = exp(-v - 1);
        d = exp(-u) * sqrt(u) * gf_quartic(num[0], v) / gf_quartic(den[0], v);
    } else if(v < 5) {
        const int r = v < 1 ? 1 : 2;
        d = gf_quartic(num[r], v) / gf_quartic(den[r], v);
    } else if(v < 300) {
        const int r = v < 12 ? 3 : (v < 50 ? 4 : 5);
        const double u = 1 / v;
        d = u * u * gf_quartic(num[r], u) / gf_quartic(den[r], u);
    } else {
        const double u = 1 / (v - v * log(v) / (v + 1));
        d = u * u * (1 + (-1.845568670 + -4.284640743 * u) * u);
    }
    return d / sigma;
}

__device__ inline bool gf_in_dalitz(double m12, double m13, double bigM, double dm1, double dm2, double dm3);

// Re w(z), z = x + i y, of the complex error function by the Matta-Reichel series as RooFit's fast ComplexErrorFunc / CERNLIB WWERF
// evaluate it and basic/VoigtianPdf.cu:41-104 restates it: twelve terms exp(-n_k) / ((d + n_k)^2 + f) on the grid n_k = (h k)^2,
// h = 0.5 (0.53 when 2|x| is within 0.02 of an integer on the real axis, where the poles of the h = 0.5 correction term sit), a pole
// correction below Im z = 2 pi, and a [6/6] Pade form of erf for |z|^2 < 1e-7. The node tables are generated (n_k and exp(-n_k))
// instead of spelled out; the Pade coefficients are the published ones.
__device__ inline double gf_faddeeva_re(double x, double y) {
    const double s = x * x + y * y;
    if(s < 1e-7) { // w = exp(-z^2) + i (2/sqrt(pi)) z exp(z^2) P(z^2) / Q(z^2)
        const double pc[7] = {65536.0, -2885792.0, 69973904.0, -791494704.0, 8962513560.0, -32794651890.0, 175685635125.0};
        const double qc[7] = {192192.0, 8648640.0, 183783600.0, 2329725600.0, 18332414100.0, 84329104860.0, 175685635125.0};
        const double zr = x * x - y * y, zi = 2 * x * y; // z^2
        double sn, cs;
        sincos(zi, &sn, &cs);
        const double ex = exp(zr), vr = ex * cs, vi = ex * sn; // v = exp(z^2)
        double cr = pc[0], ci = 0, dr = qc[0], di = 0;
        for(int i = 1; i <= 6; i++) {
            double t = cr * zr - ci * zi;
            ci = cr * zi + ci * zr;
            cr = t + pc[i];
            t = dr * zr - di * zi;
            di = dr * zi + di * zr;
            dr = t + qc[i];
        }
        const double vn = vr * vr + vi * vi, inv_r = vr / vn; // Re(1 / v)
        // i k c / d z v with k = 2 / sqrt(pi), left to right
        double ar = -1.1283791670955126 * ci, ai = 1.1283791670955126 * cr;
        const double dn = dr * dr + di * di;
        double t = (ar * dr + ai * di) / dn;
        ai = (ai * dr - ar * di) / dn;
        ar = t;
        t = ar * x - ai * y;
        ai = ar * y + ai * x;
        ar = t;
        return inv_r + (ar * vr - ai * vi);
    }
    double h = 0.5;
    if(fabs(y) < 0.01 && fabs(x) < 6.01) {
        double frac = fabs(x) * 2;
        frac -= floor(frac);
        if(frac < 0.02 || frac > 0.98) h = 0.53;
    }
    const double r = 0.31830988618379067154 * h; // h / pi
    const double d = (y - x) * (y + x), f = 4 * x * x * y * y;
    double g = 0; // (the companion sum over (s - n) u feeds Im w only)
    for(int k = 1; k <= 12; k++) {
        const double n = (h * k) * (h * k);
        const double u = exp(-n) / ((d + n) * (d + n) + f);
        g += (s + n) * u;
    }
    const double inv_s = 1 / s;
    double re = r * (y * (inv_s + 2.0 * g));
    if(y < 6.28318530717958) {
        const double k = 2.0 / r;
        double sn, cs;
        sincos(k * x, &sn, &cs);
        const double ff = exp(-(k * y)) - cs;
        const double gg = 2.0 * exp(d - k * y) / (sn * sn + ff * ff);
        double s2, c2;
        sincos(2.0 * x * y, &s2, &c2);
        re += gg * (c2 * ff - s2 * sn);
    }
    return re;
}

template <class Evt>
__device__ inline double gf_leaf(const gf_node &nd, const gf_sprog &P, const Evt &evt) {
    const double *p = P.params + nd.par_off;
    const double *c = P.consts + nd.con_off;
    const int32_t *ob = P.ints + nd.obs_off;
    switch(nd.op) {
    case GF_OP_EXP: // basic/ExpPdf.cu:7-17
        return exp(p[0] * evt(ob[0]));
    case GF_OP_EXP_OFFSET: { // ExpPdf.cu:19-30
        double x = evt(ob[0]);
        x -= p[0];
        return exp(p[1] * x);
    }
    case GF_OP_EXP_POLY: { // ExpPdf.cu:32-48
        double x = evt(ob[0]), a = 0;
        for(int i = 0; i < nd.npar; ++i) a += gf_pow_i(x, i) * p[i];
        return exp(a);
    }
    case GF_OP_EXP_POLY_OFFSET: { // ExpPdf.cu:50-67
        double x = evt(ob[0]) - p[0], a = 0;
        for(int i = 1; i < nd.npar; ++i) a += gf_pow_i(x, i) * p[i];
        return exp(a);
    }
    case GF_OP_GAUSSIAN: { // basic/GaussianPdf.cu:7-17
        double x = evt(ob[0]), mean = p[0], sigma = p[1];
        return exp(-0.5 * (x - mean) * (x - mean) / (sigma * sigma));
    }
    case GF_OP_CONSTANT: return p[nd.npar - 1]; // p * pow(x, 0): the flat background / constant efficiency of every shipped example
    case GF_OP_POLYNOMIAL: { // basic/PolynomialPdf.cu:8-25
        int lowest = (int)c[0];
        double x = evt(ob[0]), ret = 0;
        for(int i = 0; i < nd.npar; ++i) ret += p[i] * gf_pow_i(x, lowest + i);
        return ret;
    }
    case GF_OP_OFFSET_POLYNOMIAL: { // PolynomialPdf.cu:27-45 (offset itself is not applied by the reference)
        int lowest = (int)c[0];
        double x = evt(ob[0]), ret = 0;
        for(int i = 2; i < nd.npar; ++i) ret += p[i] * gf_pow_i(x, lowest + i - 2);
        return ret;
    }
    case GF_OP_MULTI_POLYNOMIAL: { // PolynomialPdf.cu:47-118
        int nobs = nd.nobs, maxDegree = (int)c[0] + 1, numBoxes = 1;
        if(maxDegree == 1) return p[nobs]; // degree 0 (the constant efficiency of the amplitude fits): one box
        for(int i = 0; i < nobs; ++i) numBoxes *= maxDegree;
        int coeff = nobs;
        double ret = p[coeff++];
        for(int i = 1; i < numBoxes; ++i) {
            double term = 1;
            int cur = i, sum = 0;
            for(int j = 0; j < nobs; ++j) {
                double x = evt(ob[j]) - p[j];
                int pw = cur % maxDegree;
                cur /= maxDegree;
                term *= gf_pow_i(x, pw);
                sum += pw;
            }
            if(sum >= maxDegree) continue;
            term *= p[coeff++];
            ret += term;
        }
        return ret;
    }
    case GF_OP_ARGUS_UPPER: { // basic/ArgusPdf.cu:7-25
        double x = evt(ob[0]), m0 = p[0], slope = p[1], power = p[2];
        double t = x / m0;
        if(t >= 1) return 0;
        t = 1 - t * t;
        return x * gf_argus_shape(t, power, slope);
    }
    case GF_OP_ARGUS_LOWER: { // ArgusPdf.cu:27-47
        double x = evt(ob[0]), m0 = p[0], slope = p[1], power = p[2];
        double t = x / m0;
        if(t <= 1) return 0;
        t *= t;
        t -= 1;
        return x * gf_argus_shape(t, power, slope);
    }
    case GF_OP_KINLIMITBW: { // basic/KinLimitBWPdf.cu:26-54
        double x = evt(ob[0]), mean = p[0], width = p[1], d0mass = c[0], pimass = c[1];
        mean += d0mass;
        x += d0mass;
        double pReal = gf_klbw_momentum(mean, pimass, d0mass);
        if(0 >= pReal) return 0;
        mean *= mean;
        double pX = gf_klbw_momentum(x, pimass, d0mass);
        double r = pX / pReal, b = gf_klbw_factor(pX) / gf_klbw_factor(pReal);
        double phsp = (r * r * r) * (b * b);
        double dm = mean - x * x;
        double massSq = dm * dm;
        double wg = width * phsp;
        double gammaSq = wg * wg;
        return (phsp * mean * width * width) / (massSq + mean * gammaSq);
    }
    case GF_OP_SCALED_GAUSSIAN: { // basic/ScaledGaussianPdf.cu:9-20
        double x = evt(ob[0]);
        double mean = p[0] + p[2], sigma = p[1] * (1 + p[3]);
        return exp(-0.5 * (x - mean) * (x - mean) / (sigma * sigma));
    }
    case GF_OP_JOHNSONSU: { // basic/JohnsonSUPdf.cu:8-24 (SQRT2PI truncated to 2.506628 as the reference)
        double x = evt(ob[0]);
        double px = (x - p[0]) / p[1];
        double sq = sqrt(1 + px * px);
        double g = p[2] + p[3] * log(px + sq);
        return p[3] / (p[1] * 2.506628 * sq) * exp(-0.5 * g * g);
    }
    case GF_OP_BW: { // basic/BWPdf.cu:6-18 (its "rootPi" is -2 atan2(-1, 0) = pi)
        double x = evt(ob[0]), mean = p[0], gamma = p[1];
        double pi_ = -2. * atan2(-1.0, 0.0);
        return (gamma / ((x - mean) * (x - mean) + gamma * gamma / 4)) / (2 * pi_);
    }
    case GF_OP_BIFURGAUSS: { // basic/BifurGaussPdf.cu:6-22
        double x = evt(ob[0]), mean = p[0];
        double sigma = x > mean ? p[2] : p[1];
        return exp(-0.5 * (x - mean) * (x - mean) / (sigma * sigma));
    }
    case GF_OP_CRYSTALBALL: { // basic/CrystalBallPdf.cu:6-35
        double x = evt(ob[0]), mean = p[0], sigma = p[1], alpha = p[2], power = p[3];
        double rx = (sigma != 0) ? (x - mean) / sigma : 0;
        if((alpha > 0 && rx <= alpha) || (alpha < 0 && rx >= alpha) || alpha == 0) return exp(-0.5 * rx * rx);
        double n_over_alpha = power / alpha;
        double a = exp(-0.5 * alpha * alpha);
        double d = (n_over_alpha - alpha) + rx;
        d = (d != 0) ? n_over_alpha / d : 0;
        return a * pow(d, power);
    }
    case GF_OP_EXPGAUS: { // basic/ExpGausPdf.cu:6-23
        double x = evt(ob[0]), mean = p[0], sigma = p[1], alpha = p[2];
        double ret = 0.5 * alpha;
        double exparg = ret * (2 * mean + alpha * sigma * sigma - 2 * x);
        double erfarg = (mean + alpha * sigma * sigma - x) / (sigma * 1.4142135623);
        ret *= exp(exparg);
        ret *= erfc(erfarg);
        return ret;
    }
    case GF_OP_NOVOSIBIRSK: { // basic/NovosibirskPdf.cu:6-38
        double mean = p[0], sigma = p[1], tail = p[2], x = evt(ob[0]);
        double qc;
        if(fabs(tail) < 1.e-7) {
            double t = (x - mean) / sigma;
            qc = 0.5 * (t * t);
        } else {
            double qa = tail * sqrt(log(4.));
            double qb = sinh(qa) / qa;
            double qx = (x - mean) / sigma * qb;
            double qy = 1. + tail * qx;
            if(qy > 1.e-7) {
                double l = log(qy) / tail;
                qc = 0.5 * (l * l + tail * tail);
            } else
                qc = 15.0;
        }
        return exp(-qc);
    }
    case GF_OP_STEP: { // basic/StepPdf.cu:6-20
        double x = evt(ob[0]);
        bool above = x > p[0];
        return ((int)c[0] == 1) == above ? 1.0 : 0.0;
    }
    case GF_OP_TRIGTHRESH_UPPER:
    case GF_OP_TRIGTHRESH_LOWER:
    case GF_OP_TRIGTHRESH2_UPPER:
    case GF_OP_TRIGTHRESH2_LOWER: { // basic/TrigThresholdPdf.cu:6-75: 1 beyond half a period from the threshold, a sine edge before
        const bool two = nd.op >= GF_OP_TRIGTHRESH2_UPPER, upper = nd.op == GF_OP_TRIGTHRESH_UPPER || nd.op == GF_OP_TRIGTHRESH2_UPPER;
        double v = two ? p[3] - evt(ob[0]) - evt(ob[1]) : evt(ob[0]);
        double distance = p[1] * (upper ? p[0] - v : v - p[0]);
        double lin = p[2];
        return distance > 0.5 ? 1.0 : (lin + (1 - lin) * sin(distance * 3.14159265));
    }
    case GF_OP_BERNSTEIN: { // basic/BernsteinPdf.cu:8-51: binomials by factorials as upstream (a / (b1 * b2) in doubles)
        double x = evt(ob[0]), lowX = c[1], uppX = c[2];
        int mindeg = (int)(unsigned int)c[3];
        int maxDegree = (int)c[0] + mindeg;
        double xn = (x - lowX) / (uppX - lowX), ret = 0;
        auto fact = [](int n) { double f = 1; for(int k = 1; k <= n; ++k) f = k * f; return f; };
        for(int i = 0; i < maxDegree - mindeg; i++) {
            double val = p[i];
            int n = maxDegree - 1;
            double bin = i > n ? 0.0 : fact(n) / (fact(i) * fact(n - i));
            val *= bin * pow(xn, (double)i) * pow(1 - xn, (double)(n - i));
            val /= (uppX - lowX);
            ret += val;
        }
        return ret;
    }
    case GF_OP_CORRGAUSS: { // basic/CorrGaussianPdf.cu:6-27: the y width grows with the x distance
        double x = evt(ob[0]), y = evt(ob[1]);
        double xd = (x - p[0]) / p[1];
        double s2 = p[3] * (1.0 + p[4] * xd * xd);
        double yd = (y - p[2]) * (s2 == 0.0 ? 0.0 : (1.0 / s2));
        return exp(-0.5 * (xd * xd + yd * yd));
    }
    case GF_OP_BINTRANSFORM: { // basic/BinTransformPdf.cu:6-27: the global bin number of the event as a value
        int ret = 0, previous = 1;
        for(int i = 0; i < nd.nobs; ++i) {
            int local = (int)floor((evt(ob[i]) - c[3 * i]) / c[3 * i + 1]);
            ret += local * previous;
            previous *= (int)c[3 * i + 2];
        }
        return (double)ret;
    }
    case GF_OP_LANDAU: return gf_landau(evt(ob[0]), p[0], p[1]);
    case GF_OP_VOIGTIAN: { // basic/VoigtianPdf.cu:285-320: the second parameter is read as the width, the third as the sigma
        double x = evt(ob[0]), m = p[0], w = p[1], s = p[2];
        if(0 == s && 0 == w) return 1;
        double arg = x - m;
        if(0 == s) return 1 / (arg * arg + 0.25 * w * w);
        double coef = -0.5 / (s * s);
        if(0 == w) return exp(coef * arg * arg);
        double cc = 0.707106781187;
        cc /= s;
        return cc * 0.5641895835477563 * gf_faddeeva_re(cc * arg, 0.5 * cc * w);
    }
    case GF_OP_VARBINTRANSFORM: { // utilities/VariableBinTransform1DPdf.cu:8-27: bin number of |x| among ascending limits
        double v = fabs(evt(ob[0]));
        int n = (int)c[0], ret = 0;
        for(int i = 0; i < n; ++i) {
            if(v < c[1 + i]) break;
            ret++;
        }
        return (double)ret;
    }
    case GF_OP_DALITZ_VETO: { // physics/DalitzVetoPdf.cu:7-33: 1 inside the Dalitz plot and outside every veto band, else 0
        double x = evt(ob[0]), y = evt(ob[1]);
        double M = p[0], m1 = p[1], m2 = p[2], m3 = p[3];
        double z = (M * M + m1 * m1 + m2 * m2 + m3 * m3) - x - y;
        double ret = gf_in_dalitz(x, y, M, m1, m2, m3) ? 1.0 : 0.0;
        const int nv = (int)(unsigned int)c[0];
        for(int i = 0; i < nv; ++i) {
            const int pair = (int)(unsigned int)c[1 + i];
            const double lo = p[4 + 2 * i], hi = p[5 + 2 * i];
            const double v = pair == 0 ? x : (pair == 1 ? y : z); // PAIR_12, PAIR_13, else PAIR_23
            ret *= (v < hi && v > lo) ? 0.0 : 1.0;
        }
        return ret;
    }
    case GF_OP_SQDALITZ_EFF: { // physics/SquareDalitzEffPdf.cu:10-74: polynomial in m23 and theta' with D -> KS0 h h masses built in
        const double x = evt(ob[0]), y = evt(ob[1]);
        const double mD = 1.86483, mK = 0.497611, mh1 = 0.13957, mh2 = 0.13957;
        if(!gf_in_dalitz(x, y, mD, mK, mh1, mh2)) return 0;
        const double m23 = mD * mD + mK * mK + mh1 * mh1 + mh2 * mh2 - x - y;
        double th;
        if(m23 < 0)
            th = 0;
        else {
            const double num = m23 * (x - y);
            const double t1 = m23 - mh1 * mh1 + mh2 * mh2, t2 = mD * mD - mK * mK - m23;
            const double l1 = t1 * t1 - 4 * m23 * mh2 * mh2, l2 = t2 * t2 - 4 * m23 * mK * mK;
            const double den = sqrt(l1) * sqrt(l2);
            th = (l1 < 0. || l2 < 0. || den == 0.) ? -99 : num / den;
        }
        if(th > 1. || th < -1.) return 0;
        if(m23 < 0 || m23 > 2) return 0;
        return p[0] * m23 * m23 + p[1] * m23 + p[2] * m23 * th * th + p[3] * th * th + p[4] * th + p[5] + p[6] * m23 * m23 * m23 * m23
               + p[7] * m23 * m23 * m23;
    }
    case GF_OP_HISTOGRAM: { // basic/SmoothHistogramPdf.cu:13-46: the content of the event's bin in the table attached by gf_table_set
        const int slot = P.ints[nd.int_off];
        const double *table = static_cast<const double *>(P.aux[slot]);
        const long n = P.aux_ld[slot];
        if(!table || n <= 0) return 0.0;
        const int nv = (int)c[0];
        long global = 0, previous = 1;
        for(int i = 0; i < nv; ++i) {
            double v = (evt(ob[i]) - c[3 * i + 4]) / c[3 * i + 5];
            global += previous * (long)floor(v);
            previous *= (long)c[3 * (i + 1)];
        }
        global = global < 0 ? 0 : (global >= n ? n - 1 : global); // upstream reads outside its table for an event outside the histogram
        return __ldg(table + global);
    }
    default: return 0.0;
    }
}

// Dalitz-plot boundary, DalitzPlotHelpers.cu:14-59
__device__ inline bool gf_in_dalitz(double m12, double m13, double bigM, double dm1, double dm2, double dm3) {
    double dm1pdm2 = dm1 + dm2, bigMmdm3 = bigM - dm3, bigMmdm1 = bigM - dm1;
    double m23 = bigM * bigM + dm1 * dm1 + dm2 * dm2 + dm3 * dm3 - m12 - m13;
    if(m23 < 0. || m23 > bigMmdm1 * bigMmdm1) return false;
    bool m12less = !(m12 < dm1pdm2 * dm1pdm2);
    bool m12grea = !(m12 > bigMmdm3 * bigMmdm3);
    double sqrtM12 = sqrt(m12), dm11 = dm1 * dm1, dm22 = dm2 * dm2, dm33 = dm3 * dm3;
    double e1star = 0.5 * (m12 - dm22 + dm11) / sqrtM12;
    double e3star = 0.5 * (bigM * bigM - m12 - dm33) / sqrtM12;
    double r1 = sqrt(e1star * e1star - dm11), r3 = sqrt(e3star * e3star - dm33);
    double es = (e1star + e3star) * (e1star + e3star);
    double minimum = es - (r1 + r3) * (r1 + r3);
    double maximum = es - (r1 - r3) * (r1 - r3);
    bool m13less = !(m13 < minimum), m13grea = !(m13 > maximum);
    return m12less && m12grea && m13less && m13grea;
}

// ---- decay-time resolution of Amp4Body_TD ------------------------------------------------------------------------
// physics/ThreeGaussResolution.cu:11-60: one Gaussian-smeared set of the four time functions
__device__ inline void gf_tres_gauss(double &P1, double &P2, double &P3, double &P4, double tau, double adjTime, double x,
                                     double y, double adjSigma, double selbias) {
    const double SQRTPIo2 = 1.0 / 1.12837916709551257390; // 1 / M_2_SQRTPI
    const double SQRT1o2PI = sqrt(0.5 * 0.31830988618379067154);
    const double R1o6 = 1.0 / 6.0;
    double _1oSqrtA = adjSigma * 1.41421356237309504880;
    double _1oSigma = 1 / adjSigma;
    double _1o2SqrtA = 0.5 * _1oSqrtA;
    double _1oSigma2 = _1oSigma * _1oSigma;
    double NormG = SQRT1o2PI * _1oSigma;
    double C = 0.5 * adjTime * adjTime * _1oSigma2;
    double Bgn = -adjTime * _1oSigma2;
    double Gamma = 1 / tau;
    double B = Gamma + Bgn;
    double u0 = _1o2SqrtA * (B + selbias);
    double u02 = u0 * u0;
    double F = _1oSqrtA * exp(-C + u02);
    double Ig0 = SQRTPIo2 * erfc(u0);
    double Ig1 = 0.5 * exp(-u02);
    double Ig2 = Ig1 * u0 + 0.5 * Ig0;
    double Ig3 = Ig1 * (u02 + 1);
    double R = x * Gamma * _1oSqrtA;
    double R2 = R * R;
    double It0 = F * Ig0;
    double It1 = F * R * (Ig1 - u0 * Ig0);
    double It2 = F * R2 * (Ig2 - u0 * Ig1 * 2 + u02 * Ig0);
    double It3 = F * R2 * R * (Ig3 - u0 * Ig2 * 3 + u02 * Ig1 * 3 - u0 * u02 * Ig0);
    P2 = NormG * (It0 - 0.5 * It2);
    P4 = NormG * (It1 - R1o6 * It3);
    double u0py = _1o2SqrtA * (B - y * Gamma + selbias);
    double u0my = _1o2SqrtA * (B + y * Gamma + selbias);
    double Fpy = _1oSqrtA * exp(-C + u0py * u0py);
    double Fmy = _1oSqrtA * exp(-C + u0my * u0my);
    double Ipy = Fpy * SQRTPIo2 * erfc(u0py);
    double Imy = Fmy * SQRTPIo2 * erfc(u0my);
    P1 = NormG * 0.5 * (Ipy + Imy);
    P3 = NormG * 0.5 * (Ipy - Imy);
}
// sin, cos and e^b - 1 for the SMALL arguments of mixing fits (x t / tau, y t / tau with x, y ~ 1e-2): Taylor polynomials in
// Horner form, exact to an ulp or two for |a| <= 0.25 (first neglected terms: a^16/16! < 2e-23 for cos, a^17/17! relative 1e-21
// for sin, b^15/15! relative 3e-21 for expm1), instead of the libm routines with their range reduction (about 150 instructions
// for the three). Larger arguments take the libm path.
__device__ __forceinline__ void gf_sincos_small(double a, double &sn, double &cs) {
    const double a2 = a * a;
    cs = fma(a2, fma(a2, fma(a2, fma(a2, fma(a2, fma(a2, fma(a2, -1.0 / 87178291200.0, 1.0 / 479001600.0), -1.0 / 3628800.0), 1.0 / 40320.0), -1.0 / 720.0),
                             1.0 / 24.0),
                     -0.5),
             1.0);
    sn = a * fma(a2, fma(a2, fma(a2, fma(a2, fma(a2, fma(a2, fma(a2, -1.0 / 1307674368000.0, 1.0 / 6227020800.0), -1.0 / 39916800.0), 1.0 / 362880.0), -1.0 / 5040.0),
                                 1.0 / 120.0),
                         -1.0 / 6.0),
                 1.0);
}
__device__ __forceinline__ double gf_expm1_small(double b) {
    double p = 1.0 / 87178291200.0; // 1/14!
    p = fma(p, b, 1.0 / 6227020800.0);
    p = fma(p, b, 1.0 / 479001600.0);
    p = fma(p, b, 1.0 / 39916800.0);
    p = fma(p, b, 1.0 / 3628800.0);
    p = fma(p, b, 1.0 / 362880.0);
    p = fma(p, b, 1.0 / 40320.0);
    p = fma(p, b, 1.0 / 5040.0);
    p = fma(p, b, 1.0 / 720.0);
    p = fma(p, b, 1.0 / 120.0);
    p = fma(p, b, 1.0 / 24.0);
    p = fma(p, b, 1.0 / 6.0);
    p = fma(p, b, 0.5);
    p = fma(p, b, 1.0);
    return p * b;
}
// op = GF_OP_TRES_TRUTH (TruthResolution.cu:8-28) or GF_OP_TRES_3GAUSS (ThreeGaussResolution.cu:62-125; p = 13 parameters)
template <bool WITH_3GAUSS = true>
__device__ inline double gf_tres(int op, const double *p, double coshterm, double costerm, double sinhterm, double sinterm,
                                 double tau, double dtime, double x, double y, double sigma) {
    if(!WITH_3GAUSS || op == GF_OP_TRES_TRUTH) {
        // same five terms in the same order as the reference; cos/sin share one argument reduction (sincos) and
        // cosh/sinh share one expm1: with u = e^a - 1, sinh a = (u + u/(u+1))/2 and cosh a = 1 + u^2/(2(u+1)), both
        // accurate to an ulp or two for the small a = y t/tau of mixing fits (no e^a - e^-a cancellation)
        double ret = 0;
        dtime *= gf_rcp(tau); // 1 / tau is the same for every event of a call: one reciprocal per thread, not a division per event
        double sn, cs, u;
        const double ax = x * dtime, by = y * dtime;
        if(fabs(ax) <= 0.25 && fabs(by) <= 0.25) {
            gf_sincos_small(ax, sn, cs);
            u = gf_expm1_small(by);
        } else {
            sincos(ax, &sn, &cs);
            u = expm1(by);
        }
        const double ui = u * gf_rcp(u + 1.0);
        const double sh = 0.5 * (u + ui), ch = 1.0 + 0.5 * u * ui;
        ret += coshterm * ch;
        ret += costerm * cs;
        ret -= 2 * sinhterm * sh;
        ret -= 2 * sinterm * sn;
        ret *= exp(-dtime);
        return ret;
    }
    double coreFraction = p[0], tailFraction = p[1];
    double outlFraction = 1 - coreFraction - tailFraction;
    double coreBias = p[2], coreScale = p[3], tailBias = p[4], tailScale = p[5], outlBias = p[6], outlScale = p[7];
    double selbias = p[8];
    coreBias += p[9];
    tailBias += p[11];
    outlBias += p[11];
    coreScale *= p[10];
    tailScale *= p[12];
    outlScale *= p[12];
    double c1, c2, c3, c4, t1, t2, t3, t4, o1, o2, o3, o4;
    gf_tres_gauss(c1, c2, c3, c4, tau, dtime - coreBias * sigma, x, y, coreScale * sigma, selbias);
    gf_tres_gauss(t1, t2, t3, t4, tau, dtime - tailBias * sigma, x, y, tailScale * sigma, selbias);
    gf_tres_gauss(o1, o2, o3, o4, tau, dtime - outlBias * sigma, x, y, outlScale * sigma, selbias);
    double P1 = coreFraction * c1 + tailFraction * t1 + outlFraction * o1;
    double P2 = coreFraction * c2 + tailFraction * t2 + outlFraction * o2;
    double P3 = coreFraction * c3 + tailFraction * t3 + outlFraction * o3;
    double P4 = coreFraction * c4 + tailFraction * t4 + outlFraction * o4;
    double ret = 0;
    ret += coshterm * P1;
    ret += costerm * P2;
    ret -= 2 * sinhterm * P3;
    ret -= 2 * sinterm * P4;
    return ret;
}
// A = sqrt(WS/RS) * sum_{flag 0,2} c A_a ; B = sum_{flag 1,2} c A_a with the reference's coefficient walk
// (Amp4Body_TD.cu:186-214, NormIntegrator_TD.cu:69-88): coefficient pair (i + k), k counting earlier flag-2 amplitudes.
template <int GF_TD_CHUNK = 4, class Load>
__device__ inline void gf_amp4td_AB(const double *p, const int32_t *flags, int namp, const Load &amp, double &Are, double &Aim,
                                    double &Bre, double &Bim) {
    Are = Aim = Bre = Bim = 0;
    int k = 0;
    // the cached amplitudes are fetched GF_TD_CHUNK at a time before the walk starts: one load per trip of a dependent
    // loop leaves a single 16-byte request per thread in flight, about half of what HBM needs at this occupancy
    for(int i0 = 0; i0 < namp; i0 += GF_TD_CHUNK) {
        double2 chunk[GF_TD_CHUNK];
#pragma unroll
        for(int j = 0; j < GF_TD_CHUNK; ++j)
            if(i0 + j < namp) chunk[j] = amp(i0 + j);
#pragma unroll
        for(int j = 0; j < GF_TD_CHUNK; ++j) {
            const int i = i0 + j;
            if(i >= namp) break;
            double2 me = chunk[j];
            int flag = flags[i];
            double cr = p[4 + 2 * (i + k)], ci = p[5 + 2 * (i + k)];
            double re = me.x * cr - me.y * ci, im = me.x * ci + me.y * cr;
            if(flag == 1) {
                Bre += re; Bim += im;
            } else {
                Are += re; Aim += im;
                if(flag == 2) {
                    ++k;
                    cr = p[4 + 2 * (i + k)]; ci = p[5 + 2 * (i + k)];
                    Bre += me.x * cr - me.y * ci;
                    Bim += me.x * ci + me.y * cr;
                }
            }
        }
    }
    Are *= p[3];
    Aim *= p[3];
}

// Which of the heavy node kinds a compiled copy of the interpreter contains. The NLL kernel is instantiated once per
// class of program (gf_core.cu picks the copy from the nodes of the installed program), so that the registers and the
// load batching of, say, the Amp4Body_TD case are not paid by a Dalitz fit whose program has no such node.
// amplitudes fetched together by the TruthResolution-only copy: 4 measured best (2: 0.403 ms, 4: 0.387, 6: 0.425, 8: 0.566,
// the larger batches spill under the 64-register cap)
#define GF_TD_CH 4
// GF_HAS_3GAUSS: the Amp4Body_TD node may carry a ThreeGaussResolution (else TruthResolution only)
enum : int { GF_HAS_CONV = 1, GF_HAS_AMP3 = 2, GF_HAS_AMP4 = 4, GF_HAS_AMP4_TD = 8, GF_HAS_3GAUSS = 16, GF_HAS_AMP3_TD = 32, GF_HAS_COMPOSITE = 64, GF_HAS_ALL = 127 };

// the event as the interpreter reads it: the caller's accessor, with one column substituted while a CompositePdf's shell runs
template <bool ON, class Evt>
struct gf_evt_sub {
    const Evt &in;
    double val;
    int obs;
    __device__ inline double operator()(int ob) const {
        if constexpr(ON) {
            if(ob == obs) return val;
        }
        return in(ob);
    }
    __device__ inline long event_number(int col) const { return in.event_number(col); }
};

// Post-order evaluation of nodes [first, last]; returns the root's raw value.
template <int HAS = GF_HAS_ALL, class Evt>
__device__ inline double gf_eval_tree(const gf_sprog &P, const gf_dev_view &V, int first, int last, const Evt &evt_in) {
    double st[GF_MAX_STACK];
    int sp = 0;
    // CompositePdf evaluates its shell on a copy of the event in which one observable holds the core function's value
    // (CompositePdf.cu:24-31): between the GF_OP_COMPOSITE_ARG and GF_OP_COMPOSITE records reads of that column are substituted
    gf_evt_sub<(HAS & GF_HAS_COMPOSITE) != 0, Evt> evt{evt_in, 0.0, -1};
    for(int k = first; k <= last; ++k) {
        const gf_node &nd = P.nodes[k];
        if(GF_IS_LEAF(nd.op)) {
            st[sp++] = gf_leaf(nd, P, evt);
            continue;
        }
        const double *p = P.params + nd.par_off;
        const int32_t *in = P.ints + nd.int_off;
        switch(nd.op) {
        case GF_OP_ADD: { // combine/AddPdf.cu:16-51 ; ints = child norm slots
            int n = nd.nchild;
            double ret = 0, tw = 0;
            const double *ch = st + (sp - n);
            for(int i = 0; i < n - 1; ++i) {
                double w = p[i];
                tw += w;
                ret += w * ch[i] * P.norms[in[i]];
            }
            ret += (1 - tw) * ch[n - 1] * P.norms[in[n - 1]];
            sp -= n;
            st[sp++] = ret;
        } break;
        case GF_OP_ADD_EXT: { // AddPdf.cu:53-80
            int n = nd.nchild;
            double ret = 0, tw = 0;
            const double *ch = st + (sp - n);
            for(int i = 0; i < n; ++i) {
                double w = p[i];
                ret += w * ch[i] * P.norms[in[i]];
                tw += w;
            }
            ret /= tw;
            sp -= n;
            st[sp++] = ret;
        } break;
        case GF_OP_PROD: { // combine/ProdPdf.cu:9-26
            int n = nd.nchild;
            double ret = 1;
            const double *ch = st + (sp - n);
            for(int i = 0; i < n; ++i) {
                double curr = ch[i] * P.norms[in[i]];
                ret *= curr;
            }
            sp -= n;
            st[sp++] = ret;
        } break;
        case GF_OP_EVTW_ADD: { // combine/EventWeightedAddPdf.cu:8-44: the weights are event columns, not parameters
            int n = nd.nchild;
            const int32_t *ob = P.ints + nd.obs_off;
            double ret = 0, tw = 0;
            const double *ch = st + (sp - n);
            for(int i = 0; i < n - 1; ++i) {
                double w = evt(ob[i]);
                tw += w;
                ret += w * ch[i] * P.norms[in[i]];
            }
            ret += (1 - tw) * ch[n - 1] * P.norms[in[n - 1]];
            sp -= n;
            st[sp++] = ret;
        } break;
        case GF_OP_EVTW_ADD_EXT: { // EventWeightedAddPdf.cu:46-91
            int n = nd.nchild;
            const int32_t *ob = P.ints + nd.obs_off;
            double ret = 0, tw = 0;
            const double *ch = st + (sp - n);
            for(int i = 0; i < n; ++i) {
                double w = evt(ob[i]);
                ret += w * ch[i] * P.norms[in[i]];
                tw += w;
            }
            ret /= tw;
            sp -= n;
            st[sp++] = ret;
        } break;
        case GF_OP_MAPPED: { // combine/MappedPdf.cu:6-38: child 0 picks which of the other children is the value
            int n = nd.nchild;
            const double *ch = st + (sp - n);
            int target = (int)floor(0.5 + ch[0]);
            target = target < 0 ? 0 : (target > n - 2 ? n - 2 : target); // upstream walks off its tables for an index out of range
            double ret = ch[1 + target] * P.norms[in[1 + target]];
            sp -= n;
            st[sp++] = ret;
        } break;
        case GF_OP_COMPOSITE_ARG: if constexpr((HAS & GF_HAS_COMPOSITE) != 0) { // the core's value stays on the stack
            evt.val = st[sp - 1];
            evt.obs = nd.nobs > 0 ? P.ints[nd.obs_off] : -1;
        } break;
        case GF_OP_COMPOSITE: if constexpr((HAS & GF_HAS_COMPOSITE) != 0) { // [core value, shell value] -> shell value
            evt.obs = -1;
            st[sp - 2] = st[sp - 1];
            sp -= 1;
        } break;
        case GF_OP_CONVOLVE: if constexpr((HAS & GF_HAS_CONV) != 0) { // combine/ConvolutionPdf.cu:27-75 via the per-offset table (gf_conv.cu)
            // ints = {aux slot, model root, resolution root}; consts = lo, hi, step
            const double *c = P.consts + nd.con_off;
            const int32_t *ob = P.ints + nd.obs_off;
            double x0 = evt(ob[0]);
            double lo = c[0], step = c[2];
            double lbo;
            if(nd.ncon >= 4)
                lbo = c[3]; // the host class stores lo / step - floor(lo / step): the same two operations, once instead of per event
            else {
                lbo = lo / step;
                lbo -= floor(lbo);
            }
            int k = (int)floor(x0 / step - lbo);
            const double *tab = static_cast<const double *>(P.aux[in[0]]);
            long pk = P.aux_ld[in[0]];
            int t = k - (int)(pk >> 32), nk = (int)(pk & 0xffffffffL);
            t = t < 0 ? 0 : (t >= nk ? nk - 1 : t);
            double ret = __ldg(tab + t);
            ret *= P.norms[P.nodes[in[1]].norm_slot];
            ret *= P.norms[P.nodes[in[2]].norm_slot];
            st[sp++] = ret;
        } break;
        case GF_OP_AMP3: if constexpr((HAS & GF_HAS_AMP3) != 0) { // physics/Amp3Body.cu:60-122 ; child = efficiency ; ints = {aux slot, nRes, ...}
            const double *c = P.consts + nd.con_off;
            const int32_t *ob = P.ints + nd.obs_off;
            double eff = st[--sp];
            double m12 = evt(ob[0]), m13 = evt(ob[1]);
            double ret = 0;
            if(gf_in_dalitz(m12, m13, c[0], c[1], c[2], c[3])) {
                long e = evt.event_number(ob[2]);
                int nres = in[1];
                const double2 *w = static_cast<const double2 *>(P.aux[in[0]]) + e;
                long ld = P.aux_ld[in[0]];
                double re = 0, im = 0;
#pragma unroll 4
                for(int i = 0; i < nres; ++i) {
                    double2 me = __ldg(w + (size_t)i * ld);
                    double ar = p[2 * i], ai = p[2 * i + 1];
                    re += ar * me.x - ai * me.y;
                    im += ar * me.y + ai * me.x;
                }
                ret = (re * re + im * im) * eff;
            }
            st[sp++] = ret;
        } break;
        case GF_OP_AMP4: if constexpr((HAS & GF_HAS_AMP4) != 0) { // physics/Amp4Body.cu:53-90 ; ints = {aux slot, nLS, nSF, nAmp, td, ...}
            const int32_t *ob = P.ints + nd.obs_off;
            double eff = st[--sp];
            long e = evt.event_number(ob[5]);
            int namp = in[3];
            const double2 *w = static_cast<const double2 *>(P.aux[in[0]]) + e;
            long ld = P.aux_ld[in[0]];
            double re = 0, im = 0;
#pragma unroll 4
            for(int i = 0; i < namp; ++i) {
                double2 me = __ldg(w + (size_t)i * ld);
                double ar = p[2 * i], ai = p[2 * i + 1];
                re += ar * me.x - ai * me.y;
                im += ar * me.y + ai * me.x;
            }
            st[sp++] = (re * re + im * im) * eff;
        } break;
        case GF_OP_AMP4_TD: if constexpr((HAS & GF_HAS_AMP4_TD) != 0) { // physics/Amp4Body_TD.cu:167-261 ; ints[4] = offset of {resolution record, flags[nAmp]}
            const int32_t *ob = P.ints + nd.obs_off;
            double eff = st[--sp];
            long e = evt.event_number(ob[5]);
            const int32_t *td = in + in[4];
            const double2 *w = static_cast<const double2 *>(P.aux[in[0]]) + e;
            long ld = P.aux_ld[in[0]];
            const double dtime = evt(ob[6]), sigma = evt(ob[7]); // requested before the amplitude walk, used after it
            double Are, Aim, Bre, Bim;
            gf_amp4td_AB<(HAS == GF_HAS_AMP4_TD) ? GF_TD_CH : ((HAS == (GF_HAS_AMP4_TD | GF_HAS_3GAUSS)) ? 4 : 1)>(p, td + 1, in[3], [&](int i) { return __ldg(w + (size_t)i * ld); }, Are, Aim,
                                                          Bre, Bim);
            double nA = Are * Are + Aim * Aim, nB = Bre * Bre + Bim * Bim;
            double t3re = Are * Bre + Aim * Bim, t3im = Aim * Bre - Are * Bim; // A conj(B)
            const gf_node &rn = P.nodes[td[0]];
            double ret = gf_tres<(HAS & GF_HAS_3GAUSS) != 0>(rn.op, P.params + rn.par_off, nA + nB, nA - nB, t3re, t3im, p[0], dtime, p[1],
                                                             p[2], sigma);
            st[sp++] = ret * eff;
        } break;
        case GF_OP_AMP3_TD: if constexpr((HAS & GF_HAS_AMP3_TD) != 0) { // physics/Amp3Body_TD.cu:40-256
            const double *c = P.consts + nd.con_off;
            const int32_t *ob = P.ints + nd.obs_off;
            double eff = st[--sp];
            double m12 = evt(ob[2]), m13 = evt(ob[3]);
            double ret = 0;
            if(gf_in_dalitz(m12, m13, c[0], c[1], c[2], c[3])) {
                long e = evt.event_number(ob[4]);
                int nres = in[1];
                const double2 *w = static_cast<const double2 *>(P.aux[in[0]]) + e;
                long ld = P.aux_ld[in[0]];
                double Ar = 0, Ai = 0, Br = 0, Bi = 0;
#pragma unroll 2
                for(int i = 0; i < nres; ++i) {
                    double2 a = __ldg(w + (size_t)(2 * i) * ld), b = __ldg(w + (size_t)(2 * i + 1) * ld);
                    double cr = p[5 + 2 * i], ci = p[6 + 2 * i];
                    Ar += a.x * cr - a.y * ci; Ai += a.x * ci + a.y * cr; // ai * amp
                    Br += b.x * cr - b.y * ci; Bi += b.x * ci + b.y * cr;
                }
                double tau = p[0], x = p[1], y = p[2];
                int tag = nd.nobs > 6 ? (int)evt(ob[6]) : 0; // D0 / D0bar: x, y shifted by +- (deltax, deltay)
                if(tag == 1) { x += p[3]; y += p[4]; }
                else if(tag == -1) { x -= p[3]; y -= p[4]; }
                double time = evt(ob[0]), sigma = evt(ob[1]);
                double term1 = (Ar * Ar + Ai * Ai) + (Br * Br + Bi * Bi), term2 = (Ar * Ar + Ai * Ai) - (Br * Br + Bi * Bi);
                double abr = Ar * Br + Ai * Bi, abi = Ai * Br - Ar * Bi; // A conj(B)
                const gf_node &rn = P.nodes[in[2 + nres]];
                const double *rp = P.params + rn.par_off;
                ret = gf_tres<(HAS & GF_HAS_3GAUSS) != 0>(rn.op, rp, term1, term2, abr, abi, tau, time, x, y, sigma);
                double mistag = nd.nobs > 5 ? evt(ob[5]) : 0.0;
                if(mistag > 0) { // the wrongly tagged fraction: A <-> B at fixed world-average mixing (Amp3Body_TD.cu:226-245)
                    const double xfix = 0.0039, yfix = 0.0065, taufix = 0.4101;
                    ret *= (1 - 2 * mistag);
                    ret += mistag * gf_tres<(HAS & GF_HAS_3GAUSS) != 0>(rn.op, rp, term1, -term2, abr, -abi, taufix, time, xfix, yfix, sigma);
                    ret += mistag * gf_tres<(HAS & GF_HAS_3GAUSS) != 0>(rn.op, rp, term1, term2, abr, abi, taufix, time, xfix, yfix, sigma);
                }
                ret *= eff;
            }
            st[sp++] = ret;
        } break;
        default: st[sp++] = 0.0; break;
        }
    }
    return st[sp - 1];
}

// metrics, MetricPointer.cu:8-58 ; binned metrics read (content, volume|error) from the last two columns
template <class Evt>
__device__ inline double gf_metric(int metric, double raw, double norm, const Evt &evt, const gf_dev_view &V,
                                   int n_cols) {
    switch(metric) {
    case GF_METRIC_EVAL: return raw;
    case GF_METRIC_NLL: raw *= norm; return raw > 0.0 ? -log(raw) : 0.0;
    case GF_METRIC_PROB: return raw * norm;
    case GF_METRIC_BINAVG: {
        raw *= norm;
        raw *= evt(n_cols - 1);
        if(raw > 0) {
            double expEvents = V.total_events * raw;
            return expEvents - evt(n_cols - 2) * log(expEvents);
        }
        return 0;
    }
    case GF_METRIC_BINERR: {
        raw -= evt(n_cols - 2);
        raw /= evt(n_cols - 1);
        return raw * raw;
    }
    case GF_METRIC_CHISQ: {
        raw *= norm;
        raw *= evt(n_cols - 1);
        double c0 = evt(n_cols - 2);
        double d = raw * V.total_events - c0;
        return d * d / (c0 > 1 ? c0 : 1);
    }
    default: return raw;
    }
}
