// gf_ops.cuh — per-node device math and the post-order stack machine.
// Each case cites the reference device function it reproduces (paths under /root/reference/src/PDFs).
#pragma once

__device__ inline double gf_pow_i(double x, int k) { return pow(x, (double)k); } // std::pow(double,int) promotes

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
        return x * pow(t, power) * exp(slope * t);
    }
    case GF_OP_ARGUS_LOWER: { // ArgusPdf.cu:27-47
        double x = evt(ob[0]), m0 = p[0], slope = p[1], power = p[2];
        double t = x / m0;
        if(t <= 1) return 0;
        t *= t;
        t -= 1;
        return x * pow(t, power) * exp(slope * t);
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

// Post-order evaluation of nodes [first, last]; returns the root's raw value.
template <class Evt>
__device__ inline double gf_eval_tree(const gf_sprog &P, const gf_dev_view &V, int first, int last, const Evt &evt) {
    double st[GF_MAX_STACK];
    int sp = 0;
    for(int k = first; k <= last; ++k) {
        const gf_node &nd = P.nodes[k];
        if(nd.op < GF_OP_ADD) {
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
        case GF_OP_CONVOLVE: { // combine/ConvolutionPdf.cu:27-75 via the per-offset table (gf_conv.cu)
            // ints = {aux slot, model root, resolution root}; consts = lo, hi, step
            const double *c = P.consts + nd.con_off;
            const int32_t *ob = P.ints + nd.obs_off;
            double x0 = evt(ob[0]);
            double lo = c[0], step = c[2];
            double lbo = lo / step;
            lbo -= floor(lbo);
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
        case GF_OP_AMP3: { // physics/Amp3Body.cu:60-122 ; child = efficiency ; ints = {aux slot, nRes, ...}
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
        case GF_OP_AMP4: { // physics/Amp4Body.cu:53-90 ; ints = {aux slot, nLS, nSF, nAmp, td, ...}
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
