// gf_kmatrix.cuh — the pi pi S-wave K-matrix amplitude (arXiv:0804.2089, as in AmpGen), shared by the three-body resonance
// (src/PDFs/physics/resonances/kMatrix.cu:19-121) and the four-body lineshape (lineshapes/kMatrix.cu:18-119), which are the
// same function of s. Helpers: src/PDFs/physics/kMatrixUtils.cu:13-186 (two-body / four-pion phase space, LU inverse without
// pivoting, propagator (1 - i adler K rho)^-1). Five channels (pi pi, K K, 4 pi, eta eta, eta eta') and five poles.
// Parameter block p[59], in the order the constructors register them: sA0, sA, s0_prod, s0_scatt, fscat[5], beta (re, im) x 5,
// f_prod (re, im) x 5, then per pole its five couplings followed by its mass.
#pragma once
#define GF_KM_N 5
struct gf_kc { double re, im; };
__device__ __forceinline__ gf_kc gf_kcm(double r, double i) { return gf_kc{r, i}; }
__device__ __forceinline__ gf_kc operator+(gf_kc a, gf_kc b) { return gf_kcm(a.re + b.re, a.im + b.im); }
__device__ __forceinline__ gf_kc operator-(gf_kc a, gf_kc b) { return gf_kcm(a.re - b.re, a.im - b.im); }
__device__ __forceinline__ gf_kc operator*(gf_kc a, gf_kc b) { return gf_kcm(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
__device__ __forceinline__ gf_kc operator*(gf_kc a, double s) { return gf_kcm(a.re * s, a.im * s); }
__device__ __forceinline__ gf_kc operator/(gf_kc a, double s) { return gf_kcm(a.re / s, a.im / s); }
__device__ __forceinline__ gf_kc operator/(gf_kc a, gf_kc b) { // as the other complex quotients of this library (gf_amp4.cuh)
    const double d = b.re * b.re + b.im * b.im;
    return gf_kcm((a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d);
}
#define GF_KM_MPI 0.139570
#define GF_KM_MK 0.493677
#define GF_KM_META 0.547862
#define GF_KM_METAP 0.96778
__device__ inline gf_kc gf_km_phsp2(double s, double m0, double m1) { // kMatrixUtils.cu:13-18
    const double v = 1. - (m0 + m1) * (m0 + m1) / s;
    return v > 0 ? gf_kcm(sqrt(v), 0) : gf_kcm(0, sqrt(-v));
}
__device__ inline gf_kc gf_km_phsp4pi(double s) { // kMatrixUtils.cu:20-35 (below threshold the polynomial itself is returned)
    if(s > 1) return gf_km_phsp2(s, 2 * GF_KM_MPI, 2 * GF_KM_MPI);
    const double v = 0.00051 + -0.01933 * s + 0.13851 * s * s + -0.20840 * s * s * s + -0.29744 * s * s * s * s + 0.13655 * s * s * s * s * s
                     + 1.07885 * s * s * s * s * s * s;
    return v >= 0 ? gf_kcm(v, 0) : gf_kcm(0, v);
}
// out of line: its 5x5 complex work matrices live in its own frame and do not set the register budget of the kernels that can reach it
static __device__ __noinline__ gf_kc gf_kmatrix(double s, const double *p) {
    int idx = 0;
    const double sA0 = p[idx++], sA = p[idx++], s0_prod = p[idx++], s0_scatt = p[idx++];
    const double *fscat = p + idx;
    idx += GF_KM_N;
    const double *beta = p + idx; // (re, im) pairs
    idx += 2 * GF_KM_N;
    const double *fprod = p + idx;
    idx += 2 * GF_KM_N;
    const double *poles = p + idx; // pole i: couplings[i][0..4], pmasses[i]
    auto coupling = [&](int i, int j) { return poles[6 * i + j]; };
    auto pmass = [&](int i) { return poles[6 * i + 5]; };
    double K[GF_KM_N][GF_KM_N];
#pragma unroll
    for(int i = 0; i < GF_KM_N; i++)
#pragma unroll
        for(int j = 0; j < GF_KM_N; j++) {
            double v = 0;
#pragma unroll
            for(int k = 0; k < GF_KM_N; k++) v += coupling(k, i) * coupling(k, j) / (pmass(k) * pmass(k) - s);
            if(i == 0 || j == 0) v += fscat[i + j] * (1 - s0_scatt) / (s - s0_scatt); // scattering term
            K[i][j] = v;
        }
    const double adler = (1. - sA0) * (s - sA * GF_KM_MPI * GF_KM_MPI / 2) / (s - sA0);
    gf_kc rho[GF_KM_N];
    rho[0] = gf_km_phsp2(s, GF_KM_MPI, GF_KM_MPI);
    rho[1] = gf_km_phsp2(s, GF_KM_MK, GF_KM_MK);
    rho[2] = gf_km_phsp4pi(s);
    rho[3] = gf_km_phsp2(s, GF_KM_META, GF_KM_META);
    rho[4] = gf_km_phsp2(s, GF_KM_META, GF_KM_METAP);
    // T = 1 - i adler K rho, inverted by LU decomposition without pivoting (kMatrixUtils.cu:100-166); only row 0 of the inverse
    // enters the amplitude, but the whole algorithm is kept so that the rounding matches the reference's
    gf_kc A[GF_KM_N][GF_KM_N], U[GF_KM_N][GF_KM_N], L[GF_KM_N][GF_KM_N], Li[GF_KM_N][GF_KM_N], F[GF_KM_N][GF_KM_N];
#pragma unroll
    for(int i = 0; i < GF_KM_N; i++)
#pragma unroll
        for(int j = 0; j < GF_KM_N; j++) {
            A[i][j] = gf_kcm(i == j ? 1. : 0., 0.) - gf_kcm(0, adler) * K[i][j] * rho[j];
            U[i][j] = L[i][j] = Li[i][j] = F[i][j] = gf_kcm(0, 0);
        }
#pragma unroll
    for(int i = 0; i < GF_KM_N; i++) {
#pragma unroll
        for(int k = i; k < GF_KM_N; k++) {
            gf_kc sum = gf_kcm(0, 0);
            for(int j = 0; j < i; j++) sum = sum + L[i][j] * U[j][k];
            U[i][k] = A[i][k] - sum;
        }
#pragma unroll
        for(int k = i; k < GF_KM_N; k++) {
            if(i == k)
                L[i][i] = gf_kcm(1, 0);
            else {
                gf_kc sum = gf_kcm(0, 0);
                for(int j = 0; j < i; j++) sum = sum + L[k][j] * U[j][i];
                L[k][i] = (A[k][i] - sum) / U[i][i];
            }
        }
    }
#pragma unroll
    for(int col = 0; col < GF_KM_N; col++)
#pragma unroll
        for(int row = 0; row < GF_KM_N; row++) {
            gf_kc sum = gf_kcm(0, 0);
#pragma unroll
            for(int i = 0; i < GF_KM_N; i++)
                if(i != row) sum = sum + L[row][i] * Li[i][col];
            Li[row][col] = (gf_kcm(row == col ? 1. : 0., 0) - sum) / L[row][row];
        }
#pragma unroll
    for(int col = 0; col < GF_KM_N; col++)
#pragma unroll
        for(int row = GF_KM_N - 1; row >= 0; row--) {
            gf_kc sum = gf_kcm(0, 0);
#pragma unroll
            for(int i = 0; i < GF_KM_N; i++)
                if(i != row) sum = sum + U[row][i] * F[i][col];
            F[row][col] = (Li[row][col] - sum) / U[row][row];
        }
    gf_kc ret = gf_kcm(0, 0);
#pragma unroll
    for(int t = 0; t < GF_KM_N; t++) {
        gf_kc M = gf_kcm(0, 0);
#pragma unroll
        for(int i = 0; i < GF_KM_N; i++) M = M + F[0][i] * coupling(t, i);
        const gf_kc pole = M / (pmass(t) * pmass(t) - s);
        ret = ret + gf_kcm(beta[2 * t], beta[2 * t + 1]) * pole;
        const gf_kc prod = F[0][t] * (1 - s0_prod) / (s - s0_prod);
        ret = ret + gf_kcm(fprod[2 * t], fprod[2 * t + 1]) * prod;
    }
    return ret;
}
