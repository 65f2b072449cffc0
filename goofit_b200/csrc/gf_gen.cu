// gf_gen.cu — phase-space generation on the device (toy generation: Amp3Body / Amp4Body / Amp4Body_TD::GenerateSig and the MC
// normalisation sample of the four-body amplitudes).
// The reference delegates this to MCBooster (mcbooster::PhaseSpace::Generate + Unweight and EvaluateArray<Dim5>:
// Amp4Body.cu:243-289,560-677, Amp4Body_TD.cu:825-1018, detail/Dim5.h:19-111), a submodule that is not part of the reference
// checkout; what is restated here is the published GENBOD algorithm (F. James, CERN 68-15: sorted uniforms -> intermediate
// masses -> sequential two-body decays, weight = product of the break-up momenta) on a COUNTER-BASED generator: event i of a
// seed is the same whatever the batch, the rank or the thread that produces it, so a sample can be generated in place on each
// GPU (or regenerated) without a host loop. One thread per event; no state is carried between events.
#include "gf_internal.cuh"
#include <mutex>
#include <string>

struct gf_v4 { double e, x, y, z; };
__device__ __forceinline__ gf_v4 operator+(gf_v4 a, gf_v4 b) { return {a.e + b.e, a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ double gf_mink(gf_v4 a, gf_v4 b) { return a.e * b.e - a.x * b.x - a.y * b.y - a.z * b.z; }
__device__ __forceinline__ double gf_dot3(gf_v4 a, gf_v4 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ gf_v4 gf_unit3(gf_v4 a) {
    const double r = 1.0 / sqrt(gf_dot3(a, a));
    return {a.e * r, a.x * r, a.y * r, a.z * r};
}
__device__ __forceinline__ gf_v4 gf_cross3(gf_v4 a, gf_v4 b) { return {0.0, a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// helicity angle and the angle between the two decay planes (detail/Dim5.h:22-92)
__device__ inline double gf_coshel(gf_v4 p, gf_v4 q, gf_v4 d) {
    const double pd = gf_mink(p, d), pq = gf_mink(p, q), qd = gf_mink(q, d);
    const double mp2 = gf_mink(p, p), mq2 = gf_mink(q, q), md2 = gf_mink(d, d);
    return (pd * mq2 - pq * qd) / sqrt((pq * pq - mq2 * mp2) * (qd * qd - mq2 * md2));
}
__device__ inline double gf_plane_angle(gf_v4 d1, gf_v4 d2, gf_v4 h1, gf_v4 h2) {
    const gf_v4 d1n = gf_unit3(d1), d2n = gf_unit3(d2), h1n = gf_unit3(h1), h2n = gf_unit3(h2);
    const gf_v4 h12 = gf_unit3(h1 + h2);
    const gf_v4 n1 = gf_unit3(gf_cross3(d1n, d2n)), n2 = gf_unit3(gf_cross3(h1n, h2n));
    const gf_v4 n3 = gf_cross3(n1, n2);
    const double cp = gf_dot3(n1, n2), sp = gf_dot3(n3, h12);
    double phi2 = acos(cp);
    if(sp < 0) phi2 *= -1;
    return phi2;
}
__host__ __device__ inline unsigned long long gf_mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// uniform in (0, 1): stream k of event `event` of `seed` (the host layer's uniform01 draws the same numbers)
__host__ __device__ inline double gf_uniform01(unsigned long long seed, unsigned long long event, unsigned long long k) {
    const unsigned long long h = gf_mix64(gf_mix64(seed ^ 0xD1B54A32D192ED03ull) + event * 0x2545F4914F6CDD1Dull + k);
    return ((h >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}
__host__ __device__ inline double gf_pdk(double a, double b, double c) {
    const double x = (a - b - c) * (a + b + c) * (a + b - c) * (a - b + c);
    return sqrt(x > 0 ? x : 0) / (2 * a);
}
__device__ __forceinline__ void gf_boost_y(gf_v4 &p, double by) { // boost along y with velocity by
    const double b2 = by * by;
    if(b2 <= 0) return;
    const double gamma = 1.0 / sqrt(1.0 - b2), bp = by * p.y, g2 = (gamma - 1.0) / b2;
    p.y += g2 * bp * by + gamma * by * p.e;
    p.e = gamma * (p.e + bp);
}

struct gf_gen_args {
    double M, m[4], wtmax;
    int nd;
    unsigned long long seed, offset;
    long n;
};
// ND daughters of one event into p[]; returns the GENBOD weight relative to the largest possible one
template <int ND>
__device__ __forceinline__ double gf_genbod_event(const gf_gen_args &G, unsigned long long ev, gf_v4 (&p)[4]) {
    unsigned long long ctr = 0;
    double rno[4] = {0, 0, 0, 0}, inv[4], pd[3], sum = 0, tecm = G.M;
#pragma unroll
    for(int k = 0; k < ND; k++) tecm -= G.m[k];
#pragma unroll
    for(int k = 1; k < ND - 1; k++) rno[k] = gf_uniform01(G.seed, ev, ctr++);
    if(ND == 4 && rno[1] > rno[2]) {
        const double t = rno[1];
        rno[1] = rno[2];
        rno[2] = t;
    }
    rno[ND - 1] = 1;
#pragma unroll
    for(int k = 0; k < ND; k++) {
        sum += G.m[k];
        inv[k] = rno[k] * tecm + sum;
    }
    double wt = 1.0 / G.wtmax;
#pragma unroll
    for(int k = 0; k < ND - 1; k++) {
        pd[k] = gf_pdk(inv[k + 1], inv[k], G.m[k + 1]);
        wt *= pd[k];
    }
    p[0] = {sqrt(pd[0] * pd[0] + G.m[0] * G.m[0]), 0, pd[0], 0};
#pragma unroll
    for(int d = 1; d < ND; d++) { // sequential two-body decays, each rotated at random and boosted into its parent's frame
        p[d] = {sqrt(pd[d - 1] * pd[d - 1] + G.m[d] * G.m[d]), 0, -pd[d - 1], 0};
        const double cZ = 2 * gf_uniform01(G.seed, ev, ctr++) - 1, sZ = sqrt(1 - cZ * cZ);
        const double ang = 2 * M_PI * gf_uniform01(G.seed, ev, ctr++);
        double sY, cY;
        sincos(ang, &sY, &cY);
#pragma unroll
        for(int j = 0; j <= d; j++) {
            double x = p[j].x;
            const double y = p[j].y;
            p[j].x = cZ * x - sZ * y;
            p[j].y = sZ * x + cZ * y;
            x = p[j].x;
            const double z = p[j].z;
            p[j].x = cY * x - sY * z;
            p[j].z = sY * x + cY * z;
        }
        if(d == ND - 1) break;
        const double beta = pd[d] / sqrt(pd[d] * pd[d] + inv[d] * inv[d]);
#pragma unroll
        for(int j = 0; j <= d; j++) gf_boost_y(p[j], beta);
    }
    return wt;
}
__device__ __forceinline__ void gf_dim5(const gf_v4 (&p)[4], double (&out)[5]) { // detail/Dim5.h:94-110
    const gf_v4 p12 = p[0] + p[1], p34 = p[2] + p[3], pM = p12 + p34;
    const double m12 = gf_mink(p12, p12), m34 = gf_mink(p34, p34);
    out[0] = m12 > 0 ? sqrt(m12) : 0.0;
    out[1] = m34 > 0 ? sqrt(m34) : 0.0;
    out[2] = gf_coshel(pM, p12, p[0]);
    out[3] = gf_coshel(pM, p34, p[2]);
    out[4] = gf_plane_angle(p[0], p[1], p[2], p[3]);
}

// particles: [nd][n][4] = (E, px, py, pz) or null; weights[n]; dim5: [5][n] or null (four daughters only)
template <int ND>
__global__ void __launch_bounds__(GF_THREADS) k_genbod(const gf_gen_args G, double *particles, double *weights, double *dim5) {
    const long stride = (long)gridDim.x * blockDim.x;
    for(long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < G.n; i += stride) {
        gf_v4 p[4];
        const double wt = gf_genbod_event<ND>(G, G.offset + (unsigned long long)i, p);
        if(particles) {
#pragma unroll
            for(int d = 0; d < ND; d++) {
                double4 *o = reinterpret_cast<double4 *>(particles + ((size_t)d * G.n + i) * 4);
                *o = make_double4(p[d].e, p[d].x, p[d].y, p[d].z);
            }
        }
        weights[i] = wt;
        if(ND == 4 && dim5) {
            double v[5];
            gf_dim5(p, v);
#pragma unroll
            for(int k = 0; k < 5; k++) dim5[(size_t)k * G.n + i] = v[k];
        }
    }
}

// GENBOD phase space of mother -> 3 or 4 daughters, events [offset, offset + n) of the stream `seed`.
// masses = mother, daughters. Host outputs (each may be null): particles[nd][n][4] = (E, px, py, pz), weights[n] relative to the
// largest possible weight, dim5[5][n] = (m12, m34, cos12, cos34, phi) for four daughters.
extern "C" int gf_genbod(gf_ctx *c, const double *masses, int nd, size_t n, uint64_t seed, uint64_t offset, double *particles, double *weights,
                         double *dim5) {
    if(!c || !masses) return gf_fail("gf_genbod: null argument");
    if(nd != 3 && nd != 4) return gf_fail("gf_genbod: 3 or 4 daughters");
    if(dim5 && nd != 4) return gf_fail("gf_genbod: the five-variable form needs four daughters");
    if(n == 0) return 0;
    cudaSetDevice(c->device);
    gf_gen_args G;
    G.M = masses[0];
    G.nd = nd;
    G.seed = seed;
    G.offset = offset;
    G.n = (long)n;
    double tecm = G.M;
    for(int k = 0; k < 4; k++) {
        G.m[k] = k < nd ? masses[1 + k] : 0.0;
        if(k < nd) tecm -= G.m[k];
    }
    if(!(tecm > 0)) return gf_fail("gf_genbod: not enough energy");
    double emmax = tecm + G.m[0], emmin = 0;
    G.wtmax = 1;
    for(int k = 1; k < nd; k++) {
        emmin += G.m[k - 1];
        emmax += G.m[k];
        G.wtmax *= gf_pdk(emmax, emmin, G.m[k]);
    }
    double *d_p = nullptr, *d_w = nullptr, *d_v = nullptr;
    auto bail = [&](const std::string &what) {
        cudaFree(d_p); cudaFree(d_w); cudaFree(d_v);
        cudaGetLastError();
        return gf_fail("gf_genbod: " + what);
    };
    if(particles && cudaMalloc(&d_p, sizeof(double) * 4 * nd * n) != cudaSuccess) return bail("device allocation failed");
    if(cudaMalloc(&d_w, sizeof(double) * n) != cudaSuccess) return bail("device allocation failed");
    if(dim5 && cudaMalloc(&d_v, sizeof(double) * 5 * n) != cudaSuccess) return bail("device allocation failed");
    const int blocks = gf_grid_blocks(c, (long)n, 8);
    if(nd == 3)
        k_genbod<3><<<blocks, GF_THREADS, 0, c->stream>>>(G, d_p, d_w, nullptr);
    else
        k_genbod<4><<<blocks, GF_THREADS, 0, c->stream>>>(G, d_p, d_w, d_v);
    c->launches++;
    cudaError_t e = cudaGetLastError();
    if(e == cudaSuccess && particles) e = cudaMemcpyAsync(particles, d_p, sizeof(double) * 4 * nd * n, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess && weights) e = cudaMemcpyAsync(weights, d_w, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess && dim5) e = cudaMemcpyAsync(dim5, d_v, sizeof(double) * 5 * n, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if(e != cudaSuccess) return bail(cudaGetErrorString(e));
    cudaFree(d_p); cudaFree(d_w); cudaFree(d_v);
    return 0;
}

// ---- gf_genbod without any allocation -----------------------------------------------------------------------------------------
// gf_genbod stages its device outputs in buffers it cudaMallocs and cudaFrees on every call. Callers that may not allocate
// (a stream under graph capture, a tight serving loop, a harness that forbids it) get the same kernel with that staging area in
// a file-scope array instead: 22 doubles per event (weight, five variables, four four-vectors) for at most
// GF_GENBOD_STATIC_MAX_EVENTS events = 5.5 MiB, held by every GPU that loads the library until the process exits.
// `gf_genbod_scratch` below is the HOST-side stand-in of the device array: host code must never hand it to a kernel. Its device
// address is resolved once per device with cudaGetSymbolAddress (that device current) and cached here.
// Every call shares the array and it keeps the previous call's contents; the kernel overwrites every element the copies read
// back (n weights, 5 n variables, 4 nd n momenta), so nothing needs clearing. Two calls that overlap on the GPU (two contexts =
// two streams on one device, or two host threads) race on it: serialise them.
static __device__ __align__(256) double gf_genbod_scratch[(size_t)22 * GF_GENBOD_STATIC_MAX_EVENTS];
#define GF_GENBOD_MAX_DEVICES 64
static double *g_genbod_scratch_ptr[GF_GENBOD_MAX_DEVICES]; // device address of gf_genbod_scratch on device d, null until resolved
static std::mutex g_genbod_scratch_mu;

extern "C" int gf_genbod_static(gf_ctx *c, const double *masses, int nd, size_t n, uint64_t seed, uint64_t offset, double *particles, double *weights,
                                double *dim5) {
    if(!c || !masses) return gf_fail("gf_genbod_static: null argument");
    if(nd != 3 && nd != 4) return gf_fail("gf_genbod_static: 3 or 4 daughters");
    if(dim5 && nd != 4) return gf_fail("gf_genbod_static: the five-variable form needs four daughters");
    if(n > (size_t)GF_GENBOD_STATIC_MAX_EVENTS)
        return gf_fail("gf_genbod_static: " + std::to_string(n) + " events exceed the static scratch array (at most " +
                       std::to_string(GF_GENBOD_STATIC_MAX_EVENTS) + " per call; split the range by `offset`, or use gf_genbod)");
    if(n == 0) return 0;
    if(c->device < 0 || c->device >= GF_GENBOD_MAX_DEVICES) return gf_fail("gf_genbod_static: device index out of range");
    cudaError_t e = cudaSetDevice(c->device);
    if(e != cudaSuccess) return gf_fail(std::string("gf_genbod_static: cudaSetDevice: ") + cudaGetErrorString(e));
    double *scratch;
    {
        std::lock_guard<std::mutex> lock(g_genbod_scratch_mu);
        if(!g_genbod_scratch_ptr[c->device]) { // set-up, once per device, with that device current
            void *ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, gf_genbod_scratch);
            if(e != cudaSuccess || !ptr) {
                cudaGetLastError();
                return gf_fail(std::string("gf_genbod_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            }
            g_genbod_scratch_ptr[c->device] = static_cast<double *>(ptr);
        }
        scratch = g_genbod_scratch_ptr[c->device];
    }
    gf_gen_args G;
    G.M = masses[0];
    G.nd = nd;
    G.seed = seed;
    G.offset = offset;
    G.n = (long)n;
    double tecm = G.M;
    for(int k = 0; k < 4; k++) {
        G.m[k] = k < nd ? masses[1 + k] : 0.0;
        if(k < nd) tecm -= G.m[k];
    }
    if(!(tecm > 0)) return gf_fail("gf_genbod_static: not enough energy");
    double emmax = tecm + G.m[0], emmin = 0;
    G.wtmax = 1;
    for(int k = 1; k < nd; k++) {
        emmin += G.m[k - 1];
        emmax += G.m[k];
        G.wtmax *= gf_pdk(emmax, emmin, G.m[k]);
    }
    // fixed partition of the array (offsets by the maximum, so every part is 256-byte aligned whatever n is)
    double *d_w = scratch;
    double *d_v = dim5 ? scratch + (size_t)GF_GENBOD_STATIC_MAX_EVENTS : nullptr;
    double *d_p = particles ? scratch + (size_t)6 * GF_GENBOD_STATIC_MAX_EVENTS : nullptr;
    const int blocks = gf_grid_blocks(c, (long)n, 8);
    if(nd == 3)
        k_genbod<3><<<blocks, GF_THREADS, 0, c->stream>>>(G, d_p, d_w, nullptr);
    else
        k_genbod<4><<<blocks, GF_THREADS, 0, c->stream>>>(G, d_p, d_w, d_v);
    c->launches++;
    e = cudaGetLastError();
    if(e == cudaSuccess && particles) e = cudaMemcpyAsync(particles, d_p, sizeof(double) * 4 * nd * n, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess && weights) e = cudaMemcpyAsync(weights, d_w, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess && dim5) e = cudaMemcpyAsync(dim5, d_v, sizeof(double) * 5 * n, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if(e != cudaSuccess) {
        cudaGetLastError();
        return gf_fail(std::string("gf_genbod_static: ") + cudaGetErrorString(e));
    }
    return 0;
}
