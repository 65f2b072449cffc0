// TEST INFRASTRUCTURE (oracle): apply an IFunctionArray functor per event (host loop; order-independent).
// Call sites: /root/reference/src/PDFs/physics/Amp4Body.cu:284-285, Amp3Body.cu:472.
#pragma once
#include <mcbooster/GContainers.h>
#include <mcbooster/GFunctional.h>
namespace mcbooster {
template <typename F> void EvaluateArray(F &fn, ParticlesSet_d &pset, VariableSet_d &vars) {
    size_t n = vars.empty() ? 0 : vars[0]->size();
    int np = static_cast<int>(pset.size());
    int nv = static_cast<int>(vars.size());
    std::vector<const Vector4R *> pp(np);
    for(int k = 0; k < np; k++) pp[k] = thrust::raw_pointer_cast(pset[k]->data());
    std::vector<GReal_t *> vv(nv);
    for(int k = 0; k < nv; k++) vv[k] = thrust::raw_pointer_cast(vars[k]->data());
#pragma omp parallel for
    for(long i = 0; i < static_cast<long>(n); i++) {
        Vector4R loc[8];
        Vector4R *ptrs[8];
        GReal_t out[16];
        for(int k = 0; k < np; k++) { loc[k] = pp[k][i]; ptrs[k] = &loc[k]; }
        F f = fn;
        f(np, ptrs, out);
        for(int k = 0; k < nv; k++) vv[k][i] = out[k];
    }
}
} // namespace mcbooster
