// placeholder until the Amp4Body kernels land
#include "gf_internal.cuh"
struct gf_amp4_state { int node; };
void gf_amp4_free(gf_ctx *c) { for(auto *s : c->amp4) delete s; c->amp4.clear(); }
