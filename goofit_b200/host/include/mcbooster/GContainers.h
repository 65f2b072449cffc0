// mcbooster/GContainers.h — the host containers of MCBooster's interface as std::vector (upstream: thrust::host_vector, same members
// for what user code does with them: size, operator[], begin / end). See GTypes.h for what this directory is.
#pragma once
#include <vector>

#include <mcbooster/GTypes.h>
#include <mcbooster/Vector4R.h>

namespace mcbooster {
typedef std::vector<GReal_t> RealVector_h;
typedef std::vector<GBool_t> BoolVector_h;
typedef std::vector<Vector4R> Particles_h;
typedef std::vector<Particles_h *> ParticlesSet_h; // one container per daughter; the receiver owns (and deletes) them, as upstream
typedef std::vector<RealVector_h *> VariableSet_h; // one container per variable; likewise
} // namespace mcbooster
