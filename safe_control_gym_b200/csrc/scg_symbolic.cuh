// scg_symbolic.cuh -- declaration of the batched SymbolicModel evaluation; the
// kernels live in scg_symbolic.cu.
#pragma once
#include "scg_params.cuh"

namespace scg {
// returns 0, or -1 when the op is not available for this env kind
template <typename real>
int symbolic_launch(const scg_config& cfg, int op, const real* x, const real* u, const real* xr, const real* ur,
                    const double* Q_host, const double* R_host, int64_t n, real dt, real* out0, real* out1, real* out2,
                    cudaStream_t st);
}  // namespace scg
