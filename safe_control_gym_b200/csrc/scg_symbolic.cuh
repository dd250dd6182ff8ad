// scg_symbolic.cuh -- batched SymbolicModel evaluation.  Filled in below.
#pragma once
#include "scg_params.cuh"
namespace scg {
template <typename real>
static int symbolic_launch(const scg_config& cfg, const Hot<real>& h, const Cold<real>* cold, int op, const real* x,
                           const real* u, int64_t n, real dt, real* out0, real* out1, cudaStream_t st) { return -1; }
}  // namespace scg
