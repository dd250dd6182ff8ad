// kernels of one (env kind, precision): see scg_inst.cuh
#include "scg_inst.cuh"
namespace scg { SCG_INSTANTIATE(SCG_QUAD_1D, float) }
