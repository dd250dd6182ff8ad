// scg_cartpole.cuh -- cart-pole env step.  Filled in below.
#pragma once
#include "scg_step.cuh"
namespace scg {
template <typename real, int MODE>
__global__ void __launch_bounds__(128)
k_cartpole(const __grid_constant__ Hot<real> h, const Cold<real>* __restrict__ cold, const Buf<real> b,
           const Roll<real> ro, const ResetInject<real> inj, const uint8_t* __restrict__ mask) {}
}  // namespace scg
