// scg_inst.cuh -- kernel launchers for one (env kind, precision); each scg_k_*.cu includes this
// and instantiates it once, so the kinds compile in parallel.
#pragma once
#include "scg_host.cuh"
#include "scg_step.cuh"

namespace scg {

template <int KIND, typename real>
int launch_reset(scg_env* e, const ResetInject<real>& inj, const uint8_t* mask, cudaStream_t st) {
    const Hot<real>& h = hot_of<real>(e);
    const Buf<real> b = typed_buf<real>(e->buf);
    const unsigned blocks = unsigned((b.n + ENV_BLOCK - 1) / ENV_BLOCK);
    const Cold<real>* cold = static_cast<const Cold<real>*>(e->cold_dev);
    const int att = e->cfg.attitude_mode;
    if constexpr (KIND == SCG_QUAD_3D || KIND == SCG_QUAD_2D) {
        if (att == SCG_ATT_ROUNDTRIP) k_reset<KIND, real, SCG_ATT_ROUNDTRIP><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
        else if (att == SCG_ATT_FAST_INTRINSIC && sizeof(real) == 4 && KIND == SCG_QUAD_3D)
            k_reset<KIND, real, (sizeof(real) == 4 && KIND == SCG_QUAD_3D) ? SCG_ATT_FAST_INTRINSIC : SCG_ATT_FAST><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
        else k_reset<KIND, real, SCG_ATT_FAST><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
    } else {
        k_reset<KIND, real, SCG_ATT_FAST><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
    }
    e->launches += 1;
    SCG_CU(cudaGetLastError());
    return SCG_OK;
}

template <int KIND, typename real, int ATT>
static void launch_rollout_att(scg_env* e, const Hot<real>& h, const Cold<real>* cold, const Buf<real>& b, const Roll<real>& ro,
                               unsigned blocks, cudaStream_t st) {
    constexpr bool F32 = sizeof(real) == 4;
    if (e->lean && e->plain && !ro.use_pid && F32) {
        if (KIND == SCG_QUAD_3D && !ro.step_outputs && (h.cfg & CFG_SPEC_MASK) == CFG_SPEC_VAL)
            k_rollout<KIND, real, ATT, true, F32, (F32 && KIND == SCG_QUAD_3D)><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro);
        else k_rollout<KIND, real, ATT, true, F32><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro);
    }
    else if (e->lean) k_rollout<KIND, real, ATT, true, false><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro);
    else k_rollout<KIND, real, ATT, false, false><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro);
}

template <int KIND, typename real>
int launch_rollout(scg_env* e, const Roll<real>& ro, cudaStream_t st) {
    const Hot<real>& h = hot_of<real>(e);
    const Buf<real> b = typed_buf<real>(e->buf);
    const unsigned blocks = unsigned((b.n + ENV_BLOCK - 1) / ENV_BLOCK);
    const Cold<real>* cold = static_cast<const Cold<real>*>(e->cold_dev);
    const int att = e->cfg.attitude_mode;
    if constexpr (KIND == SCG_QUAD_3D || KIND == SCG_QUAD_2D) {
        if (att == SCG_ATT_ROUNDTRIP) k_rollout<KIND, real, SCG_ATT_ROUNDTRIP, false, false><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro);
        else if (att == SCG_ATT_FAST_INTRINSIC && sizeof(real) == 4 && KIND == SCG_QUAD_3D)
            launch_rollout_att<KIND, real, (sizeof(real) == 4 && KIND == SCG_QUAD_3D) ? SCG_ATT_FAST_INTRINSIC : SCG_ATT_FAST>(e, h, cold, b, ro, blocks, st);
        else launch_rollout_att<KIND, real, SCG_ATT_FAST>(e, h, cold, b, ro, blocks, st);
    } else {
        launch_rollout_att<KIND, real, SCG_ATT_FAST>(e, h, cold, b, ro, blocks, st);
    }
    e->launches += 1;
    SCG_CU(cudaGetLastError());
    return SCG_OK;
}

// ConstraintList.get_values / is_violated on caller-given tensors (scg_constraints_eval)
template <int KIND, typename real>
__global__ void __launch_bounds__(128)
k_constraints(const __grid_constant__ Hot<real> h, const Cold<real>* __restrict__ cold, const real* __restrict__ state,
              const real* __restrict__ action, int64_t n, real* __restrict__ out_values, uint8_t* __restrict__ out_flags) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    real st[NX], raw[NU];
#pragma unroll
    for (int d = 0; d < NX; ++d) st[d] = state[int64_t(d) * n + i];
#pragma unroll
    for (int j = 0; j < NU; ++j) raw[j] = action ? action[i * NU + j] : real(0);
    bool violated, almost;
    eval_constraints<KIND, real>(h, cold, st, raw, action == nullptr, out_values, n, i, violated, almost, h.cfg | CFG_FULL_STATE);
    if (out_flags) out_flags[i] = (violated ? 1 : 0) | (almost ? 2 : 0);
}

template <int KIND, typename real>
int launch_constraints(scg_env* e, const real* state, const real* action, int64_t n, real* out_values, uint8_t* out_flags,
                       cudaStream_t st) {
    const Hot<real>& h = hot_of<real>(e);
    const Cold<real>* cold = static_cast<const Cold<real>*>(e->cold_dev);
    const unsigned blocks = unsigned((n + 127) / 128);
    k_constraints<KIND, real><<<blocks, 128, 0, st>>>(h, cold, state, action, n, out_values, out_flags);
    e->launches += 1;
    SCG_CU(cudaGetLastError());
    return SCG_OK;
}

#define SCG_INSTANTIATE(KIND, real)                                                                                   \
    template int launch_reset<KIND, real>(scg_env*, const ResetInject<real>&, const uint8_t*, cudaStream_t);          \
    template int launch_rollout<KIND, real>(scg_env*, const Roll<real>&, cudaStream_t);                               \
    template int launch_constraints<KIND, real>(scg_env*, const real*, const real*, int64_t, real*, uint8_t*, cudaStream_t);

}  // namespace scg
