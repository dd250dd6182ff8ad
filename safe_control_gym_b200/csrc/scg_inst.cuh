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
        else if (att == SCG_ATT_FAST_INTRINSIC && sizeof(real) == 4)
            k_reset<KIND, real, sizeof(real) == 4 ? SCG_ATT_FAST_INTRINSIC : SCG_ATT_FAST><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
        else k_reset<KIND, real, SCG_ATT_FAST><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
    } else {
        k_reset<KIND, real, SCG_ATT_FAST><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, inj, mask);
    }
    e->launches += 1;
    SCG_CU(cudaGetLastError());
    return SCG_OK;
}

// The queue pays when the launch is long and the env blocks fill the resident slots unevenly (at most a few waves):
// a static grid of 1,024 blocks leaves schedulers with three and with four warps, a grid of 1,300 runs two waves.
template <typename K>
static bool launch_queued(scg_env* e, K kernel, bool plain, const int64_t n, int k_steps, int32_t* q_chunk, int32_t* q_env_blocks,
                          int32_t* q_tickets, unsigned* grid) {
    if (e->queue_mode == 0 || k_steps < 64) return false;
    // automatic mode: PLAIN kernels only -- a range of the other kernels starts with the general spare draw (inertia,
    // disturbance offsets, normal draws), which costs more than the balance returns (level2-like config: 20.1 -> 18.1 G)
    if (e->queue_mode < 0 && !plain) return false;
    if (e->n_sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&e->n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return false;
    }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, ENV_BLOCK, 0) != cudaSuccess || occ < 1) return false;
    if (e->queue_blocks_per_sm > 0 && occ > e->queue_blocks_per_sm) occ = e->queue_blocks_per_sm;   // fewer workers than resident slots (SCG_QUEUE_BLOCKS_PER_SM)
    const int64_t resident = int64_t(occ) * e->n_sms, env_blocks = (n + ENV_BLOCK - 1) / ENV_BLOCK;
    if (e->queue_mode < 0 && (env_blocks > 6 * resident || env_blocks * 2 < resident || env_blocks % resident == 0)) return false;   // whole waves: nothing to balance
    int chunk = e->queue_chunk > 0 ? e->queue_chunk : 64;
    if (chunk > k_steps) chunk = k_steps;
    const int64_t chunks = (k_steps + chunk - 1) / chunk, tickets = chunks * env_blocks;
    if (tickets + env_blocks > 0x7ffffff0) return false;   // an item is 1 + env block + env_blocks * chunk in 31 bits
    if (e->q_cap < 2 + tickets) {   // pop counter, push counter, ready list
        if (e->q_mem) cudaFree(e->q_mem);
        e->q_mem = nullptr; e->q_cap = 0;
        if (cudaMalloc(&e->q_mem, sizeof(int32_t) * (2 + tickets)) != cudaSuccess) return false;
        e->q_cap = 2 + tickets;
    }
    *q_chunk = chunk; *q_env_blocks = int32_t(env_blocks); *q_tickets = int32_t(tickets);
    *grid = unsigned(tickets < resident ? tickets : resident);
    return true;
}

template <int KIND, typename real, int ATT, bool LEAN, bool PLAIN, int SPEC, bool DS = false>
static void launch_rollout_v(scg_env* e, const Hot<real>& h, const Cold<real>* cold, const Buf<real>& b, const Roll<real>& ro,
                             unsigned blocks, cudaStream_t st) {
    if constexpr (LEAN && sizeof(real) == 4 && SPEC != 2) {   // (SPEC 2: launches with per-step outputs are never queued)
        Roll<real> rq = ro;
        unsigned grid = 0;
        if (!ro.step_outputs && launch_queued(e, k_rollout<KIND, real, ATT, LEAN, PLAIN, SPEC, true, DS>, PLAIN && !DS, b.n, ro.k_steps, &rq.q_chunk,
                                              &rq.q_env_blocks, &rq.q_tickets, &grid)) {
            rq.q_mem = e->q_mem; rq.q_sleep_ns = e->queue_sleep_ns;
            if (cudaMemsetAsync(e->q_mem, 0, sizeof(int32_t) * (2 + size_t(rq.q_tickets)), st) == cudaSuccess) {
                k_rollout<KIND, real, ATT, LEAN, PLAIN, SPEC, true, DS><<<grid, ENV_BLOCK, 0, st>>>(h, cold, b, rq);
                e->last_queue = 1;
                return;
            }
        }
    }
    e->last_queue = 0;
    if (ro.pdl) {
        // Host-driven single steps (scg_step): programmatic dependent launch.  The kernel signals at its first instruction that
        // its dependents may be scheduled and then waits (griddepcontrol.wait) for the grid before it to complete and flush,
        // so launch latency, parameter upload and block ramp-up of step t + 1 overlap the body of step t.  Work submitted by
        // others behind it on the stream is ordered as usual (a launch without the attribute waits for full completion).
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t lc{};
        lc.gridDim = dim3(blocks); lc.blockDim = dim3(ENV_BLOCK); lc.dynamicSmemBytes = 0; lc.stream = st; lc.attrs = attr; lc.numAttrs = 1;
        if (cudaLaunchKernelEx(&lc, k_rollout<KIND, real, ATT, LEAN, PLAIN, SPEC, false, DS>, h, cold, b, ro) == cudaSuccess) return;
        (void)cudaGetLastError();   // not supported here: the plain launch below
    }
    k_rollout<KIND, real, ATT, LEAN, PLAIN, SPEC, false, DS><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro);
}

// The PLAIN + DS kernel applies (scg_step.cuh: step_env): none of the general extras, disturbances (if any) all inline, the
// restart (if it draws more than uniform init offsets) through apply_spare_simple.  Evaluated per launch: injected draws switch
// CFG_DIST_SIMPLE off (scg_set_injection).
template <int KIND, typename real, int ATT>
static void launch_rollout_att(scg_env* e, const Hot<real>& h, const Cold<real>* cold, const Buf<real>& b, const Roll<real>& ro,
                               unsigned blocks, cudaStream_t st) {
    constexpr bool F32 = sizeof(real) == 4;
    e->last_spec = 0;
    if (e->lean && e->plain && !ro.use_pid && F32) {
        if (KIND == SCG_QUAD_3D && e->spec_enabled && (h.cfg & CFG_SPEC_MASK) == CFG_SPEC_VAL) {
            e->last_spec = ro.step_outputs ? 2 : 1;
            if (ro.step_outputs) launch_rollout_v<KIND, real, ATT, true, F32, ((F32 && KIND == SCG_QUAD_3D) ? 2 : 0)>(e, h, cold, b, ro, blocks, st);
            else launch_rollout_v<KIND, real, ATT, true, F32, ((F32 && KIND == SCG_QUAD_3D) ? 1 : 0)>(e, h, cold, b, ro, blocks, st);
        } else launch_rollout_v<KIND, real, ATT, true, F32, 0>(e, h, cold, b, ro, blocks, st);
    }
    else if (e->lean && F32 && KIND != SCG_QUAD_1D && (KIND == SCG_QUAD_3D || !ro.use_pid) && ds_kernel_ok(h) &&
             !(KIND == SCG_QUAD_3D && h.dyn_uses_episode_inertia)) {   // (quad_type 3: closed-loop launches too)
        if (KIND != SCG_QUAD_1D && !ro.step_outputs && e->spec_enabled && (h.cfg & SpecCfg<KIND>::MASK) == SpecCfg<KIND>::VAL) {   // this kind's family switches compiled in (SpecCfg)
            e->last_spec = 1;
            launch_rollout_v<KIND, real, ATT, true, F32, ((F32 && KIND != SCG_QUAD_1D) ? 1 : 0), (F32 && KIND != SCG_QUAD_1D)>(e, h, cold, b, ro, blocks, st);
        } else launch_rollout_v<KIND, real, ATT, true, F32, 0, (F32 && KIND != SCG_QUAD_1D)>(e, h, cold, b, ro, blocks, st);
    }
    else if (e->lean) launch_rollout_v<KIND, real, ATT, true, false, 0>(e, h, cold, b, ro, blocks, st);
    else launch_rollout_v<KIND, real, ATT, false, false, 0>(e, h, cold, b, ro, blocks, st);
}

template <int KIND, typename real>
int launch_rollout(scg_env* e, const Roll<real>& ro, cudaStream_t st) {
    const Hot<real>& h = hot_of<real>(e);
    const Buf<real> b = typed_buf<real>(e->buf);
    const unsigned blocks = unsigned((b.n + ENV_BLOCK - 1) / ENV_BLOCK);
    const Cold<real>* cold = static_cast<const Cold<real>*>(e->cold_dev);
    const int att = e->cfg.attitude_mode;
    e->last_tlog = 0; e->last_spec = 0;
    if constexpr (KIND == SCG_QUAD_3D && sizeof(real) == 4) {
        if (ro.log_out && e->queue_mode != 1 && launch_rollout_tlog_3d_f(e, ro, st)) {   // staged (TMA) trajectory log
            e->launches += 1;
            SCG_CU(cudaGetLastError());
            return SCG_OK;
        }
    }
    if constexpr (KIND == SCG_QUAD_3D || KIND == SCG_QUAD_2D) {
        if (att == SCG_ATT_ROUNDTRIP) { e->last_queue = 0; k_rollout<KIND, real, SCG_ATT_ROUNDTRIP, false, false><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro); }
        else if (att == SCG_ATT_FAST_INTRINSIC && sizeof(real) == 4)
            launch_rollout_att<KIND, real, sizeof(real) == 4 ? SCG_ATT_FAST_INTRINSIC : SCG_ATT_FAST>(e, h, cold, b, ro, blocks, st);
        else launch_rollout_att<KIND, real, SCG_ATT_FAST>(e, h, cold, b, ro, blocks, st);
    } else if (KIND == SCG_CARTPOLE && sizeof(real) == 4 && att == SCG_ATT_FAST_INTRINSIC) {
        launch_rollout_att<KIND, real, (KIND == SCG_CARTPOLE && sizeof(real) == 4) ? SCG_ATT_FAST_INTRINSIC : SCG_ATT_FAST>(e, h, cold, b, ro, blocks, st);
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
