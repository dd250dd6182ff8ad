// Logged fused rollouts of the 3D fp32 kernels with the trajectory log written through the TMA unit (north_star: "TMA
// staging only where trajectories are logged"; the record is Logger.log's, envs/gym_pybullet_drones/Logger.py:50-66, 83-119).
// Its own translation unit: eight more instantiations of the rollout loop compile in parallel with scg_k_3d_f.cu.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdlib>

#include "scg_host.cuh"
#include "scg_step.cuh"

namespace scg {

static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    }
    return fn;
}

// log tensor [rows = logged steps * 16][n] floats; box = box_rows x 32 envs, dense in shared memory (no swizzle)
static bool encode(CUtensorMap* m, void* base, uint64_t n, uint64_t rows, uint32_t box_rows) {
    const cuuint64_t dims[2] = {n, rows};
    const cuuint64_t strides[1] = {n * sizeof(float)};
    const cuuint32_t box[2] = {32u, box_rows};
    const cuuint32_t elem[2] = {1u, 1u};
    return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int ATT, bool DS>
static void launch_att(scg_env* e, const Hot<float>& h, const Cold<float>* cold, const Buf<float>& b, const Roll<float>& ro,
                       const LogMaps& maps, unsigned blocks, cudaStream_t st) {
    e->last_spec = (e->spec_enabled && (h.cfg & CFG_SPEC_MASK) == CFG_SPEC_VAL) ? 1 : 0;
    if (e->last_spec) k_rollout_tlog<ATT, DS, true><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro, maps);
    else k_rollout_tlog<ATT, DS, false><<<blocks, ENV_BLOCK, 0, st>>>(h, cold, b, ro, maps);
}

// 1: the launch was taken (the caller counts it and checks cudaGetLastError); 0: not eligible, take the plain store path.
// Eligible: a log is wanted, the config takes a PLAIN or PLAIN + DS kernel (fast attitude mode, none of the general per-step extras), no per-step
// outputs, and a log tensor the TMA unit can address: base 16-byte aligned and a row pitch (n floats) that is a multiple of
// 16 bytes.  SCG_TLOG=0 disables it (A/B measurements).  The other LEAN configs (gate geometry, closed loop, randomised
// inertia) are compute-bound at ~1.2 TB/s of log traffic and measured SLOWER with the staging (13.3 vs 14.8 G env-steps/s
// with gate geometry: 4 KB more shared memory per block costs a resident block): they keep the plain stores.
int launch_rollout_tlog_3d_f(scg_env* e, const Roll<float>& ro, cudaStream_t st) {
    static const bool enabled = [] { const char* v = getenv("SCG_TLOG"); return !(v && atoi(v) == 0); }();
    if (!enabled || !ro.log_out || ro.step_outputs || !e->lean || ro.use_pid || e->cfg.attitude_mode == SCG_ATT_ROUNDTRIP) return 0;
    const Hot<float>& h = hot_of<float>(e);
    const bool ds = !e->plain;   // the PLAIN + DS loop (scg_inst.cuh: launch_rollout_att takes the same decision)
    if (ds && !(ds_kernel_ok(h) && !h.dyn_uses_episode_inertia)) return 0;
    const int64_t n = e->buf.n;
    if ((n & 3) != 0 || (reinterpret_cast<uintptr_t>(ro.log_out) & 15u) != 0 || !encode_fn()) return 0;
    const uint64_t rows = uint64_t((ro.k_steps + ro.log_stride - 1) / ro.log_stride) * 16u;
    if (rows >= (uint64_t(1) << 31)) return 0;
    LogMaps maps;
    if (!encode(&maps.full, ro.log_out, uint64_t(n), rows, LOG_T * 16) || !encode(&maps.one, ro.log_out, uint64_t(n), rows, 16)) return 0;
    const Buf<float> b = typed_buf<float>(e->buf);
    const unsigned blocks = unsigned((n + ENV_BLOCK - 1) / ENV_BLOCK);
    const Cold<float>* cold = static_cast<const Cold<float>*>(e->cold_dev);
    if (e->cfg.attitude_mode == SCG_ATT_FAST_INTRINSIC) {
        if (ds) launch_att<SCG_ATT_FAST_INTRINSIC, true>(e, h, cold, b, ro, maps, blocks, st);
        else launch_att<SCG_ATT_FAST_INTRINSIC, false>(e, h, cold, b, ro, maps, blocks, st);
    } else {
        if (ds) launch_att<SCG_ATT_FAST, true>(e, h, cold, b, ro, maps, blocks, st);
        else launch_att<SCG_ATT_FAST, false>(e, h, cold, b, ro, maps, blocks, st);
    }
    e->last_queue = 0;
    e->last_tlog = 1;
    return 1;
}

}  // namespace scg
