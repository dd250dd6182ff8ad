// scg_host.cuh -- host-side handle shared by the translation units of libscg_b200.so.
// scg_api.cu owns the extern "C" ABI; the kernels are instantiated per (env kind, precision) in
// scg_k_*.cu (compiled in parallel) behind the three launchers declared at the bottom.
#pragma once
#include <string>

#include "scg_params.cuh"

struct scg_env {
    scg_config cfg;
    int nx, nu, obs_dim, n_rows, n_slots;
    bool lean = false;             // the LEAN kernel specialisation applies (scg_step.cuh)
    bool plain = false;            // ... and none of the per-step extras is configured (PLAIN)
    scg::Hot<float> hf;
    scg::Hot<double> hd;
    void* cold_dev = nullptr;      // Cold<float> or Cold<double>
    double* stats_total_dev = nullptr;   // [SCG_STATS_LEN]: scg_stats_read folds the replicas here
    void* cold_host = nullptr;     // host mirror (for injection pointer updates)
    void* x_goal_dev = nullptr;
    scg_buffers buf;
    bool bound = false, did_reset = false;
    int64_t launches = 0;
    // scg_rollout_host staging
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};
    void* stage[2] = {nullptr, nullptr};
    bool stage_used[2] = {false, false};   // ev_consumed[j] has been recorded: the next copy into slot j waits for it
    size_t stage_bytes = 0;
    // time-chunk queue of the fused rollout (scg_step.cuh: k_rollout<..., QUEUE>)
    int32_t* q_mem = nullptr;      // [2 + ranges]: pop counter, push counter, ready list
    int64_t q_cap = 0;
    int n_sms = 0;
    int queue_mode = -1;           // -1 auto, 0 off, 1 on (SCG_QUEUE)
    int queue_chunk = 0;           // 0 auto (SCG_QUEUE_CHUNK)
    unsigned queue_sleep_ns = 500; // poll interval of a block that waits for its ticket's predecessor (SCG_QUEUE_SLEEP)
    int queue_blocks_per_sm = 8;   // workers of a queued launch per SM (0: every resident slot); 8 x 148 for 1,024 env blocks measured best
    int last_queue = 0;            // the last rollout launch went through the queue (scg_last_rollout_queue)
    int last_spec = 0;             // ... ran a SPEC instantiation (scg_last_rollout_spec)
    int spec_enabled = 1;          // scg_set_spec_enabled: 0 keeps every launch on the non-SPEC kernels (A/B runs, parity tests)
    int last_tlog = 0;             // ... wrote its trajectory log through the TMA unit (scg_last_rollout_tlog)
};

int scg_fail(int code, const std::string& msg);
int scg_cuda_fail(cudaError_t e, const char* what);
#define SCG_CU(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return scg_cuda_fail(_e, #x); } while (0)

namespace scg {

template <typename real> inline const Hot<real>& hot_of(const scg_env* e);
template <> inline const Hot<float>& hot_of<float>(const scg_env* e) { return e->hf; }
template <> inline const Hot<double>& hot_of<double>(const scg_env* e) { return e->hd; }

template <typename real> inline Buf<real> typed_buf(const scg_buffers& b) {
    Buf<real> t;
    t.n = b.n; t.state = (real*)b.state; t.ctrl_step = b.ctrl_step; t.episode = b.episode; t.flags = b.flags;
    t.dist_offset = b.dist_offset; t.inertia = (real*)b.inertia; t.obs = (real*)b.obs; t.env_state = (real*)b.env_state;
    t.reward = (real*)b.reward; t.done = b.done; t.info = b.info; t.mse = (real*)b.mse; t.c_values = (real*)b.c_values;
    t.terminal_obs = (real*)b.terminal_obs; t.ep_return = (real*)b.ep_return; t.ep_length = b.ep_length;
    t.ep_violations = b.ep_violations; t.stats = b.stats; t.adv_action = (const real*)b.adv_action;
    t.aero_mem = (real*)b.aero_mem;
    return t;
}

// defined (and explicitly instantiated) in scg_inst.cuh / scg_k_*.cu
template <int KIND, typename real> int launch_reset(scg_env* e, const ResetInject<real>& inj, const uint8_t* mask, cudaStream_t st);
template <int KIND, typename real> int launch_rollout(scg_env* e, const Roll<real>& ro, cudaStream_t st);
// scg_k_3d_f_tlog.cu: logged 3D fp32 rollouts with the TMA-staged log store; 1 = launched, 0 = not eligible
int launch_rollout_tlog_3d_f(scg_env* e, const Roll<float>& ro, cudaStream_t st);
template <int KIND, typename real> int launch_constraints(scg_env* e, const real* state, const real* action, int64_t n,
                                                          real* out_values, uint8_t* out_flags, cudaStream_t st);

// (see scg_inst.cuh: launch_rollout_att)
template <typename real> inline bool ds_kernel_ok(const Hot<real>& h) {
    const uint32_t c = h.cfg;
    if ((c & (CFG_COLD_CONSTRAINTS | CFG_QR_DENSE | CFG_PENALTY | CFG_AERO)) || h.normalized_action) return false;   // (gate geometry: quad_type 3 only, inline broad phase)
    if ((c & CFG_DISTURBED) && !(c & CFG_DIST_SIMPLE)) return false;
    if ((c & CFG_RESET_EXTRAS) && !((c & CFG_SPARE_SIMPLE) && ((c & CFG_DIST_SIMPLE) || h.n_dist_act == 0))) return false;
    return true;
}

}  // namespace scg
