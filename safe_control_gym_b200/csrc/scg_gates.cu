// scg_gates.cu -- info["current_target_gate_in_range" / "current_target_gate_pos"] for the whole batch
// (Quadrotor._get_info, quadrotor.py:1093-1112): the reference asks PyBullet for the closest points between the
// current gate's body and the drone within VISIBILITY_RANGE = 0.45 m and reports the true (randomised) gate pose
// when there are any, the nominal one otherwise.  Closed form here: distance from the drone (sphere r = 0.06) to the
// gate's four bars and its support, pose offsets recomputed from the episode's Philox counters.
#include "scg_env.cuh"
#include "scg_host.cuh"

namespace scg {

template <typename real> SCG_DEV real dist_box(const real (&p)[3], real cx, real cy, real cz, real hx, real hy, real hz) {
    const real q0 = m_max(m_abs(p[0] - cx) - hx, real(0)), q1 = m_max(m_abs(p[1] - cy) - hy, real(0)),
               q2 = m_max(m_abs(p[2] - cz) - hz, real(0));
    return m_sqrt(q0 * q0 + q1 * q1 + q2 * q2);
}
template <typename real> SCG_DEV real dist_vcyl(const real (&p)[3], real cz, real r, real hl) {
    const real dxy = m_max(m_sqrt(p[0] * p[0] + p[1] * p[1]) - r, real(0)), dz = m_max(m_abs(p[2] - cz) - hl, real(0));
    return m_sqrt(dxy * dxy + dz * dz);
}

template <typename real>
__global__ void __launch_bounds__(128)
k_gate_info(const __grid_constant__ Hot<real> h, const Cold<real>* __restrict__ cold, const Buf<real> b,
            real* __restrict__ out_pose, uint8_t* __restrict__ out_in_range) {
    const int64_t n = b.n;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = (b.flags[i] >> SCG_FLAG_GATE_SHIFT) & 0xFFu;
    real pose[6] = {real(0), real(0), real(0), real(0), real(0), real(0)};
    bool in_range = false;
    if (g < uint32_t(h.n_gates)) {
        const real* G = cold->gates[g];
        const RngKey key{h.seed, h.env_id_base_lo + uint32_t(i)};
        const uint32_t ep = h.reseed_on_reset ? 0u : uint32_t(b.episode[i] - 1);
        real off[3];
        pose_offsets<real>(cold->gate_rand, STREAM_RESET_GATES, g, key, ep, off);
        const real height = G[6] == real(0) ? real(1.0) : real(0.525);
        const real gx = G[0] + off[0], gy = G[1] + off[1], gz = G[2] + height, yaw = G[5] + off[2];
        real sy, cy;
        m_sincos(yaw, &sy, &cy);
        const real ex = b.state[i] - gx, ey = b.state[2 * n + i] - gy;
        const real p[3] = {cy * ex + sy * ey, -sy * ex + cy * ey, b.state[4 * n + i] - gz};
        real d = m_min(m_min(dist_box<real>(p, 0, 0, real(0.225), real(0.25), real(0.025), real(0.025)),
                             dist_box<real>(p, 0, 0, real(-0.225), real(0.25), real(0.025), real(0.025))),
                       m_min(dist_box<real>(p, real(0.225), 0, 0, real(0.025), real(0.025), real(0.25)),
                             dist_box<real>(p, real(-0.225), 0, 0, real(0.025), real(0.025), real(0.25))));
        d = m_min(d, G[6] == real(0) ? dist_vcyl<real>(p, real(-0.6), real(0.05), real(0.4))
                                     : dist_box<real>(p, 0, 0, real(-0.4), real(0.075), real(0.075), real(0.125)));
        in_range = d - Geo<real>::drone_r <= real(0.45);
        if (in_range) { pose[0] = gx; pose[1] = gy; pose[2] = gz; pose[3] = G[3]; pose[4] = G[4]; pose[5] = yaw; }
        else {
#pragma unroll
            for (int j = 0; j < 6; ++j) pose[j] = G[j];
        }
    }
    if (out_pose) {
#pragma unroll
        for (int j = 0; j < 6; ++j) out_pose[int64_t(j) * n + i] = pose[j];
    }
    if (out_in_range) out_in_range[i] = in_range ? 1 : 0;
}

template <typename real>
int launch_gate_info(scg_env* e, void* out_pose, uint8_t* out_in_range, cudaStream_t st) {
    const Hot<real>& h = hot_of<real>(e);
    const Buf<real> b = typed_buf<real>(e->buf);
    const unsigned blocks = unsigned((b.n + 127) / 128);
    k_gate_info<real><<<blocks, 128, 0, st>>>(h, static_cast<const Cold<real>*>(e->cold_dev), b, static_cast<real*>(out_pose), out_in_range);
    e->launches += 1;
    SCG_CU(cudaGetLastError());
    return SCG_OK;
}

}  // namespace scg

extern "C" int scg_gate_info(scg_env* e, void* out_pose, uint8_t* out_in_range, void* stream) {
    if (!e) return scg_fail(SCG_ERR_NULL, "scg_gate_info: null env");
    if (!e->bound) return scg_fail(SCG_ERR_STATE, "scg_gate_info: buffers not bound");
    if (e->cfg.env_kind != SCG_QUAD_3D) return scg_fail(SCG_ERR_BAD_CONFIG, "scg_gate_info: gates exist for quad_type 3 only");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return e->cfg.precision == 32 ? scg::launch_gate_info<float>(e, out_pose, out_in_range, st)
                                  : scg::launch_gate_info<double>(e, out_pose, out_in_range, st);
}
