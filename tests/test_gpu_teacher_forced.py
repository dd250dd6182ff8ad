"""fp32 per-step parity of BASELINE configs 2, 3 and 4 at the north-star bar (1e-5 per step, unit floor):
the fp32 kernel is TEACHER-FORCED from the fp64 oracle -- before every control step the sampled lanes' internal state
is overwritten with the oracle's -- so each step is compared on (the fp32 rounding of) the same state, on the same
Philox streams (key = (seed, global env id)), auto-resets included.  The free-running versions of these comparisons
(test_gpu_batch / test_gpu_cartpole / test_gpu_config4) stay as drift reports.

Integer outputs (done, violation flags, step / episode counters) must agree exactly on every step.
"""
import numpy as np
import pytest
import torch

import bench
from oracle.cartpole_oracle import CartPoleOracle
from oracle.quad_oracle import PhiloxEnvDraws, QuadOracle
from test_gpu_cartpole import CFG as CFG3
from test_gpu_config4 import CFG as CFG4
from golden_util import _pre_angles_from_quat
from test_gpu_golden import _report

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _internal_state(o, kind):
    """The kernel's persistent state of one env, from the oracle's attributes (reference state order; rate slots hold
    rpy_rates, not PyBullet's INIT_ANG_VEL)."""
    if kind == "cartpole":
        return np.asarray(o.state, dtype=np.float64)
    if kind == 2:
        return np.array([o.pos[0], o.vel[0], o.pos[2], o.vel[2], o.rpy[1], o.rpy_rates[1]])
    # attitude: angles whose quatFromEuler() is the quaternion PyBullet holds (inside the gimbal band the snapped rpy no
    # longer determines the rotation, see golden_util._pre_angles_from_quat)
    r = _pre_angles_from_quat(o.quat)
    return np.array([o.pos[0], o.vel[0], o.pos[1], o.vel[1], o.pos[2], o.vel[2], r[0], r[1], r[2],
                     o.rpy_rates[0], o.rpy_rates[1], o.rpy_rates[2]])


def _force(env, lanes, refs, kind):
    idx = torch.tensor(lanes, device=env.device)
    st = np.stack([_internal_state(refs[i], kind) for i in lanes], axis=1)
    env.buf_state[:, idx] = torch.tensor(st, dtype=env.dtype, device=env.device)
    # counters and the lagged violation flag are integers that already agree (asserted every step)


def _run(env, refs, lanes, acts, kind, near_gimbal=None):
    obs, _ = env.reset()
    for i in lanes:
        o_obs, _ = refs[i].reset()
        assert np.max(np.abs(obs[i].double().cpu().numpy() - o_obs)) <= TOL
    worst, worst_near, n_near, resets = 0.0, 0.0, 0, 0
    for t in range(acts.shape[0]):
        _force(env, lanes, refs, kind)
        near = {i: (near_gimbal(refs[i]) if near_gimbal else False) for i in lanes}
        obs, rew, done, info = env.step(acts[t])
        term = info["terminal_observation"]
        viol = info["constraint_violation"].cpu().numpy()
        obs_c, term_c, rew_c, done_c = obs.double().cpu().numpy(), term.double().cpu().numpy(), rew.double().cpu().numpy(), done.cpu().numpy()
        steps_c, eps_c = env.buf_ctrl_step.cpu().numpy(), env.buf_episode.cpu().numpy()
        for i in lanes:
            o_obs, o_rew, o_done, o_info = refs[i].step(acts[t, i])
            near[i] = near[i] or (near_gimbal(refs[i]) if near_gimbal else False)
            assert bool(done_c[i]) == bool(o_done), (t, i)
            assert int(viol[i]) == int(o_info.get("constraint_violation", 0)), (t, i)
            got = (term_c if o_done else obs_c)[i]
            err = float(np.max(np.abs(got - o_obs) / np.maximum(1.0, np.abs(o_obs))))
            err = max(err, abs(rew_c[i] - o_rew) / max(1.0, abs(o_rew)))
            if near[i]:
                n_near += 1
                worst_near = max(worst_near, err)
                assert err <= 50 * TOL, ("near gimbal", t, i, err)
            else:
                worst = max(worst, err)
                assert err <= TOL, (t, i, err)
            if o_done:
                resets += 1
                r_obs, _ = refs[i].reset()
                assert np.max(np.abs(obs_c[i] - r_obs)) <= TOL
            assert int(steps_c[i]) == refs[i].ctrl_step_counter and int(eps_c[i]) == refs[i].episode + 1, (t, i)
    return worst, worst_near, n_near, resets


@pytest.mark.parametrize("attitude", ["fast", "fast_intrinsic"])
def test_config3_cartpole_teacher_forced(attitude):
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    n, steps, base = 96, 260, 1000
    env = BatchedCartPole(num_envs=n, dtype=torch.float32, attitude=attitude, auto_reset=True, seed_offset=base, **CFG3)
    rng = np.random.default_rng(5)
    acts = rng.normal(0, 4.0, size=(steps, n, 1))
    acts[17, 3, 0] = 25.0
    lanes = [0, 3, 31, 32, 64, 95]
    kw = {k: v for k, v in CFG3.items() if k != "seed"}
    refs = {i: CartPoleOracle(draws=PhiloxEnvDraws(CFG3["seed"], base + i), **kw) for i in lanes}
    worst, _, _, resets = _run(env, refs, lanes, acts, "cartpole")
    assert resets >= 8
    print(f"config 3 (CartPole) fp32 teacher-forced: worst per-step err {worst:.2e} over {steps} steps x {len(lanes)} lanes, {resets} resets")
    _report(test="teacher_forced", scenario="config3_cartpole", attitude=attitude, dtype="fp32", tol=TOL, worst_unit_floor=worst, steps=steps, resets=resets)
    env.close()


@pytest.mark.parametrize("attitude", ["fast", "fast_intrinsic"])
def test_config4_quad2d_teacher_forced(attitude):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, steps, base = 64, 200, 500
    env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude=attitude, auto_reset=True, seed_offset=base, info_in_reset=True, **CFG4)
    kw = {k: v for k, v in CFG4.items() if k != "seed"}
    lanes = [0, 7, 31, 32, 63]
    refs = {i: QuadOracle(draws=PhiloxEnvDraws(CFG4["seed"], base + i), info_in_reset=True, **kw) for i in lanes}
    hover = env.MASS * 9.8 / 2
    acts = np.clip(hover + np.random.default_rng(9).normal(0, 0.01, size=(steps, n, 2)), 0, 4 * hover)
    worst, _, _, resets = _run(env, refs, lanes, acts, 2)
    assert resets >= 8
    print(f"config 4 (2D, impulse + step, per-episode inertia) fp32 teacher-forced: worst per-step err {worst:.2e}, {resets} resets")
    _report(test="teacher_forced", scenario="config4_quad2d", attitude=attitude, dtype="fp32", tol=TOL, worst_unit_floor=worst, steps=steps, resets=resets)
    env.close()


@pytest.mark.parametrize("attitude", ["fast", "fast_intrinsic"])
def test_config2_quad3d_teacher_forced(attitude):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, steps, base = 64, 160, 7000
    kw = bench.env_kwargs()
    seed = kw.pop("seed")
    env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude=attitude, auto_reset=True, seed=seed, seed_offset=base,
                           info_in_reset=True, **kw)
    lanes = [0, 1, 13, 31, 32, 33, 50, 63]
    refs = {i: QuadOracle(draws=PhiloxEnvDraws(seed, base + i), info_in_reset=True, **kw) for i in lanes}
    acts = bench.make_actions(torch, n, steps, torch.device("cuda:0"), 3, 0.005).cpu().numpy().astype(np.float64)
    near = lambda o: abs(o.rpy[1]) > np.pi / 2 - 0.15   # the conditioning exception of test_gpu_golden
    worst, worst_near, n_near, resets = _run(env, refs, lanes, acts, 3, near)
    assert resets >= 4
    print(f"config 2 (3D level0, {attitude}) fp32 teacher-forced: worst per-step err {worst:.2e} ({n_near} lane-steps near gimbal lock, "
          f"worst there {worst_near:.2e}), {resets} resets")
    _report(test="teacher_forced", scenario="config2_quad3d", attitude=attitude, dtype="fp32", tol=TOL, worst_unit_floor=worst,
            near_gimbal_lane_steps=n_near, worst_near_gimbal=worst_near, steps=steps, resets=resets)
    env.close()
