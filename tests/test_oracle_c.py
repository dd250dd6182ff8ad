"""The C restatement (CPU baseline / large-N checker) against the pinned numpy oracle."""
import numpy as np

from golden_util import load
from oracle.c_oracle import COracle3D
from oracle.quad_oracle import QuadOracle


def _cfg():
    cfg = dict(load("level0_dyn_s4")["cfg"])
    cfg.update(randomized_init=False, info_in_reset=False)
    return cfg


def test_c_oracle_matches_numpy_oracle_stepwise():
    g = load("level0_dyn_s4")
    cfg = _cfg()
    ref = QuadOracle(**cfg)
    ref.reset()
    co = COracle3D(QuadOracle(**cfg), n=3, auto_reset=True)
    co.reset()
    rew, done = np.zeros((1, 3)), np.zeros((1, 3), dtype=np.uint8)
    for t, a in enumerate(g["actions"]):
        obs, r, d, info = ref.step(a)
        co.rollout(np.repeat(a[None, None, :], 3, axis=1), rew, done)
        assert rew[0, 2] == r and bool(done[0, 0]) == d
        if d:   # auto-reset: the C oracle hands back the fresh episode's state (dummy_vec_env.py:38-49)
            ref.reset()
        np.testing.assert_allclose(co.obs[:, 1], ref.state, rtol=0, atol=1e-13)
    assert co.stats[0] == 3 * 2 and co.stats[9] == 3 * len(g["actions"])


def test_c_oracle_philox_reset_matches_numpy_oracle():
    from oracle.quad_oracle import PhiloxEnvDraws
    rand = {"init_x": dict(distrib="uniform", low=-0.1, high=0.1), "init_z": dict(distrib="uniform", low=0.47, high=1.47),
            "init_psi": dict(distrib="uniform", low=-0.1, high=0.1), "init_q": dict(distrib="uniform", low=-0.2, high=0.2)}
    cfg = _cfg()
    cfg.update(randomized_init=True, init_state_randomization_info=rand)
    co = COracle3D(QuadOracle(**cfg), n=5, seed=99, env_id_base=7, init_rand=rand)
    co.reset()
    for i in range(5):
        ref = QuadOracle(draws=PhiloxEnvDraws(99, 7 + i), **cfg)
        ref.reset()
        np.testing.assert_allclose(co.obs[:, i], ref.state, rtol=0, atol=1e-15)


# ---------------------------------------------------------------------------------------------------------------------
# The C ports of BASELINE configs 3 / 4 / 5 (cpu_baseline of bench.py) against the numpy oracles on the same Philox counters.
U = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)


def _run_pair(co, refs, acts, nx, tol=1e-12):
    """co: a C port over len(refs) envs; refs: numpy oracles of the same envs.  Free-running with auto-reset."""
    n = len(refs)
    co.reset()
    for o in refs:
        o.reset()
    rew, done = np.zeros((1, n)), np.zeros((1, n), dtype=np.uint8)
    resets = 0
    for t in range(acts.shape[0]):
        co.rollout(acts[t:t + 1], rew, done)
        for i, o in enumerate(refs):
            _, r, d, _ = o.step(acts[t, i])
            assert bool(done[0, i]) == bool(d), (t, i)
            assert abs(rew[0, i] - r) <= tol * max(1.0, abs(r)), (t, i, rew[0, i], r)
            if d:
                o.reset()
                resets += 1
            got = co.obs[:, i] if hasattr(co, "obs") else co.state[:, i]
            np.testing.assert_allclose(got, o.state, rtol=0, atol=tol * max(1.0, np.abs(o.state).max()), err_msg=f"step {t} env {i}")
    return resets


def test_c_port_config3_cartpole_matches_numpy_oracle():
    import bench
    from oracle.c_oracle import COracleCartPole
    from oracle.cartpole_oracle import CartPoleOracle
    from oracle.quad_oracle import PhiloxEnvDraws
    kw = bench.config3_kwargs()
    kw.update(episode_len_sec=1)   # 50-step episodes: several truncations inside the test
    seed = kw.pop("seed")
    n, base = 6, 40
    refs = [CartPoleOracle(draws=PhiloxEnvDraws(seed, base + i), **kw) for i in range(n)]
    co = COracleCartPole(CartPoleOracle(**kw), n=n, seed=seed, env_id_base=base, threads=2)
    acts = np.random.default_rng(1).normal(0, 4.0, size=(130, n, 1))
    acts[17, 3, 0] = 25.0   # beyond the input bound: violation -> done
    assert _run_pair(co, refs, acts, 4) >= 2 * n
    assert co.stats[9] == 130 * n and co.stats[0] >= 2 * n


def test_c_port_config4_quad2d_matches_numpy_oracle():
    import bench
    from oracle.c_oracle import COracleQuadExt
    from oracle.quad_oracle import PhiloxEnvDraws
    kw = bench.config4_kwargs()
    kw.update(episode_len_sec=1)
    seed = kw.pop("seed")
    n, base = 5, 900
    refs = [QuadOracle(draws=PhiloxEnvDraws(seed, base + i), **kw) for i in range(n)]
    co = COracleQuadExt(QuadOracle(**kw), n=n, seed=seed, env_id_base=base, threads=2)
    hover = refs[0].MASS * 9.8 / 2
    acts = np.clip(hover + np.random.default_rng(2).normal(0, 0.01, size=(150, n, 2)), 0, 4 * hover)
    assert _run_pair(co, refs, acts, 6) >= 2 * n


def test_c_port_config5_gates_and_log_match_numpy_oracle():
    import bench
    from oracle.c_oracle import COracleQuadExt
    from oracle.quad_oracle import PhiloxEnvDraws
    from test_gpu_gates import _kwargs as fly_through_kwargs
    for kw in (bench.config5_kwargs(), fly_through_kwargs(True)):
        kw = dict(kw)
        seed = kw.pop("seed", 5)
        n, base, k = 4, 123, 70
        refs = [QuadOracle(draws=PhiloxEnvDraws(seed, base + i), **kw) for i in range(n)]
        co = COracleQuadExt(QuadOracle(**kw), n=n, seed=seed, env_id_base=base, threads=2)
        acts = np.clip(0.03454 * 9.8 / 4 + np.random.default_rng(3).normal(0, 0.004, size=(k, n, 4)), 0.03, 0.14)
        _run_pair(co, refs, acts, 12)
        # the Logger record of one more launch: [pos3 vel3 rpy3 rpy_rates3 rpm4] per env and step
        log = np.zeros((5, 16, n))
        co.rollout(acts[:5], log_out=log)
        for t in range(5):
            for i, o in enumerate(refs):
                _, _, d, _ = o.step(acts[t, i])
                want = np.hstack([o.pos, o.vel, o.rpy, o.rpy_rates, o.last_clipped_action])
                np.testing.assert_allclose(log[t, :, i], want, rtol=0, atol=1e-9 * max(1.0, np.abs(want).max()))
                if d:
                    o.reset()
        assert kw.get("gate_geometry")
    assert co.stats[5] > 0 and co.stats[6] > 0   # the fly-through scenario passes gates and collides
