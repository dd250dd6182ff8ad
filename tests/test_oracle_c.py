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
