"""CPU baseline C2 (oracle/batched_numpy.py, the numpy-vectorised 3D step) against the single-env oracle."""
import numpy as np

import bench
from oracle.batched_numpy import BatchedQuadOracle
from oracle.quad_oracle import QuadOracle


def test_batched_numpy_step_matches_single_env_oracle():
    kw = bench.env_kwargs()
    kw.pop("seed")
    n, steps = 6, 45
    rng = np.random.default_rng(5)
    names = ("init_x", "init_x_dot", "init_y", "init_y_dot", "init_z", "init_z_dot", "init_phi", "init_theta", "init_psi")
    inits = np.array([[rng.uniform(-1, 1), 0.1, rng.uniform(-1, 1), -0.1, rng.uniform(0.5, 1.5), 0.0,
                       rng.uniform(-0.3, 0.3), rng.uniform(-0.3, 0.3), rng.uniform(-3, 3)] for _ in range(n)]).T
    inits[7, 0] = 1.5                      # one env starts next to the gimbal band
    refs = []
    for i in range(n):
        k1 = dict(kw, randomized_init=False, init_state=dict({k: float(inits[j, i]) for j, k in enumerate(names)},
                                                              init_p=0, init_q=0, init_r=0))
        refs.append(QuadOracle(**k1))
        refs[-1].reset()
    env = BatchedQuadOracle(n, kw)
    env.reset(init=inits)
    hover = 0.03454 * 9.8 / 4
    acts = hover + rng.normal(0, 0.01, size=(steps, 4, n))
    acts[30, :, 2] = 0.0                   # an input violation
    alive = np.ones(n, dtype=bool)
    for t in range(steps):
        rew, done, viol = env.step(acts[t], auto_reset=False)
        st = env.state()
        for i in range(n):
            if not alive[i]:
                continue
            _, r, d, info = refs[i].step(acts[t, :, i])
            assert np.max(np.abs(st[:, i] - refs[i].state[:9])) <= 1e-12, (t, i)
            assert float(rew[i]) == float(r) and bool(done[i]) == bool(d)
            assert int(viol[i]) == int(info["constraint_violation"])
            if d:
                alive[i] = False
    assert not alive[2] and alive.sum() >= 3
