"""CPU oracle for the batched env-step hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import or execute it, and there only as the
checker or the timed CPU baseline -- never as a fallback for the CUDA path.

Contents
--------
``bullet_math``   PyBullet's three quaternion helpers restated (pybullet ^3.2.0
                  is a third-party dependency absent from /root/reference).
``philox``        Philox4x32-10 + uniform/normal transforms (the counter-based
                  RNG contract shared with the CUDA kernels).
``quad_oracle``   numpy fp64 restatement of the reference Quadrotor DYN step
                  (single env, reference execution style).
``cartpole_oracle`` cart-pole restatement (env absent from the fork: parity
                  unpinned).
``symbolic_oracle`` closed-form fc / RK4 fd / df / loss of the prior models.
``scg_oracle.c``  the same step as plain C (OpenMP over envs) used as the
                  native CPU baseline and for large-N parity.

Parity pin status: the reference ships no numerical tests.
The oracle is pinned against the reference's OWN Python code executed in the
build container with pybullet / gym / casadi replaced by small stand-ins
(``tests/golden/make_golden.py``); the committed ``tests/golden/*.npz`` hold
those outputs.  What stays unpinned: the three PyBullet quaternion helpers
(restated from bullet3, not executed), CasADi's CVODES integrator, CartPole
(absent from this fork), and the analytic gate / obstacle geometry.
"""
