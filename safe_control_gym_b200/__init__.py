"""safe_control_gym_b200 -- B200-native batched env step for safe-control-gym's Quadrotor DYN /
CartPole path: hand-written sm_100a CUDA kernels behind a C ABI (include/scg_b200.h), with the
reference's BenchmarkEnv reset()/step(), ConstraintList and SymbolicModel surface on top."""
from .registration import make, register, get_config, spec  # noqa: F401

register(id="quadrotor", entry_point="safe_control_gym_b200.envs.quadrotor:BatchedQuadrotor",
         config_entry_point="safe_control_gym_b200.envs:quadrotor.yaml")   # envs/__init__.py:6-8
register(id="cartpole", entry_point="safe_control_gym_b200.envs.cartpole:BatchedCartPole")
