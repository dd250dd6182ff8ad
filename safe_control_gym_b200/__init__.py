"""safe_control_gym_b200 -- B200-native batched env step for safe-control-gym's Quadrotor DYN /
CartPole path: hand-written sm_100a CUDA kernels behind a C ABI (include/scg_b200.h), with the
reference's BenchmarkEnv reset()/step(), ConstraintList and SymbolicModel surface on top."""
from .registration import make, register, get_config  # noqa: F401

register(id="quadrotor", entry_point="safe_control_gym_b200.envs.quadrotor:BatchedQuadrotor")
register(id="cartpole", entry_point="safe_control_gym_b200.envs.cartpole:BatchedCartPole")
