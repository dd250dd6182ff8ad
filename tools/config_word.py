import sys, torch
sys.path.insert(0, ".")
import bench
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
from safe_control_gym_b200.envs.cartpole import BatchedCartPole
kw = bench.config3_kwargs(); kw.pop("info_in_reset")
e3 = BatchedCartPole(num_envs=64, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
e4 = BatchedQuadrotor(num_envs=64, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.config4_kwargs())
e2 = BatchedQuadrotor(num_envs=64, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.env_kwargs())
for n, e in (("3", e3), ("4", e4), ("2", e2)):
    print(n, hex(e.lib.scg_config_word(e._h)), e.lib.scg_kernel_variant(e._h))
