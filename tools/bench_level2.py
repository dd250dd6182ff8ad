import sys, torch; sys.path.insert(0,'/root/repo')
import bench
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
from tools.bench_configs import timed
kw = bench.env_kwargs()
U = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
kw.update(randomized_inertial_prop=True, inertial_prop_randomization_info={"M": U(-0.01, 0.01), "Ixx": U(-1e-6, 1e-6), "Iyy": U(-1e-6, 1e-6), "Izz": U(-1e-6, 1e-6)},
          randomized_gates_and_obstacles=True, gates_and_obstacles_randomization_info=dict(gates=U(-0.1, 0.1), obstacles=U(-0.2, 0.2)))
n, k = 65536, 200
env = BatchedQuadrotor(num_envs=n, device="cuda:0", dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
env.reset()
a = bench.make_actions(torch, n, k, torch.device("cuda:0"), 5, 0.005)
dt = timed(lambda: env.rollout(a), 5)
print("level2-like (randomised inertia, init, gates): variant", env.lib.scg_kernel_variant(env._h), f"{n*k/dt/1e9:.2f} G env-steps/s")
