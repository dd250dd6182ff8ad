import sys, torch
sys.path.insert(0, '.')
import bench
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
dev = torch.device('cuda:0')
for att in ('fast', 'roundtrip'):
    env = BatchedQuadrotor(num_envs=65536, device=dev, dtype=torch.float32, attitude=att, auto_reset=True, **bench.env_kwargs())
    acts = bench.make_actions(torch, 65536, 100, dev, 1337, 0.005)
    env.reset()
    for r in range(3):
        env.rollout(acts)
        torch.cuda.synchronize()
        s = env.buf_state
        print(att, r, 'nan envs', int(torch.isnan(s).any(0).sum()), 'inf', int(torch.isinf(s).any(0).sum()),
              'max|pitch_pre|', float(s[7].abs().max()), 'max|rates|', float(s[9:12].abs().max()),
              'frac |sin(th)|>=0.99998', float((torch.sin(s[7]).abs() >= 0.99998).float().mean()),
              'max ctrl_step', int(env.buf_ctrl_step.max()), 'z range', float(s[4].min()), float(s[4].max()))
    env.close()
