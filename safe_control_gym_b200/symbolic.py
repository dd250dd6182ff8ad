"""SymbolicModel facade (safe_control_gym/math_and_models/symbolic_systems.py) -- filled in below."""


class SymbolicModel:
    def __init__(self, env):
        self.env = env
        self.nx, self.nu, self.ny = env.state_dim, env.action_dim, env.state_dim
        self.dt = env.CTRL_TIMESTEP


def quadrotor_symbolic_model(env):
    return SymbolicModel(env)
