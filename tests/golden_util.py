"""Helpers shared by the golden-vector tests (oracle and CUDA path)."""
import json
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENARIOS = ["level0_dyn_s4", "level0_dyn_s1", "quad3d_tumble", "quad2d_disturbed",
             "quad1d_tracking", "quad3d_tracking"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False)
    g = {k: z[k] for k in z.files}
    g["cfg"] = json.loads(str(g.pop("config_json")))
    g["cfg"].pop("physics", None)
    draws, pos = [], 0
    for n in g["draw_len"]:
        draws.append(g["draw_flat"][pos:pos + n])
        pos += n
    g["draws"] = draws
    return g


def replay_oracle(g):
    """Run the numpy oracle over the golden action sequence with the recorded draws.
    Returns the list of events in the same layout make_golden.py recorded."""
    from oracle.quad_oracle import QuadOracle, ReplayDraws
    env = QuadOracle(draws=ReplayDraws(g["draws"]), **g["cfg"])
    nc = max(env.num_constraints, 1)
    ev = []

    def rec(kind, obs, rew, done, info):
        c = np.full(nc, np.nan)
        if "constraint_values" in info:
            v = np.asarray(info["constraint_values"])
            c[:len(v)] = v
        ev.append(dict(type=kind, obs=np.asarray(obs), rew=float(rew), done=int(bool(done)),
                       viol=int(info.get("constraint_violation", -1)),
                       trunc=int(info.get("TimeLimit.truncated", -1)),
                       mse=float(info.get("mse", np.nan)), cvals=c, state=env.state.copy(),
                       kin=np.hstack([env.pos, env.vel, env.rpy, env.rpy_rates])))

    def do_reset():
        ret = env.reset()
        obs, info = ret if env.INFO_IN_RESET else (ret, {})
        if not env.INFO_IN_RESET and env.constraints is not None:
            vals = [c.value(env) for c in env.constraints if c.var == "state"]
            info = {"constraint_values": np.concatenate(vals) if vals else np.zeros(0)}
        rec(0, obs, 0.0, False, info)

    do_reset()
    for t, a in enumerate(g["actions"]):
        if "adv" in g:
            env.set_adversary_control(g["adv"][t])
        obs, rew, done, info = env.step(a)
        rec(1, obs, rew, done, info)
        if done:
            do_reset()
    return ev, env
