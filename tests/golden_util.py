"""Helpers shared by the golden-vector tests (oracle and CUDA path)."""
import json
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENARIOS = ["level0_dyn_s4", "level0_dyn_s1", "quad3d_tumble", "quad2d_disturbed",
             "quad1d_tracking", "quad3d_tracking",
             # round 2: cost "quadratic" (stabilisation with goal_reached; tracking with dense Q / R), a 140-step
             # level0 trajectory without reset, competition/level2.yaml (randomised gates / obstacles / inertia)
             "quad3d_quadratic_stab", "quad3d_quadratic_track", "level0_dyn_long", "level2_dyn"]


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
    if "Q" in g:
        env.set_cost_function_param(g["Q"], g["R"])
    nc = max(env.num_constraints, 1)
    ev = []
    env.reset_log = dict(gates=[], obstacles=[], inertia=[])

    def rec(kind, obs, rew, done, info):
        c = np.full(nc, np.nan)
        if "constraint_values" in info:
            v = np.asarray(info["constraint_values"])
            c[:len(v)] = v
        ev.append(dict(type=kind, obs=np.asarray(obs), rew=float(rew), done=int(bool(done)),
                       viol=int(info.get("constraint_violation", -1)),
                       trunc=int(info.get("TimeLimit.truncated", -1)),
                       mse=float(info.get("mse", np.nan)), cvals=c, state=env.state.copy(),
                       goal=int(info["goal_reached"]) if "goal_reached" in info else -1,
                       kin=np.hstack([env.pos, env.vel, env.rpy, env.rpy_rates])))

    def do_reset():
        ret = env.reset()
        obs, info = ret if env.INFO_IN_RESET else (ret, {})
        if not env.INFO_IN_RESET and env.constraints is not None:
            vals = [c.value(env) for c in env.constraints if c.var == "state"]
            info = {"constraint_values": np.concatenate(vals) if vals else np.zeros(0)}
        rec(0, obs, 0.0, False, info)
        env.reset_log["gates"].append(np.array(env.EFFECTIVE_GATES_POSITIONS, dtype=float).reshape(len(env.GATES), 6))
        env.reset_log["obstacles"].append(np.array([[o[0], o[1], o[2], o[5]] for o in env.EFFECTIVE_OBSTACLES],
                                                   dtype=float).reshape(len(env.OBSTACLES), 4))
        env.reset_log["inertia"].append(np.array([env.OVERRIDDEN_QUAD_MASS] + list(env.OVERRIDDEN_QUAD_INERTIA), dtype=float))

    do_reset()
    for t, a in enumerate(g["actions"]):
        if "adv" in g:
            env.set_adversary_control(g["adv"][t])
        obs, rew, done, info = env.step(a)
        rec(1, obs, rew, done, info)
        if done:
            do_reset()
    return ev, env


# ----------------------------------------------------------------------------------------
# Replay through the CUDA path (C ABI via the BatchedQuadrotor facade) with injected draws.
class DrawCursor:
    def __init__(self, draws):
        self.draws, self.pos = draws, 0

    def take(self):
        v = self.draws[self.pos]
        self.pos += 1
        return np.asarray(v, dtype=np.float64).reshape(-1)


def _needs_draw(d):
    from safe_control_gym_b200 import _lib as L
    return d.kind in (L.DIST_UNIFORM, L.DIST_WHITE, L.DIST_PERIODIC)


def _inject_channel(env, mode, cur, n):
    """Consume this call's draws of one channel (in list order) and hand them to the kernel."""
    if mode not in env.disturbances:
        return
    for k, d in enumerate(env.disturbances[mode].disturbances):
        if _needs_draw(d):
            v = cur.take()
            if mode == "dynamics" and not env._dyn_applies_force:
                continue   # consumed by the reference, never reaches the DYN integrator (F5)
            env.set_injection(mode, k, np.repeat(v[:, None], n, axis=1))


def cuda_reset(env, cur):
    """Quadrotor.reset() draw order: disturbance offsets -> (gates/obstacles) -> inertia -> init
    -> observation noise of the first _get_observation."""
    from safe_control_gym_b200 import _lib as L
    n = env.num_envs
    slots = max(sum(len(v.disturbances) for v in env.disturbances.values()), 1)
    offsets = np.zeros((slots, n), dtype=np.int32)
    for mode in env.disturbances:
        for d in env.disturbances[mode].disturbances:
            if d.kind in (L.DIST_IMPULSE, L.DIST_STEP) and d.step_offset is None:
                offsets[d.slot, :] = int(cur.take()[0])
    if env.RANDOMIZED_GATES_AND_OBS:   # obstacles first, then gates: x, y, yaw each (quadrotor.py:371-421).  The kernel derives
        for _ in range(3 * (len(env.OBSTACLES) + len(env.GATES))):   # its poses from Philox counters: nothing to inject
            cur.take()
    inertia = np.zeros((4, n))
    if env.RANDOMIZED_INERTIAL_PROP:
        for j, key in enumerate(["M", "Ixx", "Iyy", "Izz"]):
            if key in env.INERTIAL_PROP_RAND_INFO:
                inertia[j, :] = cur.take()[0]
    labels = env.INIT_STATE_LABELS[env.QUAD_TYPE]
    init = np.zeros((len(labels), n))
    if env.RANDOMIZED_INIT:
        for j, name in enumerate(labels):
            if name in env.INIT_STATE_RAND_INFO:
                init[j, :] = cur.take()[0]
    _inject_channel(env, "observation", cur, n)
    return env.reset(inject_init=init, inject_inertia=inertia, inject_offsets=offsets)


def cuda_step(env, cur, action, adv=None):
    n = env.num_envs
    _inject_channel(env, "action", cur, n)
    _inject_channel(env, "dynamics", cur, n)
    _inject_channel(env, "observation", cur, n)
    if adv is not None:
        env.set_adversary_control(np.repeat(np.asarray(adv)[None, :], n, axis=0))
    return env.step(np.repeat(np.asarray(action, dtype=np.float64)[None, :], n, axis=0))


def _pre_angles_from_quat(q):
    """Euler angles whose quatFromEuler() is the stored quaternion (x, y, z, w): the general
    getEulerFromQuaternion formulas WITHOUT the gimbal-band snap.  The kernel keeps the
    pre-round-trip angles as its attitude state, PyBullet keeps the quaternion; inside the band the
    snapped rpy of the golden record no longer determines the rotation, the quaternion does."""
    x, y, z, w = (float(v) for v in q)
    sarg = max(-1.0, min(1.0, -2.0 * (x * z - w * y)))
    roll = np.arctan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z)
    yaw = np.arctan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z)
    return roll, np.arcsin(sarg), yaw


def force_state(env, kin, ctrl_step, prev_violation, n_resets, quat=None):
    """Teacher forcing: overwrite the kernel's internal state with the golden kinematics
    (pos3, vel3, attitude, rpy_rates3) of the previous event."""
    import torch
    pos, vel, rpy, rates = kin[0:3], kin[3:6], kin[6:9], kin[9:12]
    if quat is not None and env.QUAD_TYPE == 3:
        rpy = _pre_angles_from_quat(quat)
    elif quat is not None and env.QUAD_TYPE == 2:
        rpy = (0.0, 2.0 * np.arctan2(float(quat[1]), float(quat[3])), 0.0)   # physical pitch, "flipped" bit clear
    if env.QUAD_TYPE == 3:
        x = [pos[0], vel[0], pos[1], vel[1], pos[2], vel[2], rpy[0], rpy[1], rpy[2], rates[0], rates[1], rates[2]]
    elif env.QUAD_TYPE == 2:
        x = [pos[0], vel[0], pos[2], vel[2], rpy[1], rates[1]]
    else:
        x = [pos[2], vel[2]]
    env.buf_state.copy_(torch.tensor(x, dtype=env.dtype, device=env.device)[:, None].expand(-1, env.num_envs))
    env.buf_ctrl_step.fill_(int(ctrl_step))
    env.buf_flags.fill_(1 if prev_violation else 0)
    env.buf_episode.fill_(int(n_resets))
