"""Common host logic of the batched envs: buffer ownership, the C-ABI calls, and the
``BenchmarkEnv`` reset()/step() surface (safe_control_gym/envs/benchmark_env.py:63-463) in
VecEnv form (vec_env.py:16-163): ``reset() -> (obs[N, obs_dim], info)``,
``step(actions[N, nu]) -> (obs, rew[N], done[N], info)`` with ``info`` a dict of device tensors
carrying the reference's keys.  ``num_envs == 1`` with ``squeeze=True`` returns the reference's
scalar signatures (numpy obs, float reward, bool done, dict info).

torch is used as the zero-copy container for the batch memory and for streams; every
computation of the path happens in libscg_b200.so.
"""
import ctypes as C

import numpy as np
import torch

from .. import _lib as L


class LazyInfo(dict):
    """The per-step ``info`` dict with the reference's keys.  The kernel writes one packed byte per env;
    unpacking it into a dozen boolean tensors costs a dozen small launches, so each entry is built the
    first time it is read (``info["constraint_violation"]``, ``.get``, ``.items()`` ...)."""

    def __init__(self, makers):
        super().__init__()
        self._makers = makers

    def __missing__(self, key):
        if key not in self._makers:
            raise KeyError(key)
        v = self._makers[key]()
        self[key] = v
        return v

    def __contains__(self, key):
        return dict.__contains__(self, key) or key in self._makers

    def get(self, key, default=None):
        return self[key] if key in self else default

    def keys(self):
        return list(dict.fromkeys(list(dict.keys(self)) + list(self._makers)))

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return len(self.keys())

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def values(self):
        return [self[k] for k in self.keys()]


_RAW_STREAM = getattr(torch._C, "_cuda_getCurrentRawStream", None)   # the current stream's handle without a Stream object (3 us -> 0.2 us)


class _OnDevice:
    """``with torch.cuda.device(dev)`` only when ``dev`` is not already current: the guard object costs 2 us of a 6 us step()."""

    def __init__(self, index):
        self.index, self.guard = index, None

    def __enter__(self):
        if torch.cuda.current_device() != self.index:
            self.guard = torch.cuda.device(self.index)
            self.guard.__enter__()

    def __exit__(self, *exc):
        if self.guard is not None:
            g, self.guard = self.guard, None
            g.__exit__(*exc)
        return False


class BatchedEnvBase:
    NAME = "base"

    # ------------------------------------------------------------------ construction
    def _init_batch(self, cfg, num_envs, device, dtype, nx, nu, obs_dim, n_rows, n_slots, squeeze):
        if not torch.cuda.is_available():
            raise RuntimeError("safe_control_gym_b200 needs a CUDA device: there is no CPU fallback for the env step")
        self.lib = L.load()
        self.num_envs = int(num_envs)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("device must be a CUDA device")
        self.dtype = dtype
        self._squeeze = bool(squeeze) and self.num_envs == 1
        self._cfg = cfg
        self._np_real = np.float32 if dtype == torch.float32 else np.float64
        n = self.num_envs
        dev = self.device
        with torch.cuda.device(dev):
            handle = C.c_void_p()
            L.check(self.lib.scg_create(C.byref(cfg), C.byref(handle)))
        self._h = handle

        def z(*shape, dt=dtype):
            return torch.zeros(*shape, dtype=dt, device=dev)
        self.buf_state = z(nx, n)
        self.buf_ctrl_step = z(n, dt=torch.int32)
        self.buf_episode = z(n, dt=torch.int32)
        self.buf_flags = z(n, dt=torch.int32)
        self.buf_dist_offset = z(max(n_slots, 1), n, dt=torch.int32)
        self.buf_inertia = z(4, n)
        self.buf_obs = z(obs_dim, n)
        self.buf_env_state = z(nx, n)
        self.buf_reward = z(n)
        self.buf_done = z(n, dt=torch.uint8)
        self.buf_info = z(n, dt=torch.uint8)
        self.buf_mse = z(n)
        self.buf_c_values = z(max(n_rows, 1), n)
        self.buf_terminal_obs = z(obs_dim, n)
        self.buf_ep_return = z(n)
        self.buf_ep_length = z(n, dt=torch.int32)
        self.buf_ep_violations = z(n, dt=torch.int32)
        self.buf_aero = z(n)   # dyn_aero drag: rotor speed sum of the previous substep's action
        self.buf_stats = z(L.STAT_REPLICAS, L.STATS_LEN, dt=torch.float64)   # per-warp-group copies; totals: stats_vector()
        self._adv = None
        self._obs_view = self.buf_obs.t()
        self._done_view = self.buf_done.view(torch.bool)
        self._dev_index = dev.index if dev.index is not None else torch.cuda.current_device()
        self._on_device = _OnDevice(self._dev_index)
        self._step_makers = None   # the lazy info entries of step(): built once (they close over buffers that never move)
        self._bind()
        self.initial_reset = False
        self._keep = []        # reset()'s temporaries: alive until the next reset
        self._inject_keep = {}   # (channel, k) -> the injection tensor the device block points at

    def _bind(self, with_step_outputs=True):
        b = L.Buffers()
        b.n = self.num_envs
        b.state, b.ctrl_step, b.episode = self.buf_state.data_ptr(), self.buf_ctrl_step.data_ptr(), self.buf_episode.data_ptr()
        b.flags, b.dist_offset, b.inertia = self.buf_flags.data_ptr(), self.buf_dist_offset.data_ptr(), self.buf_inertia.data_ptr()
        b.obs, b.env_state, b.reward = self.buf_obs.data_ptr(), self.buf_env_state.data_ptr(), self.buf_reward.data_ptr()
        b.done, b.info = self.buf_done.data_ptr(), self.buf_info.data_ptr()
        if with_step_outputs:
            b.mse, b.c_values, b.terminal_obs = self.buf_mse.data_ptr(), self.buf_c_values.data_ptr(), self.buf_terminal_obs.data_ptr()
            b.ep_return, b.ep_length, b.ep_violations = (self.buf_ep_return.data_ptr(), self.buf_ep_length.data_ptr(),
                                                         self.buf_ep_violations.data_ptr())
        b.stats = self.buf_stats.data_ptr()
        b.aero_mem = self.buf_aero.data_ptr()
        b.adv_action = self._adv.data_ptr() if self._adv is not None else None
        self._buffers = b
        L.check(self.lib.scg_bind_buffers(self._h, C.byref(b)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.scg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _stream(self):
        if _RAW_STREAM is not None:
            return C.c_void_p(_RAW_STREAM(self._dev_index))
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _as_real(self, x, shape=None):
        t = torch.as_tensor(x, dtype=self.dtype, device=self.device)
        if shape is not None:
            t = t.reshape(shape)
        return t.contiguous()

    @property
    def launch_count(self):
        return int(self.lib.scg_launch_count(self._h))

    @property
    def lean(self):
        """True when the config selected the LEAN instantiation of the step kernel (scg_kernel_variant)."""
        return bool(self.lib.scg_kernel_variant(self._h))

    def set_rollout_queue(self, mode=-1, chunk_steps=0):
        """Scheduling of the fused rollout launches (scg_set_rollout_queue): -1 automatic, 0 one block per 64 envs for the
        whole launch, 1 always through the time-chunk ticket queue; ``chunk_steps`` 0 = default.  Same results either way."""
        L.check(self.lib.scg_set_rollout_queue(self._h, int(mode), int(chunk_steps)))

    @property
    def last_rollout_queued(self):
        """True when the last fused rollout launch went through the ticket queue."""
        return bool(self.lib.scg_last_rollout_queue(self._h))

    @property
    def last_rollout_tlog(self):
        """True when the last fused rollout wrote its trajectory log through the TMA unit (scg_last_rollout_tlog)."""
        return bool(self.lib.scg_last_rollout_tlog(self._h))

    @property
    def last_rollout_spec(self):
        """True when the last fused rollout ran a SPEC instantiation (the scenario switches of its family compiled in)."""
        return bool(self.lib.scg_last_rollout_spec(self._h))

    def set_spec_enabled(self, enabled):
        """Keep this env's launches on the non-SPEC kernels (``False``): same results; parity tests and A/B measurements."""
        L.check(self.lib.scg_set_spec_enabled(self._h, int(bool(enabled))))

    def set_injection(self, channel, k, tensor):
        """Feed pre-drawn noise for disturbance ``k`` of ``channel`` ([dim, N]); None = Philox."""
        ch = {"observation": L.CH_OBSERVATION, "action": L.CH_ACTION, "dynamics": L.CH_DYNAMICS}[channel]
        if tensor is None:
            L.check(self.lib.scg_set_injection(self._h, ch, k, None))
            self._inject_keep.pop((ch, k), None)
            return
        t = self._as_real(tensor)
        L.check(self.lib.scg_set_injection(self._h, ch, k, C.c_void_p(t.data_ptr())))
        # the device block keeps the pointer until it is replaced: so does this dict (the previous tensor of the slot may
        # still be read by a launch in flight on this stream; torch's caching allocator orders its reuse behind that)
        self._inject_keep[(ch, k)] = t

    # ------------------------------------------------------------------ reset / step
    def _info_from_buffers(self, reset=False):
        bits = self.buf_info
        if reset:   # after_reset: state constraints only, in YAML order (benchmark_env.py:376-383)
            return {"constraint_values": self.buf_c_values[self._state_row_idx].t()}

        if self._step_makers is not None:
            return LazyInfo(self._step_makers)

        def bit(mask, as_int=False):
            return (lambda: (bits & mask).ne(0).to(torch.int32)) if as_int else (lambda: (bits & mask).ne(0))
        extra = self._extra_info()
        self._step_makers = {
            **extra,
            "constraint_values": lambda: self.buf_c_values[:self._n_rows].t(),
            "constraint_violation": bit(L.INFO_VIOLATION, True),
            "TimeLimit.truncated": bit(L.INFO_TRUNCATED),
            "mse": lambda: self.buf_mse,
            "collision": bit(L.INFO_COLLISION),
            "stepped_through_gate": bit(L.INFO_GATE_PASSED),
            "at_goal_position": bit(L.INFO_AT_GOAL),
            "task_completed": bit(L.INFO_COMPLETED),
            "goal_reached": bit(L.INFO_GOAL_REACHED),
            "out_of_bound": bit(L.INFO_OUT_OF_BOUND),
            "current_target_gate_id": self._gate_id,
            "terminal_observation": lambda: self.buf_terminal_obs.t(),
            "state": lambda: self.buf_env_state.t(),
            "info_bits": lambda: bits,
        }
        return LazyInfo(self._step_makers)

    def _extra_info(self):
        """Env-specific lazy info keys (name -> maker)."""
        return {}

    def _gate_id(self):
        """info["current_target_gate_id"]: the gate index, -1 once all gates are passed (quadrotor.py:1093-1112)."""
        g = ((self.buf_flags >> L.FLAG_GATE_SHIFT) & 0xFF).to(torch.int32)
        return torch.where(g >= int(getattr(self, "NUM_GATES", 0)), torch.full_like(g, -1), g)

    def reset(self, mask=None, inject_init=None, inject_inertia=None, inject_offsets=None):
        """Quadrotor.reset() for every env (or those with ``mask[i] != 0``)."""
        self.before_reset()
        keep = []
        def ptr(t, dt):
            if t is None:
                return None
            t = torch.as_tensor(t, dtype=dt, device=self.device).contiguous()
            keep.append(t)
            return C.c_void_p(t.data_ptr())
        m = ptr(mask, torch.uint8)
        with torch.cuda.device(self.device):
            L.check(self.lib.scg_reset(self._h, m, ptr(inject_init, self.dtype), ptr(inject_inertia, self.dtype),
                                       ptr(inject_offsets, torch.int32), self._stream()))
        self._keep = keep
        obs = self.buf_obs.t()
        info = self._info_from_buffers(reset=True)
        info.update(self._reset_info())
        if self._squeeze:
            obs = obs[0].cpu().numpy().astype(np.float64)
            info = dict(info)
            info["constraint_values"] = info["constraint_values"][0].cpu().numpy().astype(np.float64)
            return (obs, info) if self.INFO_IN_RESET else obs
        return (obs, info) if self.INFO_IN_RESET else obs

    def before_reset(self):
        self.initial_reset = True

    def _reset_info(self):
        return {}

    def step(self, action):
        """Quadrotor.step(action) for the whole batch: ONE kernel launch."""
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        if isinstance(action, torch.Tensor) and action.dtype == self.dtype and action.device == self.device \
                and action.is_contiguous() and action.numel() == self.num_envs * self.action_dim:
            a = action   # already the [N, nu] device tensor the kernel reads
        else:
            a = self._as_real(action, (self.num_envs, self.action_dim))
        self.current_raw_input_action = a
        with self._on_device:
            rc = self.lib.scg_step(self._h, C.c_void_p(a.data_ptr()), self._stream())
        if rc != 0:
            L.check(rc)
        self._adv_consumed()
        obs, rew, done = self._obs_view, self.buf_reward, self._done_view
        info = self._info_from_buffers()
        if self._squeeze:
            i = {k: (v[0].cpu().numpy() if v.dim() > 1 else v[0].item()) for k, v in info.items()}
            i["constraint_values"] = np.asarray(i["constraint_values"], dtype=np.float64)
            i["collision"] = (None, bool(i["collision"]))   # (object id, flag) (quadrotor.py:1046-1057; ids are PyBullet's)
            if isinstance(i.get("current_target_gate_pos"), np.ndarray):
                i["current_target_gate_pos"] = i["current_target_gate_pos"].tolist() if i["current_target_gate_id"] >= 0 else []
            return obs[0].cpu().numpy().astype(np.float64), float(rew[0].item()), bool(done[0].item()), i
        return obs, rew, done, info

    def _adv_consumed(self):
        pass

    def rollout(self, actions, reward_out=None, done_out=None, log_out=None, log_stride=1):
        """K fused control steps with auto-reset; ``actions`` is a device tensor [K, N, nu].
        Returns the final observation view; per-episode statistics accumulate in ``stats()``."""
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        a = torch.as_tensor(actions, dtype=self.dtype, device=self.device).contiguous()
        k = a.shape[0]
        assert a.shape[1:] == (self.num_envs, self.action_dim), "actions must be [K, N, nu]"
        args = L.RolloutArgs()
        args.actions, args.k_steps, args.log_stride = a.data_ptr(), k, int(log_stride)
        args.reward_out = reward_out.data_ptr() if reward_out is not None else None
        args.done_out = done_out.data_ptr() if done_out is not None else None
        args.log_out = log_out.data_ptr() if log_out is not None else None
        with self._on_device:
            L.check(self.lib.scg_rollout(self._h, C.byref(args), self._stream()))
        return self.buf_obs.t()

    def rollout_noalloc(self, actions, reward_out=None, done_out=None, log_out=None, log_stride=1):
        """``rollout`` through scg_rollout_noalloc: the library allocates nothing (the ticket queue's scratch is a static
        device array, see include/scg_b200.h for the limits).  ``actions`` must already be the [K, N, nu] device tensor of
        the env's dtype -- nothing is converted or copied here either."""
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        a = actions
        assert a.is_cuda and a.dtype == self.dtype and a.is_contiguous() and a.shape[1:] == (self.num_envs, self.action_dim)
        args = L.RolloutArgs()
        args.actions, args.k_steps, args.log_stride = a.data_ptr(), a.shape[0], int(log_stride)
        args.reward_out = reward_out.data_ptr() if reward_out is not None else None
        args.done_out = done_out.data_ptr() if done_out is not None else None
        args.log_out = log_out.data_ptr() if log_out is not None else None
        with self._on_device:
            L.check(self.lib.scg_rollout_noalloc(self._h, C.byref(args), self._stream()))
        return self._obs_view

    def rollout_pid(self, k_steps, pid=None, pid_mem=None, actions_out=None, reward_out=None, done_out=None, log_out=None,
                    log_stride=1):
        """K fused CLOSED-LOOP control steps: the action of every step comes from the on-device DSL PID
        (quadrotor_utils.py:70-278) tracking ``X_GOAL`` -- the loop of competition_utils.thrusts /
        controllers/pid/pid.py in one launch, no action table.  ``pid``: ``_lib.PidConfig`` (default gains);
        ``pid_mem``: [9, N] controller memory, kept on the env between calls when omitted."""
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        pid = pid if pid is not None else L.PidConfig()
        if pid_mem is None:
            if getattr(self, "pid_mem", None) is None:
                self.pid_mem = torch.zeros(9, self.num_envs, dtype=self.dtype, device=self.device)
            pid_mem = self.pid_mem
        args = L.RolloutArgs()
        args.actions, args.k_steps, args.log_stride = None, int(k_steps), int(log_stride)
        args.reward_out = reward_out.data_ptr() if reward_out is not None else None
        args.done_out = done_out.data_ptr() if done_out is not None else None
        args.log_out = log_out.data_ptr() if log_out is not None else None
        out = C.c_void_p(actions_out.data_ptr()) if actions_out is not None else None
        with self._on_device:
            L.check(self.lib.scg_rollout_pid(self._h, C.byref(pid), C.c_void_p(pid_mem.data_ptr()), C.byref(args), out, self._stream()))
        return self.buf_obs.t()

    def rollout_host(self, host_actions, chunk_steps=50):
        """Same as ``rollout`` with the action table in pinned HOST memory [K, N, nu]; the
        library pipelines the host->device copies against the kernel launches."""
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        assert host_actions.dtype == self.dtype and host_actions.is_contiguous() and not host_actions.is_cuda
        k = host_actions.shape[0]
        with self._on_device:
            L.check(self.lib.scg_rollout_host(self._h, C.c_void_p(host_actions.data_ptr()), k, int(chunk_steps), self._stream()))
        return self.buf_obs.t()

    def stats_vector(self):
        """float64[STATS_LEN] device tensor: the accumulators summed over their replicas (include/scg_b200.h: SCG_STAT_REPLICAS)."""
        return self.buf_stats.sum(0)

    def stats(self, reset=False):
        """Episode statistics accumulated on the device (VecRecordEpisodeStatistics semantics)."""
        out = self.stats_vector().cpu().numpy().copy()
        if reset:
            self.buf_stats.zero_()
        return {name: float(out[i]) for i, name in enumerate(L.STAT_NAMES)}

    # ------------------------------------------------------------------ state access
    @property
    def state(self):
        s = self.buf_env_state.t()
        return s[0].cpu().numpy().astype(np.float64) if self._squeeze else s

    @property
    def ctrl_step_counter(self):
        return int(self.buf_ctrl_step[0].item()) if self._squeeze else self.buf_ctrl_step

    @property
    def pyb_step_counter(self):
        c = self.ctrl_step_counter
        return c * self.PYB_STEPS_PER_CTRL

    # ------------------------------------------------------------------ VecEnv attribute / method access
    def _get_indices(self, indices):
        """vec_env.py:154-163: None -> all envs, an int -> [int]."""
        if indices is None:
            return list(range(self.num_envs))
        return [indices] if isinstance(indices, (int, np.integer)) else list(indices)

    def get_attr(self, attr_name, indices=None):
        """dummy_vec_env.py:102-110: one value per selected env.  Per-env tensors ([N, ...]) are indexed, scenario
        constants (shared by the lock-step batch) are repeated."""
        idx = self._get_indices(indices)
        v = getattr(self, attr_name)
        if isinstance(v, torch.Tensor) and v.dim() >= 1 and v.shape[0] == self.num_envs:
            return [v[i] for i in idx]
        return [v for _ in idx]

    def set_attr(self, attr_name, values, indices=None):
        """dummy_vec_env.py:112-123.  The batch shares its scenario constants, so one value (or equal values) for ALL
        envs is the only assignment that has a meaning; anything else raises."""
        idx = self._get_indices(indices)
        vals = list(values) if isinstance(values, (list, tuple)) else [values] * len(idx)
        if len(vals) != len(idx):
            raise ValueError("set_attr: one value per selected env is required")
        if len(idx) != self.num_envs or any(not np.array_equal(np.asarray(v), np.asarray(vals[0])) for v in vals[1:]):
            raise ValueError("set_attr: the envs of a batch share their attributes; set one value for all of them")
        setattr(self, attr_name, vals[0])

    def env_method(self, method_name, *method_args, indices=None, **method_kwargs):
        """dummy_vec_env.py:125-143: the method runs once on the batch; its result is repeated per selected env."""
        idx = self._get_indices(indices)
        out = getattr(self, method_name)(*method_args, **method_kwargs)
        return [out for _ in idx]

    def get_env_random_state(self):
        """dummy_vec_env.py:88-92.  The env's randomness is counter-based: (seed, first global env id) plus the per-env
        episode counters determine every future draw."""
        return [{"seed": int(self._cfg.seed), "env_id_base": int(self._cfg.env_id_base), "episode": self.buf_episode.clone()}]

    def set_env_random_state(self, worker_random_states):
        """dummy_vec_env.py:94-100."""
        st = worker_random_states[0]
        if int(st["seed"]) != int(self._cfg.seed) or int(st["env_id_base"]) != int(self._cfg.env_id_base):
            raise ValueError("set_env_random_state: the state belongs to an env with another seed / env id range")
        self.buf_episode.copy_(st["episode"])

    def state_dict(self):
        """Checkpoint of the batch: a handful of tensors."""
        names = ["state", "ctrl_step", "episode", "flags", "dist_offset", "inertia", "ep_return", "ep_length",
                 "ep_violations", "stats", "obs", "env_state"]
        return {k: getattr(self, "buf_" + k).clone() for k in names}

    def load_state_dict(self, sd):
        for k, v in sd.items():
            getattr(self, "buf_" + k).copy_(v)
        self.initial_reset = True
