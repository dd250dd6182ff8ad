"""Trajectory export in the reference Logger's formats.

The rollout kernel writes the Logger record -- pos3, vel3, rpy3, ang_vel3, rpm4, in the re-ordered layout of
``Logger.log`` (envs/gym_pybullet_drones/Logger.py:50-66, 115) -- as ``log[T][16][N]`` on the device
(``env.rollout(..., log_out=log, log_stride=s)``).  ``TrajectoryLog`` turns a slice of envs of that tensor into the
arrays and files the reference's plotting / sim-to-real tools read:

* ``states``  ``[num_drones, 16, T]`` and ``timestamps`` ``[num_drones, T]`` (Logger.py:46-48);
* ``save()``  one ``.npy``-named npz archive with ``timestamps, states, controls`` (Logger.py:123-128);
* ``save_as_csv()``  per-signal two-column CSVs ``x0.csv, y0.csv, ... rpm3-0.csv, pwm3-0.csv`` (Logger.py:131-204).
"""
import os
from datetime import datetime

import numpy as np
import torch

_SIGNALS = [("x", 0), ("y", 1), ("z", 2), ("r", 6), ("p", 7), ("ya", 8), ("vx", 3), ("vy", 4), ("vz", 5),
            ("wx", 9), ("wy", 10), ("wz", 11)]


class TrajectoryLog:
    def __init__(self, log, logging_freq_hz, envs=None, output_folder="results", t0=0.0):
        """log: ``[T][16][N]`` tensor (device or host) or array; envs: indices of the envs to export (default: all --
        meant for a handful of envs, the files are per env)."""
        log = log.detach() if isinstance(log, torch.Tensor) else torch.as_tensor(np.asarray(log))
        if log.dim() != 3 or log.shape[1] != 16:
            raise ValueError("TrajectoryLog: expected a [T][16][N] Logger record")
        if envs is not None:
            log = log[:, :, torch.as_tensor(list(envs), device=log.device, dtype=torch.long)]
        self.LOGGING_FREQ_HZ = int(logging_freq_hz)
        self.OUTPUT_FOLDER = output_folder
        self.states = log.permute(2, 1, 0).to(torch.float64).cpu().numpy().copy()          # [num_drones, 16, T]
        self.NUM_DRONES, _, T = self.states.shape
        self.timestamps = np.tile(t0 + np.arange(T) / self.LOGGING_FREQ_HZ, (self.NUM_DRONES, 1))
        self.controls = np.zeros((self.NUM_DRONES, 12, T))
        self.counters = np.full(self.NUM_DRONES, T)

    def _stamp(self):
        return datetime.now().strftime("%m.%d.%Y_%H.%M.%S")

    def save(self):
        """Logger.save (Logger.py:123-128): np.savez into a file named ``save-flight-<date>.npy``."""
        os.makedirs(self.OUTPUT_FOLDER, exist_ok=True)
        path = os.path.join(self.OUTPUT_FOLDER, "save-flight-" + self._stamp() + ".npy")
        with open(path, "wb") as out_file:
            np.savez(out_file, timestamps=self.timestamps, states=self.states, controls=self.controls)
        return path

    def save_as_csv(self, comment=""):
        """Logger.save_as_csv (Logger.py:131-204): one two-column (t, value) CSV per signal and drone."""
        csv_dir = os.path.join(self.OUTPUT_FOLDER, "save-flight-" + comment + "-" + self._stamp())
        os.makedirs(csv_dir, exist_ok=True)
        f = self.LOGGING_FREQ_HZ
        t = np.arange(0, self.timestamps.shape[1] / f, 1 / f)[:self.states.shape[2]]

        def put(name, i, col):
            with open(os.path.join(csv_dir, name + str(i) + ".csv"), "wb") as out_file:
                np.savetxt(out_file, np.transpose(np.vstack([t, col])), delimiter=",")
        for i in range(self.NUM_DRONES):
            s = self.states[i]
            for name, row in _SIGNALS:
                put(name, i, s[row])
            for name, row in (("rr", 6), ("pr", 7), ("yar", 8)):   # finite-difference angle rates, 0 first
                put(name, i, np.hstack([0, (s[row, 1:] - s[row, 0:-1]) * f]))
            for j in range(4):
                put("rpm%d-" % j, i, s[12 + j])
                put("pwm%d-" % j, i, (s[12 + j] - 4070.3) / 0.2685)
        return csv_dir
