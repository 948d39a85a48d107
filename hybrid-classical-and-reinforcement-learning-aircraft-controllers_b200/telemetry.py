"""Trajectory export for a sampled subset of the fleet, in the formats the reference's analysis tools read.

``TelemetryLogger`` (visualization/logger.py:9-156 in the reference) appends one Python dict per aircraft and step; here
the k sampled aircraft are gathered on the device (``acb200_telemetry_record``) into a ring ``[capacity][21][k]`` that
leaves the GPU once, and ``save`` writes the reference's layouts:

* JSON (logger.py:138-145): ``{"metadata": {id: {...}}, "data": {id: {"states": [{time, position, velocity, attitude,
  angular_rate, airspeed, altitude}], "commands": [{time, mode}], "surfaces": [{time, aileron, elevator, rudder,
  throttle}], "times": [...]}}}`` with ``indent=2`` -- what ``MultiAircraftReplay`` loads (visualization/replay.py:47-57);
* HDF5 (logger.py:126-137): one group per aircraft, metadata as string attributes, datasets ``times``, ``positions``,
  ``attitudes``.  ``h5py`` is optional exactly as in the reference: without it a ``.hdf5`` path falls back to JSON
  with the suffix replaced (logger.py:30-38).
"""
from __future__ import annotations

import ctypes as C
import json
from pathlib import Path
from typing import Dict, Optional, Sequence

import numpy as np

FIELDS = 21          # state 13 | derived 4 | controls 4 (include/acb200.h: ACB200_TELEMETRY_FIELDS)


def write_log(filepath: str, aircraft_ids: Sequence[str], records: np.ndarray, metadata: Optional[Dict[str, dict]] = None,
              command_mode: Optional[str] = None) -> Path:
    """records [T][21][k] (any float dtype) -> the reference's log file.  Returns the path written."""
    records = np.asarray(records, dtype=np.float64)
    T, nf, k = records.shape
    if nf != FIELDS or k != len(aircraft_ids):
        raise ValueError("records must be [T][21][len(aircraft_ids)]")
    path = Path(filepath)
    path.parent.mkdir(parents=True, exist_ok=True)
    metadata = metadata or {}
    use_hdf5 = str(filepath).endswith(".hdf5") or str(filepath).endswith(".h5")
    if use_hdf5:
        try:
            import h5py
        except ImportError:
            print("Warning: h5py not available, falling back to JSON logging")      # logger.py:34
            use_hdf5 = False
            path = path.with_suffix(".json")
    if use_hdf5:
        with h5py.File(path, "w") as f:
            for j, aid in enumerate(aircraft_ids):
                grp = f.create_group(aid)
                for key, val in (metadata.get(aid) or {}).items():
                    grp.attrs[key] = str(val)
                if T:
                    grp.create_dataset("times", data=records[:, 12, j])
                    grp.create_dataset("positions", data=records[:, 0:3, j])
                    grp.create_dataset("attitudes", data=records[:, 6:9, j])
        return path
    data = {}
    for j, aid in enumerate(aircraft_ids):
        r = records[:, :, j]
        states, surfaces, commands, times = [], [], [], []
        for t in range(T):
            tm = float(r[t, 12])
            states.append({"time": tm, "position": r[t, 0:3].tolist(), "velocity": r[t, 3:6].tolist(),
                           "attitude": r[t, 6:9].tolist(), "angular_rate": r[t, 9:12].tolist(), "airspeed": float(r[t, 13]),
                           "altitude": float(r[t, 14])})
            times.append(tm)
            # controls rows are elevator, aileron, rudder, throttle (ControlSurfaces.to_array); the log's key order is the logger's
            surfaces.append({"time": tm, "aileron": float(r[t, 18]), "elevator": float(r[t, 17]), "rudder": float(r[t, 19]),
                             "throttle": float(r[t, 20])})
            if command_mode is not None:
                commands.append({"time": tm, "mode": command_mode})
        data[aid] = {"states": states, "commands": commands, "surfaces": surfaces, "times": times}
    with open(path, "w") as f:
        json.dump({"metadata": {aid: (metadata.get(aid) or {}) for aid in aircraft_ids}, "data": data}, f, indent=2)
    return path


class TelemetryRecorderB200:
    """Samples ``ids`` (global indices into the fleet) every ``record()``; ``save(path)`` writes the log.

    ``capacity`` steps fit in the device ring (21 x k values each); recording past it raises, like a full disk would --
    sample fewer aircraft or save and ``clear()``."""

    def __init__(self, fleet, ids, capacity: int, names: Optional[Sequence[str]] = None, metadata: Optional[Dict[str, dict]] = None,
                 command_mode: Optional[str] = None):
        import torch
        from . import _cabi
        self._torch = torch
        self.fleet = fleet                     # AircraftFleet or BatchedRateControlEnv: needs .ctx, .state, .derived, .controls
        self.ctx = fleet.ctx
        dev = fleet.state.device
        self.ids = torch.as_tensor(ids, dtype=torch.int64, device=dev).contiguous()
        self.k, self.capacity = int(self.ids.numel()), int(capacity)
        self.names = list(names) if names is not None else [f"aircraft_{int(i):07d}" for i in self.ids.cpu().tolist()]
        if len(self.names) != self.k or len(set(self.names)) != self.k:
            raise ValueError("one distinct name per sampled aircraft")
        self.metadata = metadata or {}
        self.command_mode = command_mode
        self.ring = torch.zeros((self.capacity, FIELDS, self.k), dtype=fleet.state.dtype, device=dev)
        self.count = 0
        f = getattr(fleet, "c", None)
        self._fleet_c = f.fleet if hasattr(f, "fleet") else f          # EnvC embeds the FleetC
        if not isinstance(self._fleet_c, _cabi.FleetC):
            raise TypeError("fleet must be an AircraftFleet or a BatchedRateControlEnv")

    def record(self) -> None:
        if self.count >= self.capacity:
            raise RuntimeError("telemetry ring is full: save() and clear(), or sample fewer aircraft / steps")
        torch = self._torch
        self.ctx.set_stream(torch.cuda.current_stream(self.ring.device).cuda_stream)
        self.ctx.check(self.ctx.lib.acb200_telemetry_record(self.ctx.handle, C.byref(self._fleet_c), C.c_void_p(self.ids.data_ptr()),
                                                            self.k, C.c_void_p(self.ring[self.count].data_ptr())))
        self.count += 1

    def clear(self) -> None:
        self.count = 0

    def records(self) -> np.ndarray:
        """[count][21][k] float64 on the host (one device-to-host copy)."""
        return self.ring[: self.count].double().cpu().numpy()

    def save(self, filepath: str) -> Path:
        return write_log(filepath, self.names, self.records(), self.metadata, self.command_mode)
