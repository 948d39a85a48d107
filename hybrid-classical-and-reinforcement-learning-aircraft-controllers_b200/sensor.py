"""Sensor models for the fleet: the reference's ``SensorInterface`` family (interfaces/sensor.py:17-263) over device
tensors.

``NoisySensorB200`` keeps ``NoisySensorInterface``'s surface -- ``update(true_state)``, ``get_state()``, ``reset()``,
``get_sensor_type()``, ``is_perfect()``, ``get_noise_parameters()`` and the same ``noise_config`` dict -- for N aircraft
at once; all arithmetic is ``acb200_sensor_update`` (csrc/sensor.cu).  The noise of aircraft ``g`` at update ``t`` is a
pure function of ``(seed, g, t)``, so it does not depend on how the fleet is sharded (``id_offset``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi
from .fleet import AircraftFleet, _dtype, _ptr

_DEFAULTS = dict(zip(_cabi.SENSOR_STDDEV_KEYS, (0.01, 0.1, 1.0, 0.1, 0.5, 0.5, 0.01)))   # sensor.py:176-182


class PerfectSensorB200:
    """``PerfectSensorInterface`` (sensor.py:105-134): hands back the fleet's own tensors."""

    def __init__(self):
        self._fleet = None

    def update(self, fleet: AircraftFleet) -> None:
        self._fleet = fleet

    def get_state(self):
        if self._fleet is None:
            raise RuntimeError("Sensor not yet updated with state")
        return self._fleet.state, self._fleet.derived

    def reset(self) -> None:
        self._fleet = None

    def get_sensor_type(self) -> str:
        return "perfect"

    def is_perfect(self) -> bool:
        return True


class NoisySensorB200:
    """``NoisySensorInterface`` for a whole ``AircraftFleet``.

    ``measured_state`` [13][n] (N,E,D,u,v,w,roll,pitch,yaw,p,q,r,time) and ``measured_derived`` [4][n] (airspeed,
    altitude, ground_speed, heading) are rewritten by every ``update``; ``bias`` [6][n] holds gyro_bias[3], accel_bias[3].
    """

    def __init__(self, noise_config: Optional[dict], n: int, precision: int = 64, device=0, id_offset: int = 0,
                 ctx: Optional[_cabi.Context] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("acb200 needs a CUDA device (no CPU fallback)")
        nc = dict(noise_config or {})
        self._config = nc
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.n, self.precision, self.dtype = int(n), precision, _dtype(precision)
        self.ctx = ctx or _cabi.Context(self.device.index or 0)
        c = _cabi.SensorConfigC()
        self.ctx.lib.acb200_default_sensor_config(C.byref(c))
        c.enabled = int(bool(nc.get("enabled", True)))
        for k in _cabi.SENSOR_STDDEV_KEYS:
            setattr(c, k, float(nc.get(k, _DEFAULTS[k])))
        seed = nc.get("seed")
        if seed is None:                     # default_rng(None): fresh OS entropy in the reference (sensor.py:189-190)
            seed = int.from_bytes(__import__("os").urandom(8), "little")
        c.seed, c.id_offset = int(seed) & 0xFFFFFFFFFFFFFFFF, int(id_offset)
        self.c = c
        kw = dict(device=self.device, dtype=self.dtype)
        self.bias = torch.zeros((6, n), **kw)
        self.measured_state = torch.zeros((13, n), **kw)
        self.measured_derived = torch.zeros((4, n), **kw)
        self.tick = 0
        self._updated = False

    # -- SensorInterface ---------------------------------------------------------------------------
    def update(self, fleet: AircraftFleet) -> None:
        """sensor.update(true_state) for every aircraft of ``fleet`` (sensor.py:200-243)."""
        if fleet.n != self.n or fleet.precision != self.precision:
            raise ValueError("sensor and fleet differ in size or precision")
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        self.ctx.check(self.ctx.lib.acb200_sensor_update(self.ctx.handle, C.byref(fleet.c), C.byref(self.c), _ptr(self.bias),
                                                         self.tick & 0xFFFFFFFF, _ptr(self.measured_state),
                                                         _ptr(self.measured_derived)))
        self.tick += 1
        self._updated = True

    def get_state(self):
        if not self._updated:
            raise RuntimeError("Sensor not yet updated with state")
        return self.measured_state, self.measured_derived

    def reset(self, mask: Optional[torch.Tensor] = None) -> None:
        """Zero the biases (sensor.py:245-251); ``mask`` restricts it to some aircraft (per-episode resets)."""
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        if mask is not None:
            mask = mask.to(self.device, torch.uint8).contiguous()
        self.ctx.check(self.ctx.lib.acb200_sensor_reset(self.ctx.handle, self.n, self.precision, _ptr(self.bias), _ptr(mask)))
        if mask is None:
            self._updated = False

    def get_sensor_type(self) -> str:
        return "noisy"

    def is_perfect(self) -> bool:
        return False

    def get_noise_parameters(self) -> dict:
        return {"enabled": bool(self.c.enabled), **{k: getattr(self.c, k) for k in _cabi.SENSOR_STDDEV_KEYS}}

    @property
    def gyro_bias(self) -> torch.Tensor:
        return self.bias[0:3]

    @property
    def accel_bias(self) -> torch.Tensor:
        return self.bias[3:6]

    def __repr__(self) -> str:
        return f"NoisySensorB200(type=noisy, n={self.n})"


class NoisySensorInterfaceB200:
    """One-aircraft drop-in for ``NoisySensorInterface``: takes and returns ``AircraftState`` dataclasses."""

    def __init__(self, noise_config: dict, precision: int = 64, device=0):
        self._fleet = AircraftFleet(1, precision=precision, device=device)
        self._sensor = NoisySensorB200(noise_config, 1, precision=precision, device=device, ctx=self._fleet.ctx)
        self._state = None

    def update(self, true_state) -> None:
        import numpy as np
        from dataclasses import replace
        f, s = self._fleet, self._sensor
        v = np.concatenate([np.asarray(true_state.position, float), np.asarray(true_state.velocity, float),
                            np.asarray(true_state.attitude, float), np.asarray(true_state.angular_rate, float),
                            [float(true_state.time)]])
        d = np.array([true_state.airspeed, true_state.altitude, true_state.ground_speed, true_state.heading], float)
        f.state.copy_(torch.as_tensor(v.reshape(13, 1)).to(f.dtype)); f.derived.copy_(torch.as_tensor(d.reshape(4, 1)).to(f.dtype))
        s.update(f)
        m = s.measured_state[:, 0].double().cpu().numpy(); md = s.measured_derived[:, 0].double().cpu().numpy()
        self._state = replace(true_state, position=m[0:3].copy(), velocity=m[3:6].copy(), attitude=m[6:9].copy(),
                              angular_rate=m[9:12].copy(), airspeed=float(md[0]), altitude=float(md[1]))

    def get_state(self):
        if self._state is None:
            raise RuntimeError("Sensor not yet updated with state")
        return self._state

    def reset(self) -> None:
        self._sensor.reset()
        self._state = None

    @property
    def _gyro_bias(self):
        return self._sensor.bias[0:3, 0].double().cpu().numpy()

    @property
    def _accel_bias(self):
        return self._sensor.bias[3:6, 0].double().cpu().numpy()

    def get_sensor_type(self) -> str:
        return "noisy"

    def is_perfect(self) -> bool:
        return False

    def get_noise_parameters(self) -> dict:
        return self._sensor.get_noise_parameters()
