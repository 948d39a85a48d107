"""Drop-in for the reference's pybind11 module ``aircraft_controls_bindings`` (cpp/bindings/bindings.cpp:13-164):
same class names, constructor signatures, keyword names, defaults, ``__repr__`` shapes and ``__version__``.

The arithmetic is the library's own fp32 PID core through the C ABI (``acb200_pid_compute_f32`` /
``acb200_pid_reset``, csrc/pid.cuh) -- the very code the CUDA kernels inline -- so a sequence fed to this module
and to the device agents gives bit-identical floats.  Values cross the boundary the way pybind11 moves them:
Python floats are narrowed to C ``float`` on entry and widened on return.
"""
from __future__ import annotations

import ctypes as C

from . import _cabi

__version__ = "1.1.0"          # bindings.cpp:163


def _f32(x) -> float:
    return C.c_float(x).value


class PIDGains:
    """cpp/include/pid_controller.h:22-29"""

    def __init__(self, kp: float = 0.0, ki: float = 0.0, kd: float = 0.0):
        self.kp, self.ki, self.kd = kp, ki, kd

    kp = property(lambda s: s._kp, lambda s, v: setattr(s, "_kp", _f32(v)))
    ki = property(lambda s: s._ki, lambda s, v: setattr(s, "_ki", _f32(v)))
    kd = property(lambda s: s._kd, lambda s, v: setattr(s, "_kd", _f32(v)))

    def __repr__(self):
        return f"PIDGains(kp={self.kp:f}, ki={self.ki:f}, kd={self.kd:f})"


class PIDConfig:
    """cpp/include/pid_controller.h:34-49: output [-1, 1], integral [-10, 10], derivative_filter_alpha 0.1"""
    _FIELDS = ("output_min", "output_max", "integral_min", "integral_max", "derivative_filter_alpha")

    def __init__(self):
        self.gains = PIDGains()
        self.output_min, self.output_max = -1.0, 1.0
        self.integral_min, self.integral_max = -10.0, 10.0
        self.derivative_filter_alpha = 0.1

    def __setattr__(self, k, v):
        object.__setattr__(self, k, _f32(v) if k in self._FIELDS else v)

    def _c(self) -> _cabi.PidConfigC:
        g = self.gains
        return _cabi.PidConfigC(g.kp, g.ki, g.kd, self.output_min, self.output_max, self.integral_min, self.integral_max,
                                self.derivative_filter_alpha)

    def _copy(self) -> "PIDConfig":
        c = PIDConfig()
        c.gains = PIDGains(self.gains.kp, self.gains.ki, self.gains.kd)
        for f in self._FIELDS:
            setattr(c, f, getattr(self, f))
        return c

    def __repr__(self):
        return (f"PIDConfig(gains={self.gains!r}, output=[{self.output_min:f}, {self.output_max:f}], "
                f"integral=[{self.integral_min:f}, {self.integral_max:f}], alpha={self.derivative_filter_alpha:f})")


class PIDController:
    """cpp/src/pid_controller.cpp:10-82"""

    def __init__(self, config: PIDConfig | None = None):
        self._cfg = (config or PIDConfig())._copy()      # the C++ constructor copies the struct
        self._st = _cabi.PidStateC()
        self._lib = _cabi.load()

    def compute(self, setpoint: float, measurement: float, dt: float) -> float:
        c = self._cfg._c()
        return self._lib.acb200_pid_compute_f32(C.byref(c), C.byref(self._st), setpoint, measurement, dt)

    def reset(self) -> None:
        self._lib.acb200_pid_reset(C.byref(self._st))

    def set_gains(self, gains: PIDGains) -> None:
        self._cfg.gains = PIDGains(gains.kp, gains.ki, gains.kd)

    def get_gains(self) -> PIDGains:
        g = self._cfg.gains
        return PIDGains(g.kp, g.ki, g.kd)

    def get_error(self) -> float:
        return self._st.error

    def get_integral(self) -> float:
        return self._st.integral

    def get_derivative(self) -> float:
        return self._st.derivative_filtered      # getDerivative() returns the filtered term (pid_controller.h:113)

    def get_output(self) -> float:
        return self._st.output

    def __repr__(self):
        g = self._cfg.gains
        return f"PIDController(kp={g.kp:f}, ki={g.ki:f}, kd={g.kd:f})"


class Vector3:
    """cpp/include/pid_controller.h:142-149"""

    def __init__(self, x: float = 0.0, y: float = 0.0, z: float = 0.0):
        self.x, self.y, self.z = _f32(x), _f32(y), _f32(z)

    def __repr__(self):
        return f"Vector3(x={self.x:f}, y={self.y:f}, z={self.z:f})"


class ControlOutput:
    """cpp/include/pid_controller.h:154-161"""

    def __init__(self, roll: float = 0.0, pitch: float = 0.0, yaw: float = 0.0):
        self.roll, self.pitch, self.yaw = _f32(roll), _f32(pitch), _f32(yaw)

    def __repr__(self):
        return f"ControlOutput(roll={self.roll:f}, pitch={self.pitch:f}, yaw={self.yaw:f})"


class MultiAxisPIDController:
    """cpp/src/pid_controller.cpp:88-152: three independent PIDController (roll, pitch, yaw)"""

    def __init__(self, roll_config: PIDConfig | None = None, pitch_config: PIDConfig | None = None,
                 yaw_config: PIDConfig | None = None):
        self._axes = [PIDController(roll_config), PIDController(pitch_config), PIDController(yaw_config)]

    def compute(self, setpoint: Vector3, measurement: Vector3, dt: float) -> ControlOutput:
        r, p, y = self._axes
        return ControlOutput(r.compute(setpoint.x, measurement.x, dt), p.compute(setpoint.y, measurement.y, dt),
                             y.compute(setpoint.z, measurement.z, dt))

    def reset(self) -> None:
        for a in self._axes:
            a.reset()

    def set_gains(self, axis: int, gains: PIDGains) -> None:
        if axis in (0, 1, 2):                      # invalid axis: do nothing (pid_controller.cpp:120-127)
            self._axes[axis].set_gains(gains)

    def get_gains(self, axis: int) -> PIDGains:
        return self._axes[axis].get_gains() if axis in (0, 1, 2) else PIDGains()   # pid_controller.cpp:129-136

    def get_error(self) -> Vector3:
        return Vector3(*(a.get_error() for a in self._axes))

    def get_integral(self) -> Vector3:
        return Vector3(*(a.get_integral() for a in self._axes))

    def get_output(self) -> ControlOutput:
        return ControlOutput(*(a.get_output() for a in self._axes))

    def __repr__(self):
        k = [a.get_gains().kp for a in self._axes]
        return f"MultiAxisPIDController(roll_kp={k[0]:f}, pitch_kp={k[1]:f}, yaw_kp={k[2]:f})"
