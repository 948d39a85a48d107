"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/acb200.h declares,
its descriptors match the header, and its host PID core reproduces the reference's known answers."""
import ctypes as C
import os
import re

import numpy as np

from conftest import GOLDEN, ROOT


def _lib():
    import acb200
    return acb200, acb200._cabi.load()


def test_library_exports_every_declared_symbol():
    acb200, lib = _lib()
    hdr = open(os.path.join(ROOT, "include", "acb200.h")).read()
    declared = set(re.findall(r"\b(acb200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(acb200._cabi.SYMBOLS), declared ^ set(acb200._cabi.SYMBOLS)
    for s in declared:
        assert getattr(lib, s) is not None
    assert lib.acb200_version() == b"0.1.0"


def test_struct_sizes_match_header():
    acb200, lib = _lib()
    cab = acb200._cabi
    assert C.sizeof(cab.AircraftParamsC) == 39 * 8
    assert C.sizeof(cab.PidConfigC) == 32 and C.sizeof(cab.PidStateC) == 24
    assert C.sizeof(cab.ControllerParamsC) == 9 * 32 + 20 * 8 and C.sizeof(cab.MissionC) == 48          # + the statistics pointer (r02)
    assert C.sizeof(cab.FleetC) == 8 + 8 + 4 * 8
    assert C.sizeof(cab.EnvC) == C.sizeof(cab.FleetC) + 19 * 8
    assert C.sizeof(cab.EnvConfigC) == 64


def test_default_params_equal_oracle_and_reference_values():
    acb200, lib = _lib()
    from oracle import oracle as orc
    p, o = acb200.default_aircraft_params(), orc.default_aircraft_params()
    for f in acb200._cabi.AIRCRAFT_FIELDS:
        assert getattr(p, f) == getattr(o, f), f
    assert (p.mass, p.cl_alpha, p.max_thrust, p.max_pitch_angle) == (8.0, 5.0, 50.0, 85.0)
    c, oc = acb200.default_controller_params(), orc.default_controller_params()
    assert bytes(c)[: 6 * 32] == bytes(oc)[: 6 * 32]
    assert c.max_yaw_rate == oc.max_yaw_rate == np.radians(160.0)
    cessna = acb200.default_aircraft_params("cessna")
    assert (cessna.mass, cessna.wing_span, cessna.max_thrust) == (15.0, 3.0, 80.0)


def test_host_pid_known_answers_bit_exact():
    acb200, lib = _lib()
    cab = acb200._cabi
    g = np.load(os.path.join(GOLDEN, "pid_kat.npz"))
    for ci in range(g["cfg"].shape[0]):
        kp, ki, kd, omin, omax, imin, imax, alpha = [float(x) for x in g["cfg"][ci]]
        cfg = cab.PidConfigC(kp, ki, kd, omin, omax, imin, imax, alpha)
        st = cab.PidStateC()
        for t in range(g["sp"].shape[1]):
            if np.isnan(g["sp"][ci, t]):
                lib.acb200_pid_reset(C.byref(st))
                assert (st.error, st.integral, st.output) == (0.0, 0.0, 0.0)
                continue
            o = lib.acb200_pid_compute_f32(C.byref(cfg), C.byref(st), g["sp"][ci, t], g["meas"][ci, t], g["dt"][ci, t])
            got = np.array([o, st.error, st.integral, st.derivative_filtered], np.float32)
            exp = g["out"][ci, t]
            assert np.array_equal(got, exp, equal_nan=True), (ci, t, got, exp)


def test_no_gpu_means_loud_failure():
    import torch
    acb200, lib = _lib()
    if torch.cuda.is_available():
        return
    import pytest
    with pytest.raises(RuntimeError):
        acb200.Context(0)
    with pytest.raises(RuntimeError):
        acb200.AircraftFleet(4)
    with pytest.raises(RuntimeError):
        acb200.BatchedRateControlEnv(4)


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "hybrid-classical-and-reinforcement-learning-aircraft-controllers_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "acb_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f
