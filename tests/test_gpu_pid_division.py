"""The PID derivative's division on the device: (error - error_prev) / dt is evaluated as a fused multiply-add sequence on a
host-side reciprocal (csrc/pid.cuh, PidDivFused) and must equal IEEE division bit for bit -- the reference's
PIDController::compute divides (cpp/src/pid_controller.cpp:37-41).  Every exponent, both signs, zeros, subnormals, the window
edges at 2^-100 / 2^100, infinities and NaN go through the roll-rate loop of a real fleet step; the quotient is read back
from the controller memory (derivative filter alpha = 1, previous filtered value -0 so that the sign of a zero quotient
survives the filter's addition)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def numerators(rng):
    exps = np.arange(0, 256, dtype=np.uint32)
    man = np.concatenate([[0, 1, 0x7FFFFF, 0x400000, 0x555555], rng.integers(0, 1 << 23, 27)]).astype(np.uint32)
    grid = ((exps[:, None] << 23) | man[None, :]).ravel()
    dense = rng.integers(0, 1 << 31, 150_000, dtype=np.uint32)                       # any magnitude
    flight = (rng.integers(96, 140, 150_000).astype(np.uint32) << 23) | rng.integers(0, 1 << 23, 150_000).astype(np.uint32)
    edges = np.array([0x0D7FFFFF, 0x0D800000, 0x0D800001, 0x717FFFFF, 0x71800000, 0x71800001, 0x7F7FFFFF, 0x7F800000, 0x7FC00000,
                      0x00000001, 0x007FFFFF, 0x00800000, 0], np.uint32)
    pos = np.concatenate([grid, dense, flight, edges])
    bits = np.concatenate([pos, pos | np.uint32(0x80000000)])
    if bits.size & 1:
        bits = np.concatenate([bits, bits[:1]])
    return bits


@pytest.mark.parametrize("precision,pair", [(32, "0"), (32, "1"), (64, "0")])      # one-aircraft / pair kernel of the fp32 build, fp64 build
@pytest.mark.parametrize("dt", [0.01, 0.02, 0.001, 1.0 / 60.0, 0.005, 1e-4, 2e-6, 0.3, 16.0, 20.0, 1000.0])
def test_pid_quotient_equals_ieee_division(dt, precision, pair):
    import torch
    import acb200
    bits = numerators(np.random.default_rng(17))
    n = bits.size
    x = bits.view(np.float32)
    cp = acb200.default_controller_params()
    for k, v in (("kp", 0.0), ("ki", 0.0), ("kd", 0.0), ("derivative_filter_alpha", 1.0)):
        setattr(cp.rate[0], k, v)
    os.environ["ACB200_ENV_PAIR"] = pair
    try:
        fleet = acb200.AircraftFleet(n, precision=precision, config={"controller_params": cp})
        fleet.reset()
        # error_prev of the roll-rate loop: error (0 - 0) - (-x) = x; a numerator of -0 needs the error itself to be -0
        # (a roll-rate command of -0 against a roll rate of +0) and error_prev = +0
        zero = x == 0
        fleet.pid[1] = torch.as_tensor(np.where(zero, np.float32(0.0), -x), device="cuda")
        fleet.pid[2] = -0.0
        cmd = torch.zeros(4, n, device="cuda", dtype=fleet.dtype)
        cmd[0] = torch.as_tensor(np.where(zero & np.signbit(x), -0.0, 0.0), device="cuda", dtype=fleet.dtype)
        cmd[3] = 0.5
        fleet.step_rate(cmd, 0.01, dt_ctrl=dt)
        torch.cuda.synchronize()
        got = fleet.pid[2].cpu().numpy()
    finally:
        os.environ.pop("ACB200_ENV_PAIR", None)
    with np.errstate(all="ignore"):
        want = x / np.float32(dt)
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    bad = (got.view(np.uint32) != want.view(np.uint32)) & ~nan
    assert not bad.any(), f"{int(bad.sum())} quotients differ, first x = {x[bad][0]!r}: got {got[bad][0]!r}, IEEE {want[bad][0]!r}"
