"""Sensor noise on the GPU (acb200_sensor_update, csrc/sensor.cu) against the CPU restatement on the same
(seed, id, tick) stream, plus the behaviours tests/test_interfaces.py:343-470 of the reference checks."""
import numpy as np
import pytest
import torch

from helpers import sample_initial_states
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

# fp64 build: the only differences are libm vs CUDA log / sincospi (a few ulp of a normal <= 6.7) times sigma
F64_TOL = 1e-12
# fp32 build: float log / sincospif / sqrt and a float accumulate: 2e-6 of max(|value|, sigma)
F32_TOL = 2e-6


def _truth(n, seed=0):
    rng = np.random.default_rng(seed)
    s = sample_initial_states(rng, n)
    s[0:2] = rng.uniform(-500, 500, (2, n)); s[12] = 1.5
    d = np.stack([rng.uniform(12, 30, n), -s[2], rng.uniform(10, 30, n), rng.uniform(0, 2 * np.pi, n)])
    return s, d


def _fleet(n, precision, s, d):
    import acb200
    f = acb200.AircraftFleet(n, precision=precision)
    f.state.copy_(torch.as_tensor(s)); f.derived.copy_(torch.as_tensor(d))
    return f


NOISE = {"enabled": True, "imu_gyro_stddev": 0.02, "gps_position_stddev": 2.5, "attitude_stddev": 0.05, "seed": 42}


@pytest.mark.parametrize("precision", [64, 32])
def test_updates_match_the_restatement(precision):
    import acb200
    n, T = 1000, 40
    s, d = _truth(n)
    f = _fleet(n, precision, s, d)
    sensor = acb200.NoisySensorB200(NOISE, n, precision=precision, id_offset=123456789012)
    cfg = orc.sensor_config(NOISE, id_offset=123456789012)
    bias = np.zeros((6, n))
    tol = F64_TOL if precision == 64 else F32_TOL
    s_in = f.state.double().cpu().numpy(); d_in = f.derived.double().cpu().numpy()    # the truth as the build holds it
    for t in range(T):
        sensor.update(f)
        ms, md = orc.sensor_update(cfg, s_in, d_in, bias, t)
        got_s = sensor.measured_state.double().cpu().numpy(); got_d = sensor.measured_derived.double().cpu().numpy()
        scale_s = np.maximum(np.abs(ms), 1.0); scale_d = np.maximum(np.abs(md), 1.0)
        assert np.max(np.abs(got_s - ms) / scale_s) < tol * 8, (t, np.max(np.abs(got_s - ms) / scale_s))
        assert np.max(np.abs(got_d - md) / scale_d) < tol * 8
        # random-walk accumulation in float drifts with t; fp64 stays at rounding level
        btol = 1e-15 if precision == 64 else 2e-9 * (t + 4)
        assert np.max(np.abs(sensor.bias.double().cpu().numpy() - bias)) < btol + 1e-3 * tol
    assert torch.equal(sensor.measured_state[12], f.state[12])                    # time is copied
    assert torch.equal(sensor.measured_derived[2:], f.derived[2:])                # ground speed, heading are copied
    assert not torch.equal(sensor.measured_state[:12], f.state[:12])


def test_f32_and_f64_builds_draw_the_same_noise():
    import acb200
    n = 4096
    s, d = _truth(n, 3)
    out = {}
    for precision in (64, 32):
        f = _fleet(n, precision, s, d)
        sensor = acb200.NoisySensorB200(NOISE, n, precision=precision)
        for _ in range(5):
            sensor.update(f)
        out[precision] = ((sensor.measured_state - f.state).double().cpu().numpy(), sensor.bias.double().cpu().numpy())
    # positions ~500 m in float carry 3e-5 of rounding; compare the noise itself loosely and the biases tightly
    assert np.max(np.abs(out[64][0][:3] - out[32][0][:3])) < 1e-4
    assert np.max(np.abs(out[64][0][6:12] - out[32][0][6:12])) < 1e-6
    assert np.max(np.abs(out[64][1] - out[32][1])) < 2e-8


def test_disabled_returns_ground_truth():
    """tests/test_interfaces.py:389-404"""
    import acb200
    n = 257
    s, d = _truth(n, 1)
    f = _fleet(n, 64, s, d)
    sensor = acb200.NoisySensorB200({"enabled": False, "imu_gyro_stddev": 100.0}, n)
    with pytest.raises(RuntimeError):
        sensor.get_state()
    sensor.update(f)
    ms, md = sensor.get_state()
    assert torch.equal(ms, f.state) and torch.equal(md, f.derived)
    assert not sensor.bias.any()


def test_reproducible_with_seed_and_sharding_invariant():
    """tests/test_interfaces.py:406-427, plus: a shard [lo, hi) of the fleet sees the same noise as the whole fleet."""
    import acb200
    n, lo = 1024, 384
    s, d = _truth(n, 2)
    f = _fleet(n, 32, s, d)
    a = acb200.NoisySensorB200({"gps_position_stddev": 1.0, "seed": 999}, n, precision=32)
    b = acb200.NoisySensorB200({"gps_position_stddev": 1.0, "seed": 999}, n, precision=32)
    fs = _fleet(n - lo, 32, s[:, lo:], d[:, lo:])
    c = acb200.NoisySensorB200({"gps_position_stddev": 1.0, "seed": 999}, n - lo, precision=32, id_offset=lo)
    for _ in range(3):
        a.update(f); b.update(f); c.update(fs)
    assert torch.equal(a.measured_state, b.measured_state) and torch.equal(a.bias, b.bias)
    assert torch.equal(a.measured_state[:, lo:], c.measured_state) and torch.equal(a.bias[:, lo:], c.bias)
    other = acb200.NoisySensorB200({"gps_position_stddev": 1.0, "seed": 1000}, n, precision=32)
    other.update(f)
    assert not torch.allclose(other.measured_state[:3], a.measured_state[:3])


def test_bias_drifts_and_reset_clears_it():
    """tests/test_interfaces.py:429-468"""
    import acb200
    n = 512
    f = acb200.AircraftFleet(n, precision=64)
    sensor = acb200.NoisySensorB200({"enabled": True, "seed": 42}, n)
    for _ in range(100):
        sensor.update(f)
    gb, ab = sensor.gyro_bias.cpu().numpy(), sensor.accel_bias.cpu().numpy()
    assert abs(gb.std() / 1e-3 - 1.0) < 0.15 and abs(ab.std() / 1e-2 - 1.0) < 0.15      # step * sqrt(100)
    mask = torch.zeros(n, dtype=torch.uint8); mask[::2] = 1
    sensor.reset(mask)
    assert not sensor.bias[:, ::2].any() and sensor.bias[:, 1::2].any()
    sensor.reset()
    assert not sensor.bias.any()
    with pytest.raises(RuntimeError):
        sensor.get_state()


def test_noise_statistics_on_a_large_fleet():
    """2^20 aircraft, one update on a zero truth: per-channel standard deviation within 5 sigma of sampling error of the
    configured value (the reference's defaults, sensor.py:176-182), means within 5 sigma of zero."""
    import acb200
    n = 1 << 20
    f = acb200.AircraftFleet(n, precision=32)
    f.state.zero_(); f.derived.zero_()
    sensor = acb200.NoisySensorB200({"seed": 1}, n, precision=32)
    sensor.update(f)
    noise = torch.cat([sensor.measured_state[:12], sensor.measured_derived[:2]]).double()
    want = torch.tensor([1.0] * 3 + [0.1] * 3 + [0.01] * 3 + [0.01] * 3 + [0.5, 0.5], dtype=torch.float64, device=noise.device)
    assert torch.all((noise.std(dim=1) / want - 1.0).abs() < 5.0 / np.sqrt(2 * n))
    assert torch.all(noise.mean(dim=1).abs() < 5.0 * want / np.sqrt(n))
    assert float(noise.abs().max() / 1.0) < 7.0


def test_single_aircraft_drop_in_and_argument_errors():
    """tests/test_interfaces.py:346-387 through the dataclass interface; bad stddev -> ValueError like Generator.normal."""
    import acb200
    from acb200.adapters import AircraftState
    sensor = acb200.NoisySensorInterfaceB200({"enabled": True, "imu_gyro_stddev": 0.1, "gps_position_stddev": 5.0,
                                              "attitude_stddev": 0.05, "airspeed_stddev": 1.0, "seed": 123})
    assert sensor.get_sensor_type() == "noisy" and not sensor.is_perfect()
    assert sensor.get_noise_parameters()["gps_position_stddev"] == 5.0
    truth = AircraftState(position=np.array([100.0, 200.0, -150.0]), velocity=np.array([20.0, 5.0, -2.0]),
                          attitude=np.array([0.1, 0.2, 0.0]), angular_rate=np.array([0.05, 0.01, 0.02]), airspeed=25.0,
                          altitude=150.0)
    with pytest.raises(RuntimeError):
        sensor.get_state()
    sensor.update(truth)
    m = sensor.get_state()
    assert not np.allclose(m.position, truth.position, atol=0.1)
    assert not np.allclose(m.attitude, truth.attitude, atol=0.01)
    # against the restatement for global id 0, tick 0
    cfg = orc.sensor_config({"imu_gyro_stddev": 0.1, "gps_position_stddev": 5.0, "attitude_stddev": 0.05,
                             "airspeed_stddev": 1.0, "seed": 123})
    s = np.concatenate([truth.position, truth.velocity, truth.attitude, truth.angular_rate, [0.0]]).reshape(13, 1)
    ms, md = orc.sensor_update(cfg, s, np.array([[25.0], [150.0], [0.0], [0.0]]), np.zeros((6, 1)), 0)
    assert np.allclose(m.position, ms[0:3, 0], rtol=0, atol=1e-10) and abs(m.airspeed - md[0, 0]) < 1e-10
    for _ in range(50):
        sensor.update(truth)
    assert np.any(sensor._gyro_bias != 0)
    sensor.reset()
    assert np.allclose(sensor._gyro_bias, 0) and np.allclose(sensor._accel_bias, 0)
    bad = acb200.NoisySensorB200({"gps_position_stddev": -1.0}, 4)
    with pytest.raises(ValueError):
        bad.update(acb200.AircraftFleet(4))
    with pytest.raises(ValueError):
        acb200.NoisySensorB200({}, 4).update(acb200.AircraftFleet(8))
