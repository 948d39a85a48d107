"""Sensor noise, CPU side: the restatement of NoisySensorInterface.update (oracle/acb_oracle.c
orc_sensor_update) against the reference's own outputs, and the Philox/Box-Muller normal stream against the reference's
noise statistics.  Fixture: tests/golden/sensor.npz, written by tests/golden/make_golden.py sensor from the unmodified
interfaces/sensor.py."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import oracle as orc

CASES = ["default_seed42", "custom_seed123", "gps_seed999", "disabled"]
KEYS = orc.SENSOR_STDDEV_KEYS


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(GOLDEN, "sensor.npz"))


def _cfg(params):
    nc = {"enabled": bool(params[0]), "seed": int(params[8])}
    nc.update({k: float(v) for k, v in zip(KEYS, params[1:8])})
    return orc.sensor_config(nc)


@pytest.mark.parametrize("case", CASES)
def test_restatement_equals_reference_bit_for_bit(golden, case):
    """Same standard normals in (the reference's PCG64 draws, stored in the fixture) -> same measured states and the same
    bias history out, bit for bit: addition order, bias-before-walk, walk scales (sensor.py:209-231)."""
    truth, meas, bias_ref, z = (golden[f"{case}/{k}"] for k in ("truth", "measured", "bias", "z"))
    cfg = _cfg(golden[f"{case}/params"])
    bias = np.zeros((6, 1))
    for t in range(truth.shape[0]):
        s = np.ascontiguousarray(truth[t, :13].reshape(13, 1)); d = np.ascontiguousarray(truth[t, 13:].reshape(4, 1))
        ms, md = orc.sensor_update(cfg, s, d, bias, t, z=np.ascontiguousarray(z[t].reshape(20, 1)))
        got = np.concatenate([ms[:12, 0], [ms[12, 0]], md[:, 0]])
        assert np.array_equal(got, meas[t]), (case, t, np.abs(got - meas[t]).max())
        assert np.array_equal(bias[:, 0], bias_ref[t]), (case, t)
    if case == "disabled":
        assert np.array_equal(meas, truth) and not bias.any()


def test_fixture_normals_are_numpy_pcg64(golden):
    """The stored normals are what default_rng(seed) yields here too (same numpy stream as the reference run)."""
    z = golden["default_seed42/z"]
    live = np.random.default_rng(42).standard_normal(z.shape)
    if not np.array_equal(live, z):
        pytest.skip("this numpy's PCG64 normal stream differs from the one the fixture was made with")


def test_philox_normals_are_standard_normal():
    """20 channels x 200k draws: mean, variance, kurtosis, tails and cross-channel correlation of the kernel's stream."""
    ids = np.arange(4000); ticks = np.arange(50)
    z = np.array([orc.sensor_normals(11, int(i), int(t)) for i in ids for t in ticks])
    m = z.shape[0]
    assert np.all(np.abs(z.mean(0)) < 5.0 / np.sqrt(m))
    assert np.all(np.abs(z.var(0) - 1.0) < 5.0 * np.sqrt(2.0 / m))
    assert np.all(np.abs((z ** 4).mean(0) - 3.0) < 5.0 * np.sqrt(96.0 / m))
    assert abs(np.mean(np.abs(z) > 1.959964) - 0.05) < 5.0 * np.sqrt(0.05 * 0.95 / z.size)
    corr = np.corrcoef(z.T) - np.eye(20)
    assert np.abs(corr).max() < 5.0 / np.sqrt(m)
    # successive ticks and neighbouring aircraft are uncorrelated too
    zz = z.reshape(len(ids), len(ticks), 20)
    assert abs(np.mean(zz[:, 1:, 0] * zz[:, :-1, 0])) < 5.0 / np.sqrt(zz[:, 1:, 0].size)
    assert abs(np.mean(zz[1:, :, 0] * zz[:-1, :, 0])) < 5.0 / np.sqrt(zz[1:, :, 0].size)


def test_noise_statistics_match_the_reference(golden):
    """The fleet restatement at the default config, 4000 aircraft x 64 updates on a zero truth, against the per-channel
    standard deviations measured on 4000 reference sensors: measurement noise and the random-walk spread of both biases
    agree within sampling error (two independent estimates: 5 sigma of sqrt(2) * s / sqrt(2R))."""
    R, U = int(golden["stats/sensors"][0]), int(golden["stats/updates"][0])
    cfg = orc.sensor_config({"enabled": True, "seed": 77})
    state = np.zeros((13, R)); der = np.zeros((4, R)); bias = np.zeros((6, R))
    for t in range(U):
        ms, md = orc.sensor_update(cfg, state, der, bias, t)
    noise = np.concatenate([ms[:12], md[:2]])
    tol = 5.0 / np.sqrt(R)
    ref_std = golden["stats/noise_std"]
    assert np.all(np.abs(noise.std(axis=1) / ref_std - 1.0) < tol), noise.std(axis=1) / ref_std
    assert np.all(np.abs(bias.std(axis=1) / golden["stats/walk_std"] - 1.0) < tol)
    assert np.all(np.abs(noise.mean(axis=1)) < 5.0 * ref_std / np.sqrt(R))
    # and against the closed forms: sigma per channel; walk = step * sqrt(U)
    want = np.array([1.0] * 3 + [0.1] * 3 + [0.01] * 3 + [np.sqrt(0.01 ** 2 + (U - 1) * 1e-8)] * 3 + [0.5, 0.5])
    assert np.all(np.abs(noise.std(axis=1) / want - 1.0) < tol)
    assert np.all(np.abs(bias.std(axis=1) / (np.array([1e-4] * 3 + [1e-3] * 3) * np.sqrt(U)) - 1.0) < tol)


def test_noise_is_a_function_of_seed_id_tick_only():
    """Sharding invariance on the CPU side: aircraft g's noise is the same in a fleet of 64 and in shard [32,64)."""
    rng = np.random.default_rng(0)
    state = rng.normal(size=(13, 64)); der = rng.normal(size=(4, 64))
    full = orc.sensor_config({"seed": 5}); shard = orc.sensor_config({"seed": 5}, id_offset=32)
    b0 = np.zeros((6, 64)); b1 = np.zeros((6, 32))
    for t in range(3):
        ms0, md0 = orc.sensor_update(full, state, der, b0, t)
        ms1, md1 = orc.sensor_update(shard, np.ascontiguousarray(state[:, 32:]), np.ascontiguousarray(der[:, 32:]), b1, t)
    assert np.array_equal(ms0[:, 32:], ms1) and np.array_equal(md0[:, 32:], md1) and np.array_equal(b0[:, 32:], b1)
    other = orc.sensor_config({"seed": 6})
    ms2, _ = orc.sensor_update(other, state, der, np.zeros((6, 64)), 2)
    assert not np.allclose(ms2[:3], ms0[:3])


def test_bad_stddev_raises_like_generator_normal():
    cfg = orc.sensor_config({"gps_position_stddev": -1.0})
    with pytest.raises(ValueError):
        orc.sensor_update(cfg, np.zeros((13, 1)), np.zeros((4, 1)), np.zeros((6, 1)), 0)
