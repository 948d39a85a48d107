"""Telemetry export.  CPU: the writer against the file the UNMODIFIED reference TelemetryLogger wrote
for the same records (tests/golden/telemetry.json, made by tests/golden/make_golden.py telemetry) -- byte for byte -- and the
loading side (visualization/replay.py:47-57 reads "data" / "metadata").  GPU: the device ring against the fleet's own
tensors along a flown mission."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

IDS = ["lead", "wing_left", "wing_right"]
META = {"lead": {"type": "rc_plane", "formation_index": 0}, "wing_left": {"type": "rc_plane", "formation_index": 1}}


def test_json_log_equals_the_reference_loggers_file(tmp_path):
    from acb200 import telemetry
    rec = np.load(os.path.join(GOLDEN, "telemetry_records.npy"))
    out = telemetry.write_log(str(tmp_path / "log.json"), IDS, rec, META, command_mode="WAYPOINT")
    assert out.read_text() == open(os.path.join(GOLDEN, "telemetry.json")).read()
    # what MultiAircraftReplay._load_data does with it
    content = json.load(open(out))
    data = content.get("data", {})
    assert list(data) == IDS and content["metadata"]["wing_right"] == {}
    t = data["lead"]["times"]
    assert len(t) == rec.shape[0] and abs((t[-1] - t[0]) - 0.05) < 1e-12
    assert data["wing_left"]["surfaces"][2]["elevator"] == rec[2, 17, 1] and data["wing_left"]["surfaces"][2]["aileron"] == rec[2, 18, 1]


def test_hdf5_path_falls_back_to_json_like_the_reference(tmp_path, capsys):
    from acb200 import telemetry
    try:
        import h5py  # noqa: F401
        pytest.skip("h5py is installed here: the HDF5 branch is taken instead")
    except ImportError:
        pass
    rec = np.load(os.path.join(GOLDEN, "telemetry_records.npy"))
    out = telemetry.write_log(str(tmp_path / "sub" / "log.hdf5"), IDS, rec, META, command_mode="WAYPOINT")
    assert out.suffix == ".json" and out.exists() and "falling back to JSON" in capsys.readouterr().out   # logger.py:30-38
    assert out.read_text() == open(os.path.join(GOLDEN, "telemetry.json")).read()


def test_writer_rejects_mismatched_shapes(tmp_path):
    from acb200 import telemetry
    with pytest.raises(ValueError):
        telemetry.write_log(str(tmp_path / "x.json"), IDS, np.zeros((4, 20, 3)))
    with pytest.raises(ValueError):
        telemetry.write_log(str(tmp_path / "x.json"), IDS[:2], np.zeros((4, 21, 3)))
    out = telemetry.write_log(str(tmp_path / "empty.json"), IDS, np.zeros((0, 21, 3)))
    assert json.load(open(out))["data"]["lead"] == {"states": [], "commands": [], "surfaces": [], "times": []}


@pytest.mark.gpu
@pytest.mark.parametrize("precision", [64, 32])
def test_recorder_follows_the_fleet(tmp_path, precision):
    import torch
    import acb200
    n, steps = 4096, 25
    fleet = acb200.AircraftFleet(n, precision=precision, config={"dt_physics": 0.01})
    fleet.set_mission([[300.0, 0.0, -100.0, 15.0], [300.0, 300.0, -100.0, 15.0]], acceptance_radius=40.0, guidance="PP")
    ids = [0, 17, 2048, 4095]
    rec = acb200.TelemetryRecorderB200(fleet, ids, capacity=steps, command_mode="WAYPOINT",
                                       metadata={"aircraft_0000017": {"type": "rc_plane"}})
    want = []
    l0 = fleet.ctx.launch_count
    for _ in range(steps):
        fleet.step_waypoint(0.01)
        rec.record()
        want.append(torch.cat([fleet.state[:, ids], fleet.derived[:, ids], fleet.controls[:, ids]]).double().cpu().numpy())
    assert fleet.ctx.launch_count - l0 == 2 * steps                 # one gather launch per record
    got = rec.records()
    assert got.shape == (steps, 21, 4) and np.array_equal(got, np.stack(want))
    with pytest.raises(RuntimeError):
        rec.record()                                                # ring full
    out = rec.save(str(tmp_path / "mission.json"))
    data = json.load(open(out))
    assert list(data["data"]) == ["aircraft_0000000", "aircraft_0000017", "aircraft_0002048", "aircraft_0004095"]
    s = data["data"]["aircraft_0002048"]
    assert len(s["states"]) == steps and s["commands"][0]["mode"] == "WAYPOINT"
    assert s["states"][-1]["position"] == got[-1, 0:3, 2].tolist() and s["states"][-1]["altitude"] == got[-1, 14, 2]
    assert abs(s["times"][-1] - steps * 0.01) < 1e-5
    assert data["metadata"]["aircraft_0000017"] == {"type": "rc_plane"}
    rec.clear(); rec.record()
    assert rec.count == 1


@pytest.mark.gpu
def test_recorder_on_the_env_and_bad_ids():
    import torch
    import acb200
    env = acb200.BatchedRateControlEnv(1024, controller="rate_pid", precision=32)
    env.reset()
    rec = acb200.TelemetryRecorderB200(env, [5, 1023, 5000, -1], capacity=3, names=["a", "b", "outside", "negative"])
    env.step(); rec.record()
    got = rec.records()[0]
    assert np.array_equal(got[:13, 0], env.state[:, 5].double().cpu().numpy())
    assert np.array_equal(got[17:, 1], env.controls[:, 1023].double().cpu().numpy())
    assert np.isnan(got[:, 2]).all() and np.isnan(got[:, 3]).all()     # ids outside the fleet log NaN, they do not fault
    with pytest.raises(ValueError):
        acb200.TelemetryRecorderB200(env, [1, 2], capacity=1, names=["x", "x"])
