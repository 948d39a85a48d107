"""acb200.flight_episode against an episode the unmodified reference logged (tests/golden/flight_episode.json, made by
tests/golden/make_golden.py::gen_flight_episode with FlightLoggerCallback._extract_flight_data + FlightLogger): same step
records, same episode dictionary (wall-clock fields aside), same TensorBoard summary (tag, plugin name, JSON content)."""
import json
import os

import numpy as np
import pytest

import acb200
from acb200.flight_episode import FlightEpisodeWriterB200, flight_step_record, read_flight_events
from conftest import GOLDEN


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(GOLDEN, "flight_episode.json")))


def _info(d):
    info = dict(d)
    info["position"] = np.asarray(info["position"])
    return info


def test_step_records_equal_the_reference_callback(golden):
    for inp, exp in zip(golden["inputs"], golden["records"]):
        rec = flight_step_record(np.asarray(inp["obs"], np.float32), np.asarray(inp["action"], np.float32), inp["reward"],
                                 _info(inp["info"]), inp["step"], "golden_agent")
        assert json.loads(json.dumps(rec, default=lambda o: o.tolist())) == exp
    assert flight_step_record(np.zeros(5), np.zeros(4), 0.0, {}, 0) is None            # callbacks.py:176-177


def test_episode_equals_the_reference_logger(golden, tmp_path):
    w = FlightEpisodeWriterB200(tmp_path, agent_id="golden_agent")
    for inp in golden["inputs"]:
        w.log_step(np.asarray(inp["obs"], np.float32), np.asarray(inp["action"], np.float32), inp["reward"], _info(inp["info"]), inp["step"])
    ep = w.end_episode(success=False, termination_reason="crash_altitude", config={"episode_num": 7}, tags=["training_step_1040"])
    w.close()
    exp = golden["episode"]["content"]
    wall = ("start_time", "duration")
    assert {k: v for k, v in ep.items() if k not in wall} == {k: v for k, v in exp.items() if k not in wall}
    assert list(ep) == list(exp)                                                        # same key order in the JSON
    assert ep["trajectory"][-1]["events"][0]["event_type"] == "termination" and ep["trajectory"][-1]["events"][0]["severity"] == "error"
    # the plain-file copy and (tensorboard is installed here) the event-file record the plugin reads
    lines = open(tmp_path / "flight_episodes.jsonl").read().splitlines()
    assert len(lines) == 1 and json.loads(lines[0]) == ep
    pytest.importorskip("tensorboard")
    got = read_flight_events(tmp_path)
    assert len(got) == 1 and got[0] == ep
    from tensorboard.backend.event_processing.event_file_loader import EventFileLoader
    tags = [(v.tag, ev.step) for f in tmp_path.glob("events.out.tfevents.*") for ev in EventFileLoader(str(f)).Load() for v in ev.summary.value]
    assert tags == [(golden["episode"]["tag"], golden["episode"]["step"])]


def test_writer_protocol_errors(tmp_path):
    w = FlightEpisodeWriterB200(tmp_path, use_tensorboard=False)
    with pytest.raises(RuntimeError, match="No episode in progress"):
        w.end_episode()
    w.start_episode("a")
    with pytest.raises(RuntimeError, match="already in progress"):
        w.start_episode("a")
    with pytest.raises(ValueError, match="Agent ID mismatch"):
        w.log_flight_data(step=0, agent_id="b", position=(0, 0, 0), orientation=(0, 0, 0), velocity=(0, 0, 0), telemetry={},
                          rl_metrics={"reward": 0.0, "action": [0, 0, 0, 0]})
    w.close()
    assert json.loads(open(tmp_path / "flight_episodes.jsonl").read())["termination_reason"] == "incomplete"
