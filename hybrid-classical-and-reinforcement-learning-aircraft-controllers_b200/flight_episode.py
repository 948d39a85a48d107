"""Flight-plugin episode export (second half): the reference's TensorBoard "flight" plugin format.

What the reference writes: ``FlightLoggerCallback._extract_flight_data`` (tensorboard_flight_plugin/src/tensorboard_flight/
callbacks.py:152-263) turns one env step -- the 18-float observation, the action, the reward and the env's ``info`` dict --
into a step record; ``FlightLogger`` (logger.py:93-293) collects the records of an episode into the ``FlightEpisode`` layout
of data/schema.py:170-215 and stores ``json.dumps(episode.to_dict())`` as the plugin content of ONE summary value with tag
``flight/<agent_id>/episode`` and ``plugin_name = "flight"`` in a TensorBoard event file.  The plugin's UI reads exactly that.

Here: the same step mapping and the same episode dictionary, built from what ``RateControlEnvB200`` / ``B200VecEnv`` return
(their ``info`` carries the keys the mapping reads: ``position``, ``time``, ``reward_components``).  ``FlightEpisodeWriterB200``
always appends the episode JSON to ``<log_dir>/flight_episodes.jsonl`` (plain file, no dependency) and, when ``tensorboard``
is importable, also writes the event-file record the plugin loads.  The UI itself is out of scope (graphics: unavailable
on this box); this module only produces its input.  Pinned on an episode logged by the unmodified reference classes
(tests/golden/flight_episode.json, tests/test_flight_episode.py).
"""
from __future__ import annotations

import json
import time
from pathlib import Path
from typing import Any, Dict, List, Optional

import numpy as np


def _py(v: Any) -> Any:
    """numpy -> native types, as data/schema.py:8-18 does before serialising"""
    if isinstance(v, (np.integer, np.floating)):
        return v.item()
    if isinstance(v, np.ndarray):
        return v.tolist()
    if isinstance(v, dict):
        return {k: _py(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return [_py(x) for x in v]
    return v


def flight_step_record(obs, action, reward, info: Dict[str, Any], step: int, agent_id: str = "sb3_agent") -> Optional[dict]:
    """callbacks.py:152-263: observation layout [p,q,r, cmd(3), err(3), airspeed, altitude, roll, pitch, yaw, prev_action(4)];
    velocity approximated from airspeed and attitude in NED; angles in degrees; control surfaces from the action."""
    if len(obs) < 18:
        return None
    p, q, r = obs[0], obs[1], obs[2]
    airspeed, altitude = float(obs[9]), float(obs[10])
    roll, pitch, yaw = float(obs[11]), float(obs[12]), float(obs[13])
    position = info.get("position", np.array([0.0, 0.0, -altitude]))
    if not isinstance(position, (list, np.ndarray)):
        position = np.array([0.0, 0.0, -altitude])
    position = tuple(float(x) for x in position[:3])
    vx = airspeed * np.cos(pitch) * np.cos(yaw)
    vy = airspeed * np.cos(pitch) * np.sin(yaw)
    vz = -airspeed * np.sin(pitch)
    telemetry = {"airspeed": airspeed, "altitude": altitude, "g_force": 1.0,
                 "throttle": float(action[3]) if len(action) > 3 else 0.5, "aoa": 0.0, "aos": 0.0,
                 "heading": float(np.degrees(yaw)), "vertical_speed": float(vz), "turn_rate": float(r),
                 "bank_angle": float(np.degrees(roll)),
                 "aileron": float(action[0]) if len(action) > 0 else 0.0, "elevator": float(action[1]) if len(action) > 1 else 0.0,
                 "rudder": float(action[2]) if len(action) > 2 else 0.0}
    rl = {"reward": float(reward), "action": action.tolist() if isinstance(action, np.ndarray) else list(action),
          "value_estimate": info.get("value", None), "policy_logprob": info.get("log_prob", None)}
    if "reward_components" in info:
        rl["reward_components"] = info["reward_components"]
    events = []
    if "termination_reason" in info and info.get("dones", False):
        events.append({"timestamp": info.get("time", 0.0), "type": "termination",
                       "severity": "error" if "crash" in info["termination_reason"] else "info", "message": info["termination_reason"]})
    return {"step": step, "agent_id": agent_id, "position": position,
            "orientation": (float(np.degrees(roll)), float(np.degrees(pitch)), float(np.degrees(yaw))),
            "velocity": (float(vx), float(vy), float(vz)), "angular_velocity": (float(p), float(q), float(r)),
            "telemetry": telemetry, "rl_metrics": rl, "events": events if events else None,
            "timestamp": info.get("time", step * 0.02)}


def _data_point(rec: dict, cumulative_reward: float) -> dict:
    """logger.py:93-196 + FlightDataPoint.to_dict (schema.py:131-167): optional fields appear only when set."""
    t, rl = rec["telemetry"], rec["rl_metrics"]
    telemetry = {k: _py(t.get(k, d)) for k, d in (("airspeed", 0.0), ("altitude", 0.0), ("g_force", 1.0), ("throttle", 0.0),
                                                 ("aoa", 0.0), ("aos", 0.0), ("heading", 0.0), ("vertical_speed", 0.0),
                                                 ("turn_rate", 0.0))}
    telemetry["bank_angle"] = _py(t.get("bank_angle", rec["orientation"][0]))
    for k in ("aileron", "elevator", "rudder"):
        if t.get(k) is not None:
            telemetry[k] = _py(t[k])
    action = rl["action"] if isinstance(rl["action"], list) else list(rl["action"])
    metrics = {"reward": _py(rl["reward"]), "cumulative_reward": _py(cumulative_reward), "action": _py(action)}
    for k in ("policy_logprob", "value_estimate", "advantage", "entropy", "reward_components"):
        if rl.get(k) is not None:
            metrics[k] = _py(rl[k])
    ts = rec.get("timestamp")
    ts = float(rec["step"]) if ts is None else ts
    point = {"timestamp": _py(ts), "step": int(rec["step"]), "position": _py(np.array(rec["position"])),
             "orientation": {"roll": _py(rec["orientation"][0]), "pitch": _py(rec["orientation"][1]), "yaw": _py(rec["orientation"][2])},
             "velocity": _py(np.array(rec["velocity"])),
             "angular_velocity": _py(np.array(rec["angular_velocity"] if rec.get("angular_velocity") else [0.0, 0.0, 0.0])),
             "telemetry": telemetry, "rl_metrics": metrics}
    if rec.get("events"):
        point["events"] = [dict({"timestamp": _py(e.get("timestamp", ts)), "event_type": e.get("type", "custom"),
                                 "severity": e.get("severity", "info"), "message": e.get("message", "")},
                                **({"metadata": _py(e["metadata"])} if e.get("metadata") is not None else {})) for e in rec["events"]]
    return point


class FlightEpisodeWriterB200:
    """``FlightLogger``'s episode protocol (logger.py:79-262): ``start_episode``, ``log_flight_data(**record)`` /
    ``log_step(obs, action, reward, info, step)``, ``end_episode(success, termination_reason, config, tags)``."""

    def __init__(self, log_dir, agent_id: str = "sb3_agent", use_tensorboard: Optional[bool] = None):
        self.log_dir = Path(log_dir)
        self.log_dir.mkdir(parents=True, exist_ok=True)
        self.agent_id = agent_id
        self.episode_number = 0
        self.current_episode: Optional[List[dict]] = None
        self.cumulative_reward = 0.0
        self.episode_start_time: Optional[float] = None
        self.jsonl_path = self.log_dir / "flight_episodes.jsonl"
        self._writer = None
        if use_tensorboard is not False:
            try:
                from tensorboard.summary.writer.event_file_writer import EventFileWriter
                self._writer = EventFileWriter(str(self.log_dir))
            except ImportError:
                if use_tensorboard:
                    raise

    def start_episode(self, agent_id: Optional[str] = None) -> None:
        if self.current_episode is not None:
            raise RuntimeError("Episode already in progress. Call end_episode() first.")
        self.current_episode, self.cumulative_reward, self.episode_start_time = [], 0.0, time.time()
        if agent_id is not None:
            self.agent_id = agent_id

    def log_flight_data(self, **rec) -> None:
        if self.current_episode is None:
            self.start_episode(rec.get("agent_id"))
        if rec.get("agent_id", self.agent_id) != self.agent_id:
            raise ValueError(f"Agent ID mismatch: {rec.get('agent_id')} != {self.agent_id}. Call end_episode() before switching agents.")
        self.cumulative_reward += rec["rl_metrics"].get("reward", 0.0)
        self.current_episode.append(_data_point(rec, self.cumulative_reward))

    def log_step(self, obs, action, reward, info, step: int) -> None:
        rec = flight_step_record(np.asarray(obs), np.asarray(action), reward, info, step, self.agent_id)
        if rec is not None:
            self.log_flight_data(**rec)

    def end_episode(self, success: bool = False, termination_reason: str = "unknown", config: Optional[dict] = None,
                    tags: Optional[List[str]] = None) -> dict:
        if self.current_episode is None:
            raise RuntimeError("No episode in progress. Call start_episode() first.")
        ep = {"episode_id": f"{self.agent_id}_ep{self.episode_number}", "agent_id": self.agent_id,
              "episode_number": int(self.episode_number), "start_time": _py(self.episode_start_time),
              "duration": _py(time.time() - self.episode_start_time), "total_steps": int(len(self.current_episode)),
              "total_reward": _py(self.cumulative_reward), "success": bool(success), "termination_reason": termination_reason,
              "trajectory": self.current_episode}
        if config is not None:
            ep["config"] = _py(config)
        if tags is not None:
            ep["tags"] = tags
        self._write(ep)
        self.current_episode, self.episode_start_time, self.cumulative_reward = None, None, 0.0
        self.episode_number += 1
        return ep

    def _write(self, ep: dict) -> None:
        content = json.dumps(ep)
        with open(self.jsonl_path, "a") as f:
            f.write(content + "\n")
        if self._writer is not None:          # logger.py:265-293: one summary value, the JSON as the plugin's content
            from tensorboard.compat.proto.event_pb2 import Event
            from tensorboard.compat.proto.summary_pb2 import Summary
            summary = Summary()
            value = summary.value.add()
            value.tag = f"flight/{ep['agent_id']}/episode"
            value.metadata.plugin_data.plugin_name = "flight"
            value.metadata.plugin_data.content = content.encode("utf-8")
            self._writer.add_event(Event(wall_time=time.time(), step=ep["episode_number"], summary=summary))
            self._writer.flush()

    def flush(self) -> None:
        if self._writer is not None:
            self._writer.flush()

    def close(self) -> None:
        if self.current_episode is not None:
            self.end_episode(termination_reason="incomplete")
        if self._writer is not None:
            self._writer.flush(); self._writer.close()
            self._writer = None


def read_flight_events(log_dir) -> List[dict]:
    """The episodes a TensorBoard event directory holds for the flight plugin (what plugin.py decodes): tag
    ``flight/*/episode``, plugin_name ``flight``, JSON content."""
    from tensorboard.backend.event_processing.event_file_loader import EventFileLoader
    out = []
    for path in sorted(Path(log_dir).glob("events.out.tfevents.*")):
        for ev in EventFileLoader(str(path)).Load():
            for v in ev.summary.value:
                if v.metadata.plugin_data.plugin_name == "flight":
                    out.append(json.loads(v.metadata.plugin_data.content.decode("utf-8")))
    return out
