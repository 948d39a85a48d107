"""pytest plugin for running the REFERENCE's own test files against this repository's drop-ins (TEST INFRASTRUCTURE).

    python -m pytest -p acb200_dropin_plugin /root/reference/tests/test_pid_bindings.py ...

ACB200_DROPIN selects what is swapped in before the reference's test modules are imported:
  * ``pid``    -- ``sys.modules["aircraft_controls_bindings"]`` = ``acb200.pid_bindings`` (the ctypes mirror of the pybind
                  module over ``acb200_pid_compute_f32``): the reference's Python agents then run on this library's PID core;
  * ``agents`` -- additionally the agent / mission classes of ``controllers.*`` are replaced by ``acb200.adapters`` /
                  ``acb200.guidance`` ones (reference physics backend, reference dataclasses, our controllers).
Every swapped class counts its constructions; the totals are written to $ACB200_DROPIN_REPORT so that the caller can
assert the drop-ins were what actually ran.  Nothing of the reference is copied or modified: it is imported from where it
lies and its names are rebound in memory.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import acb200  # noqa: E402

MODE = os.environ.get("ACB200_DROPIN", "pid")
COUNTS = {}


def _counted(cls, name):
    class Counted(cls):
        def __init__(self, *a, **kw):
            COUNTS[name] = COUNTS.get(name, 0) + 1
            super().__init__(*a, **kw)
    Counted.__name__ = cls.__name__
    Counted.__qualname__ = cls.__qualname__
    return Counted


sys.modules["aircraft_controls_bindings"] = acb200.pid_bindings
_orig_compute = acb200.pid_bindings.PIDController.compute


def _compute(self, *a, **kw):
    COUNTS["pid_compute"] = COUNTS.get("pid_compute", 0) + 1
    return _orig_compute(self, *a, **kw)


acb200.pid_bindings.PIDController.compute = _compute

if MODE == "agents":
    import controllers.attitude_agent as _att
    import controllers.hsa_agent as _hsa
    import controllers.mission_planner as _mp
    import controllers.rate_agent as _rate
    import controllers.surface_agent as _surf
    import controllers.waypoint_agent as _wp
    import controllers as _pkg
    from acb200 import adapters, guidance
    swaps = [(_rate, "RateAgent", adapters.RateAgent), (_att, "AttitudeAgent", adapters.AttitudeAgent),
             (_surf, "SurfaceAgent", adapters.SurfaceAgent), (_hsa, "HSAAgent", guidance.HSAAgent),
             (_wp, "WaypointAgent", guidance.WaypointAgent), (_mp, "MissionPlanner", guidance.MissionPlanner)]
    for mod, name, ours in swaps:
        c = _counted(ours, name)
        setattr(mod, name, c)
        if hasattr(_pkg, name):
            setattr(_pkg, name, c)


def pytest_sessionfinish(session, exitstatus):
    path = os.environ.get("ACB200_DROPIN_REPORT")
    if path:
        worker = os.environ.get("PYTEST_XDIST_WORKER", "main")
        with open(f"{path}.{worker}", "w") as f:
            json.dump(COUNTS, f)
