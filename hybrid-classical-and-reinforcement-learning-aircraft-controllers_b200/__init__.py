"""acb200: B200-native batched flight-control path (6-DOF step, cascaded PID, rate-env) behind the
reference's interfaces.  Import as ``acb200`` (see acb200.py at the repository root)."""
from . import _cabi
from ._cabi import Context, default_aircraft_params, default_controller_params, kernel_info
from .envs import BatchedRateControlEnv
from .fleet import AircraftFleet
from . import adapters, pid_bindings
from .adapters import SimulationAircraftBackendB200
from . import sharding, vecenv
from .vecenv import B200VecEnv
from . import policy
from .policy import LearnedPolicyB200
from . import rollout
from .rollout import RolloutBufferB200
from . import sensor, telemetry
from . import config_loader, guidance
from . import gym_env
from . import flight_episode
from .flight_episode import FlightEpisodeWriterB200
from .gym_env import RateControlEnvB200, ResidualRateControlEnvB200
from .guidance import HSAAgent, MissionPlanner, WaypointAgent
from .telemetry import TelemetryRecorderB200
from .sensor import NoisySensorB200, NoisySensorInterfaceB200, PerfectSensorB200

__all__ = ["flight_episode", "FlightEpisodeWriterB200", "gym_env", "RateControlEnvB200", "ResidualRateControlEnvB200", "config_loader", "guidance", "HSAAgent", "WaypointAgent", "MissionPlanner", "telemetry", "TelemetryRecorderB200", "sensor", "NoisySensorB200", "NoisySensorInterfaceB200", "PerfectSensorB200", "policy", "LearnedPolicyB200", "rollout", "RolloutBufferB200", "_cabi", "adapters", "pid_bindings", "SimulationAircraftBackendB200", "vecenv", "sharding", "B200VecEnv", "Context", "AircraftFleet", "BatchedRateControlEnv", "default_aircraft_params",
           "default_controller_params", "kernel_info"]
__version__ = "0.1.0"
