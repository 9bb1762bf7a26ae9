"""B200-native batched physics + operational-space-control backend for the
RearrangementEnv hot path of peterdavidfagan/mujoco_robot_environments."""
__version__ = "0.1.0"
