from .vec_env import MicroRTSGridModeVecEnv, unit_type_table  # noqa: F401
