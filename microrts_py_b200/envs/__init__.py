from .vec_env import MicroRTSBotVecEnv, MicroRTSGridModeVecEnv, unit_type_table  # noqa: F401
