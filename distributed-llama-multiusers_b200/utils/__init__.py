"""Small shared helpers: the environment-variable table (envvars)."""
from .envvars import ENV_VARS, table as env_table  # noqa: F401
