"""mjrl/samplers/evaluation_sampler.py: the module name the reference imports from."""
from .base_sampler import do_evaluation_rollout, do_evaluation_rollout_star  # noqa: F401
