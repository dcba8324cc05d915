"""Drop-in for the reference's `from detr.main import build_ACT_model_and_optimizer` (policy.py:6)."""
from act_plus_plus_b200.policy import build_ACT_model_and_optimizer  # noqa: F401
