"""Put this directory first on sys.path and the reference's `from policy import ACTPolicy` (imitate_episodes.py:20)
resolves to the B200 engine."""
from act_plus_plus_b200.policy import ACTPolicy, kl_divergence  # noqa: F401


class CNNMLPPolicy:
    def __init__(self, *a, **k):
        raise NotImplementedError("CNNMLPPolicy is outside the B200 hot path (SURVEY.md section 2, item 13)")


class DiffusionPolicy:
    def __init__(self, *a, **k):
        raise NotImplementedError("DiffusionPolicy is outside the B200 hot path (SURVEY.md section 2, item 14)")
