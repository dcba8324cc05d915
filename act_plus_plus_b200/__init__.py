"""act_plus_plus_b200 -- B200-native (sm_100a) implementation of the ACT DETRVAE policy hot path.

The compute lives in ``libact_b200.so`` (hand-written CUDA: tcgen05/TMA contractions + fused bandwidth
kernels) behind the C ABI of ``include/act_b200.h``; this package is the Python host side mirroring the
reference's ``policy.ACTPolicy`` / ``detr.main.build_ACT_model_and_optimizer`` interface.
There is no CPU fallback: importing the compute ops without the built library raises.
"""
__version__ = "0.1.0"
