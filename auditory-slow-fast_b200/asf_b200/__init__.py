"""asf_b200 -- B200-native (sm_100a) implementation of the Auditory SlowFast hot path behind the
reference's Python API (log-mel extraction, pathway packing, build_model, AudioSlowFast/ResNet forward +
backward, data-parallel training step).  See DESIGN.md / INTEGRATION.md at the repository root."""
__version__ = "0.1.0"
