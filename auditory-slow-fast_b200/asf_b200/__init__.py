"""asf_b200 -- B200-native (sm_100a) implementation of the Auditory SlowFast hot path behind the
reference's Python API (log-mel extraction, pathway packing, build_model, AudioSlowFast/ResNet forward +
backward, data-parallel training step).  See DESIGN.md / INTEGRATION.md at the repository root."""
__version__ = "0.1.0"

import os as _os

# The engine runs one step on ~12 CUDA streams (two pathways, weight-gradient streams, copies, collectives).  With the
# default of 8 hardware work queues several streams share a queue and a kernel of one stream can sit behind an
# unrelated kernel of another ("false dependencies"): that costs overlap everywhere, and for the SyncBN peer exchange
# (csrc/peer.cu), whose kernels wait for the same kernel on the other GPUs, it can turn into a cross-rank wait cycle.
# 32 queues give every stream its own.  Read by the driver when the CUDA context is created: set before any CUDA call.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
