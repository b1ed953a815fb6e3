from .audio_model_builder import AudioSlowFast, ResNet  # noqa: F401
from .build import MODEL_REGISTRY, build_model  # noqa: F401
