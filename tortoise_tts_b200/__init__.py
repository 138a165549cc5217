"""tortoise_tts_b200 -- B200-native (sm_100a) UnifiedVoice autoregressive sampler behind the reference's
model-call signature.  Only the AR hot path lives here; see DESIGN.md."""
from .unified_voice import UnifiedVoiceB200, ARGeometry, patch_reference_module  # noqa: F401
from . import _lib  # noqa: F401

__all__ = ["UnifiedVoiceB200", "ARGeometry", "patch_reference_module"]
