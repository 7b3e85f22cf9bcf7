"""speech_feature_extraction.speechpy.processing drop-in (load_data.py:5, utils.py:6)."""
from speaker_verification_b200.speechpy.processing import *  # noqa: F401,F403
from speaker_verification_b200.speechpy.processing import __all__  # noqa: F401
