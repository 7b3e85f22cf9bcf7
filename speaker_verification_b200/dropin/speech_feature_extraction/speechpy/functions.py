"""speech_feature_extraction.speechpy.functions drop-in (load_data.py:5, utils.py:6)."""
from speaker_verification_b200.speechpy.functions import *  # noqa: F401,F403
from speaker_verification_b200.speechpy.functions import __all__  # noqa: F401
