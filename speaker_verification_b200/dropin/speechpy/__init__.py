"""``import speechpy`` drop-in (speech_feature_extraction/example/test_package.py:8,18)."""
from speaker_verification_b200.speechpy import feature, functions, processing  # noqa: F401
