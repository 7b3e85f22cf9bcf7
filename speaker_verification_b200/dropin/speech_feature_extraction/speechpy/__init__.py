"""``import speech_feature_extraction.speechpy as speech`` drop-in.

The reference's directory has no __init__.py (namespace package) and utils.py:388-395
reaches ``speech.feature`` / ``speech.processing`` only because load_data.py:5 imported the
submodules first; importing them here makes both spellings work in any order.
"""
from . import feature, functions, processing  # noqa: F401
