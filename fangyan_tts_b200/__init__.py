"""fangyan_tts_b200: B200-native stage 1+2 extraction (speaker embedding + speech tokens) for CosyVoice3 data prep."""
__version__ = "0.1"
