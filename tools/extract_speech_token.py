#!/usr/bin/env python3
"""Drop-in for the reference's CosyVoice/tools/extract_speech_token.py (same flags, same files); see fangyan_tts_b200/cli.py."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from fangyan_tts_b200.cli import main_speech_token, run_tool  # noqa: E402

if __name__ == "__main__":
    run_tool(main_speech_token)
