#!/usr/bin/env python3
"""Drop-in for the reference's tools/make_parquet_list.py (same flags, same files): Arrow-direct stage 3 packing."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fangyan_tts_b200.parquet_pack import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
