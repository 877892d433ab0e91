"""CPU oracle of the stage 1+2 extraction path -- TEST INFRASTRUCTURE ONLY.

Imported by tests/, by __graft_entry__.smoke() and by bench.py's CPU legs (cpu_baseline / parity / --impl reference);
never by the product path (fangyan_tts_b200/), which has no CPU fallback.  What each module follows and what it is
pinned against (details in the module headers and in DESIGN.md section 2):

    frontend.py     torchaudio kaldi.fbank / Resample (the reference's literal callees: pinned bit-exact / <= 1e-6),
                    whisper.log_mel_spectrogram (pinned on transformers' WhisperFeatureExtractor; unpinned against
                    whisper itself), matcha mel_spectrogram (golden vectors from the reference's own function)
    campplus.py     published CAM++ architecture -- PARITY UNPINNED (no campplus.onnx, no onnxruntime offline)
    s3tokenizer.py  published S3Tokenizer v2/v3 architecture -- PARITY UNPINNED as a network; its RoPE and FSMN-memory
                    blocks are pinned bit-equal on third-party code in the image (tests/test_oracle_pins_cpu.py)
    stage3.py       the pandas `job` of tools/make_parquet_list.py:26-62, restated for the stage-3 packing tests
"""
