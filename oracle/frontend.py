"""CPU oracle for the audio front-ends of stages 1-2 (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product path never does.

Parity status: the reference holds no tests or golden vectors for this path
(SURVEY.md section 4), so the restatements here are pinned against the third-party
callees the reference uses, where they import in the build container:
  * kaldi_fbank      <-> torchaudio.compliance.kaldi.fbank   (torchaudio 2.11.0; ref pins 2.3.1)
  * resample         <-> torchaudio.transforms.Resample       (same)
  * whisper_log_mel  <-> transformers.WhisperFeatureExtractor (openai-whisper is absent:
                         "parity unpinned" against whisper itself)
tests/test_oracle_cpu.py performs those checks when the libraries import.

Reference call sites restated here (paths relative to /root/reference/CosyVoice):
  tools/extract_embedding.py:26-33       load -> resample -> kaldi.fbank(80, dither=0) -> CMN
  tools/extract_speech_token.py:28-38    load -> resample -> mono -> 30 s gate -> whisper log-mel(128)
  cosyvoice/utils/file_utils.py:44-50    load_wav
"""
from __future__ import annotations

import math
import wave

import numpy as np
import torch

SAMPLE_RATE = 16000

# ----------------------------------------------------------------------------- A1 decode


def load_wav_pcm(path):
    """RIFF PCM16 WAV -> (float32 [C, N] in [-1, 1), sample_rate).

    Mirrors torchaudio.load(normalize=True) for 16-bit PCM: int16 / 32768
    (tools/extract_embedding.py:26).  Only PCM16 is supported by the oracle.
    """
    with wave.open(str(path), "rb") as w:
        ch, sw, sr, n = w.getnchannels(), w.getsampwidth(), w.getframerate(), w.getnframes()
        if sw != 2:
            raise ValueError(f"{path}: only PCM16 supported by the oracle (sampwidth={sw})")
        raw = w.readframes(n)
    pcm = np.frombuffer(raw, dtype="<i2").reshape(-1, ch).T
    return torch.from_numpy(pcm.astype(np.float32) / 32768.0), sr


# ----------------------------------------------------------------------------- A2 resample


def resample_kernel(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """Polyphase windowed-sinc filter bank of torchaudio's default Resample
    (sinc_interp_hann), built in float64 and cast to float32.

    Follows torchaudio/functional/functional.py:1305-1402 (_get_sinc_resample_kernel).
    Returns (kernel float32 [new, 2*width+orig], width, orig, new) with orig/new reduced by gcd.
    """
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    base = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base)
    idx = np.arange(-width, width + orig, dtype=np.float64)[None, :] / orig
    # torchaudio builds the phase offsets from an int64 arange, so `/ new` is a float32 division (its `dtype`
    # argument is None at this call); only `idx` is float64.  Reproduced here or 22.05/44.1 kHz differ by ~7e-6.
    phase = (np.arange(0, -new, -1).astype(np.float32) / np.float32(new)).astype(np.float64)
    t = phase[:, None] + idx
    t = t * base
    t = np.clip(t, -lowpass_filter_width, lowpass_filter_width)
    window = np.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base / orig
    safe = np.where(t == 0, 1.0, t)
    kern = np.where(t == 0, 1.0, np.sin(safe) / safe) * window * scale
    return kern.astype(np.float32), width, orig, new


def resample(wave_: torch.Tensor, orig_freq: int, new_freq: int = SAMPLE_RATE) -> torch.Tensor:
    """[C, N] float32 -> [C, ceil(new*N/orig)] float32 (functional.py:1405-1432)."""
    if orig_freq == new_freq:
        return wave_
    kern, width, orig, new = resample_kernel(orig_freq, new_freq)
    c, n = wave_.shape
    x = torch.nn.functional.pad(wave_, (width, width + orig))
    y = torch.nn.functional.conv1d(x[:, None, :], torch.from_numpy(kern)[:, None, :], stride=orig)
    y = y.transpose(1, 2).reshape(c, -1)
    target = int(math.ceil(new * n / orig))
    return y[:, :target]


# ----------------------------------------------------------------------------- A3/A4 kaldi fbank + CMN

FBANK_FRAME = 400
FBANK_SHIFT = 160
FBANK_NFFT = 512
FBANK_BINS = 80
FBANK_PREEMPH = 0.97
FBANK_EPS = float(np.finfo(np.float32).eps)  # 1.1920929e-07


def kaldi_num_frames(n: int) -> int:
    """snip_edges=True frame count (kaldi.py:44-83)."""
    return 0 if n < FBANK_FRAME else 1 + (n - FBANK_FRAME) // FBANK_SHIFT


def povey_window() -> np.ndarray:
    """hann(400, periodic=False) ** 0.85 (kaldi.py:98-100), float32."""
    w = torch.hann_window(FBANK_FRAME, periodic=False, dtype=torch.float32).pow(0.85)
    return w.numpy()


def kaldi_mel_banks(num_bins=FBANK_BINS, nfft=FBANK_NFFT, sr=SAMPLE_RATE, low=20.0, high=0.0) -> np.ndarray:
    """Triangular mel bank [num_bins, nfft/2] (Nyquist column dropped; kaldi.py:436-511),
    computed in float32 like the callee so the weights are bit-identical."""
    nyq = 0.5 * sr
    if high <= 0.0:
        high += nyq
    nbin_fft = nfft // 2
    fft_bin_width = sr / nfft

    def mel(f):
        return 1127.0 * math.log(1.0 + f / 700.0)

    mel_low, mel_high = mel(low), mel(high)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = torch.arange(num_bins, dtype=torch.float32).unsqueeze(1)
    left = mel_low + b * delta
    center = mel_low + (b + 1.0) * delta
    right = mel_low + (b + 2.0) * delta
    f = fft_bin_width * torch.arange(nbin_fft, dtype=torch.float32)
    m = (1127.0 * (1.0 + f / 700.0).log()).unsqueeze(0)
    up = (m - left) / (center - left)
    down = (right - m) / (right - center)
    bank = torch.max(torch.zeros(1), torch.min(up, down))
    return bank.numpy()


def kaldi_fbank(wave_: torch.Tensor) -> torch.Tensor:
    """[C, N] float32 -> [M, 80] float32 using channel 0, exactly the arguments of
    tools/extract_embedding.py:29-32 (num_mel_bins=80, dither=0, sample_frequency=16000;
    everything else at the kaldi.fbank defaults, kaldi.py:514-645)."""
    x = wave_[0]
    m = kaldi_num_frames(x.numel())
    if m == 0:
        return torch.empty(0, FBANK_BINS)
    frames = x.unfold(0, FBANK_FRAME, FBANK_SHIFT)[:m].clone()          # [M, 400]
    frames = frames - frames.mean(dim=1, keepdim=True)                  # remove_dc_offset
    prev = torch.cat([frames[:, :1], frames[:, :-1]], dim=1)            # replicate-left
    frames = frames - FBANK_PREEMPH * prev
    frames = frames * torch.from_numpy(povey_window())
    frames = torch.nn.functional.pad(frames, (0, FBANK_NFFT - FBANK_FRAME))
    spec = torch.fft.rfft(frames).abs().pow(2.0)                        # [M, 257] power
    bank = torch.from_numpy(kaldi_mel_banks())
    bank = torch.nn.functional.pad(bank, (0, 1))                        # zero Nyquist column
    mel = spec @ bank.T
    return torch.max(mel, torch.tensor(FBANK_EPS)).log()


def cmn(feat: torch.Tensor) -> torch.Tensor:
    """tools/extract_embedding.py:33."""
    return feat - feat.mean(dim=0, keepdim=True)


# ----------------------------------------------------------------------------- A5 whisper log-mel

WHISPER_NFFT = 400
WHISPER_HOP = 160
WHISPER_BINS = 128


def slaney_mel_filters(n_mels=WHISPER_BINS, n_fft=WHISPER_NFFT, sr=SAMPLE_RATE) -> np.ndarray:
    """librosa.filters.mel(sr, n_fft, n_mels) with its defaults (Slaney scale, Slaney area
    norm, fmin 0, fmax sr/2): the table openai-whisper ships as assets/mel_filters.npz
    (whisper/audio.py::mel_filters).  Returns float32 [n_mels, n_fft/2+1]."""
    def hz_to_mel(f):
        f = np.asarray(f, dtype=np.float64)
        f_sp = 200.0 / 3
        mels = f / f_sp
        min_log_hz = 1000.0
        min_log_mel = min_log_hz / f_sp
        logstep = math.log(6.4) / 27.0
        return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, mels)

    def mel_to_hz(m):
        m = np.asarray(m, dtype=np.float64)
        f_sp = 200.0 / 3
        min_log_hz = 1000.0
        min_log_mel = min_log_hz / f_sp
        logstep = math.log(6.4) / 27.0
        return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)

    fftfreqs = np.linspace(0, sr / 2, n_fft // 2 + 1)
    mel_pts = np.linspace(hz_to_mel(0.0), hz_to_mel(sr / 2), n_mels + 2)
    hz_pts = mel_to_hz(mel_pts)
    fdiff = np.diff(hz_pts)
    ramps = hz_pts[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, n_fft // 2 + 1))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (hz_pts[2:n_mels + 2] - hz_pts[:n_mels])
    w *= enorm[:, None]
    return w.astype(np.float32)


def whisper_num_frames(n: int) -> int:
    return n // WHISPER_HOP


def whisper_log_mel(audio: torch.Tensor) -> torch.Tensor:
    """[1, N] float32 -> [1, 128, N//160] float32.

    Restates whisper.log_mel_spectrogram(audio, n_mels=128) of openai-whisper 20231117
    (whisper/audio.py), the callee of tools/extract_speech_token.py:38: centred reflect-padded
    STFT with a periodic Hann window, last frame dropped, power, Slaney mel, log10 floor 1e-10,
    clamp to (global max - 8), (x + 4) / 4.
    """
    window = torch.hann_window(WHISPER_NFFT)
    stft = torch.stft(audio, WHISPER_NFFT, WHISPER_HOP, window=window, return_complex=True)
    mag = stft[..., :-1].abs() ** 2
    filt = torch.from_numpy(slaney_mel_filters())
    mel = filt @ mag
    log_spec = torch.clamp(mel, min=1e-10).log10()
    log_spec = torch.maximum(log_spec, log_spec.max() - 8.0)
    return (log_spec + 4.0) / 4.0


# ----------------------------------------------------------------------------- N4 matcha 24 kHz mel

MATCHA_SR, MATCHA_NFFT, MATCHA_HOP, MATCHA_MELS = 24000, 1920, 480, 80


def matcha_mel(audio: torch.Tensor) -> torch.Tensor:
    """[1, N] float32 at 24 kHz -> [1, 80, T] float32, T = 1 + (N - 480) // 480.

    Restates matcha.utils.audio.mel_spectrogram(y, n_fft=1920, num_mels=80, sampling_rate=24000, hop_size=480,
    win_size=1920, fmin=0, fmax=None, center=False) (third_party/Matcha-TTS/matcha/utils/audio.py:44-88), the
    `feat_extractor` of conf/cosyvoice3.yaml:140-151 called at cosyvoice/cli/frontend.py:122 and
    cosyvoice/dataset/processor.py:174: reflect-pad (n_fft - hop) / 2 on both sides, periodic-Hann STFT without
    centring, sqrt(re^2 + im^2 + 1e-9), librosa Slaney mel basis, log(clamp(., 1e-5)).
    Pinned by tests/test_oracle_cpu.py against that function itself (imported from /root/reference with
    librosa.filters.mel supplied by slaney_mel_filters, itself pinned on transformers' mel_filter_bank) and by the
    golden vectors tests/golden/matcha.npz generated the same way."""
    pad = (MATCHA_NFFT - MATCHA_HOP) // 2
    y = torch.nn.functional.pad(audio.unsqueeze(1), (pad, pad), mode="reflect").squeeze(1)
    spec = torch.stft(y, MATCHA_NFFT, hop_length=MATCHA_HOP, win_length=MATCHA_NFFT,
                      window=torch.hann_window(MATCHA_NFFT), center=False, normalized=False, onesided=True,
                      return_complex=True)
    mag = torch.sqrt(spec.real ** 2 + spec.imag ** 2 + 1e-9)
    basis = torch.from_numpy(slaney_mel_filters(MATCHA_MELS, MATCHA_NFFT, MATCHA_SR))
    return torch.log(torch.clamp(basis @ mag, min=1e-5))


# ----------------------------------------------------------------------------- A8 gate


def stage2_gate(num_samples_16k: int) -> bool:
    """True when the utterance is tokenised; False -> [] (extract_speech_token.py:34-36).
    Strictly greater than 30 s is dropped."""
    return not (num_samples_16k / 16000 > 30)
