"""Audio packaging shared by the AudioLDM and Bark callbacks (reference: swarm/audio/audioldm.py:27-33,
swarm/audio/bark.py:23-35 write a float32 WAV with scipy then transcode to mp3 with pydub/ffmpeg)."""
from __future__ import annotations

import io
import shutil
import subprocess
import tempfile
from io import BytesIO

import numpy as np


def wav_bytes(audio: np.ndarray, rate: int) -> bytes:
    import scipy.io.wavfile

    buf = io.BytesIO()
    scipy.io.wavfile.write(buf, rate=rate, data=audio.astype(np.float32))
    return buf.getvalue()


def encode_audio(audio: np.ndarray, rate: int, content_type: str = "audio/mpeg"):
    """Returns (BytesIO, content_type). mp3 needs ffmpeg (as in the reference); when ffmpeg is not
    installed the WAV itself is returned with content type audio/wav rather than failing."""
    wav = wav_bytes(audio, rate)
    if content_type == "audio/mpeg" and shutil.which("ffmpeg"):
        with tempfile.TemporaryDirectory() as tmp:
            src, dst = f"{tmp}/a.wav", f"{tmp}/a.mp3"
            with open(src, "wb") as fh:
                fh.write(wav)
            r = subprocess.run(["ffmpeg", "-y", "-loglevel", "error", "-i", src, dst])
            if r.returncode == 0:
                with open(dst, "rb") as fh:
                    return BytesIO(fh.read()), "audio/mpeg"
    return BytesIO(wav), ("audio/wav" if content_type == "audio/mpeg" else content_type)
