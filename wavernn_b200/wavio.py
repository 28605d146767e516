"""Dependency-free float32 WAV writer.

Replaces `save_wav` (`utils/dsp.py:22-23`), which calls `librosa.output.write_wav` — an API that
no longer exists in current librosa and is absent on this box.  Same output contract: float32
samples, `sample_rate` Hz, mono (IEEE-float WAVE, format tag 3).
"""
import struct

import numpy as np


def save_wav(x, path, sample_rate):
    data = np.asarray(x, dtype=np.float32).tobytes()
    hdr = b'RIFF' + struct.pack('<I', 50 + len(data)) + b'WAVE'
    hdr += b'fmt ' + struct.pack('<IHHIIHHH', 18, 3, 1, int(sample_rate), int(sample_rate) * 4, 4, 32, 0)
    hdr += b'fact' + struct.pack('<II', 4, len(data) // 4)
    hdr += b'data' + struct.pack('<I', len(data))
    with open(str(path), 'wb') as f:
        f.write(hdr + data)


class WavStreamWriter:
    """The same file as `save_wav`, written chunk by chunk as fold groups finish (`Vocoder.generate_stream`): the header
    goes out first with zero lengths, every chunk is appended and flushed, `close` patches the three length fields."""

    def __init__(self, path, sample_rate):
        self.f = open(str(path), 'wb')
        self.n = 0
        self.rate = int(sample_rate)
        self.f.write(self._header(0))

    def _header(self, nbytes):
        hdr = b'RIFF' + struct.pack('<I', 50 + nbytes) + b'WAVE'
        hdr += b'fmt ' + struct.pack('<IHHIIHHH', 18, 3, 1, self.rate, self.rate * 4, 4, 32, 0)
        hdr += b'fact' + struct.pack('<II', 4, nbytes // 4)
        hdr += b'data' + struct.pack('<I', nbytes)
        return hdr

    def write(self, chunk):
        data = np.asarray(chunk, dtype=np.float32).tobytes()
        self.f.write(data)
        self.f.flush()
        self.n += len(data)

    def close(self):
        if self.f:
            self.f.seek(0)
            self.f.write(self._header(self.n))
            self.f.close()
            self.f = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def load_wav(path):
    """Read back a file written by `save_wav` -> (float32 array, sample_rate)."""
    raw = open(str(path), 'rb').read()
    assert raw[:4] == b'RIFF' and raw[8:12] == b'WAVE'
    pos, rate, data = 12, None, None
    while pos + 8 <= len(raw):
        tag, size = raw[pos:pos + 4], struct.unpack('<I', raw[pos + 4:pos + 8])[0]
        body = raw[pos + 8:pos + 8 + size]
        if tag == b'fmt ':
            rate = struct.unpack('<I', body[4:8])[0]
        elif tag == b'data':
            data = np.frombuffer(body, dtype=np.float32).copy()
        pos += 8 + size + (size & 1)
    return data, rate
