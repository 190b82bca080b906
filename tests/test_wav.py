"""Audio decode + staging: wav header walk, sample decode, fused 16-bit ingest.

CPU: the oracle (oracle/np_wav.py) against two independent wav readers (stdlib ``wave``,
``scipy.io.wavfile``); the C ABI's ``afe_wav_parse`` and the host-emulated ``afe_wav_decode`` /
``afe_resample_batch_wav16`` against the oracle.  GPU (``-m gpu``): the same through the product
library and the reference-facing ``load_audio`` / ``validate_file`` / ``process_calls``."""
import io
import struct
import wave

import numpy as np
import pytest

from oracle import np_resample, np_wav
from tests import emu_util


# ------------------------------------------------------------------ builders
def _wav_bytes(frames: bytes, channels: int, rate: int, bits: int, tag: int = 1, extra_chunks=(), extensible=False,
               data_size=None, riff_size=None) -> bytes:
    align = channels * bits // 8
    if extensible:
        guid = struct.pack("<H", tag) + bytes.fromhex("000000001000800000aa00389b71")
        fmt = struct.pack("<HHIIHHHHI", 0xFFFE, channels, rate, rate * align, align, bits, 22, bits, 3) + guid
    else:
        fmt = struct.pack("<HHIIHH", tag, channels, rate, rate * align, align, bits)
    body = b"WAVE" + b"fmt " + struct.pack("<I", len(fmt)) + fmt
    for cid, payload in extra_chunks:
        body += cid + struct.pack("<I", len(payload)) + payload + (b"\0" if len(payload) & 1 else b"")
    body += b"data" + struct.pack("<I", len(frames) if data_size is None else data_size) + frames
    return b"RIFF" + struct.pack("<I", len(body) if riff_size is None else riff_size) + body


def _pcm(rng, n, ch, bits):
    if bits == 8:
        return rng.integers(0, 256, (n, ch), dtype=np.uint8).tobytes()
    if bits == 16:
        return rng.integers(-32768, 32768, (n, ch), dtype=np.int16).astype("<i2").tobytes()
    if bits == 32:
        return rng.integers(-2**31, 2**31, (n, ch), dtype=np.int64).astype("<i4").tobytes()
    v = rng.integers(-2**23, 2**23, (n, ch), dtype=np.int64)           # 24-bit: three little-endian bytes
    u = (v & 0xFFFFFF).astype(np.uint32)
    return np.stack([u & 255, (u >> 8) & 255, (u >> 16) & 255], -1).astype(np.uint8).tobytes()


CASES = [(8, 1, 1, np_wav.U8), (16, 1, 2, np_wav.S16), (24, 1, 2, np_wav.S24), (32, 1, 2, np_wav.S32),
         (32, 3, 2, np_wav.F32), (64, 3, 1, np_wav.F64), (16, 1, 1, np_wav.S16), (16, 1, 6, np_wav.S16)]


def _case_bytes(bits, tag, ch, n=1237, seed=0, **kw):
    rng = np.random.default_rng(seed + bits + ch)
    if tag == 3:
        frames = rng.uniform(-1.5, 1.5, (n, ch)).astype("<f4" if bits == 32 else "<f8").tobytes()
    else:
        frames = _pcm(rng, n, ch, bits)
    return _wav_bytes(frames, ch, 8000, bits, tag, **kw), frames


# ------------------------------------------------------------------ oracle vs independent readers
@pytest.mark.parametrize("bits,tag,ch,code", CASES)
def test_oracle_matches_wave_and_scipy(bits, tag, ch, code):
    from scipy.io import wavfile
    data, frames = _case_bytes(bits, tag, ch)
    h = np_wav.parse(data)
    assert (h["format"], h["channels"], h["sample_rate"], h["bits_per_sample"], h["num_frames"]) == (code, ch, 8000, bits, 1237)
    assert data[h["data_offset"]:h["data_offset"] + h["data_bytes"]] == frames
    if tag == 1:                                                     # stdlib wave reads PCM only
        with wave.open(io.BytesIO(data)) as w:
            assert (w.getnchannels(), w.getframerate(), w.getsampwidth() * 8, w.getnframes()) == (ch, 8000, bits, 1237)
            assert w.readframes(1237) == frames
    rate, x = wavfile.read(io.BytesIO(data))                          # scipy: raw values, (frames, channels)
    x = x.reshape(1237, ch)
    y, sr = np_wav.load(data)
    assert sr == rate == 8000 and y.dtype == np.float32 and y.shape == (ch, 1237)
    if code == np_wav.U8:
        ref = (x.astype(np.float64) - 128.0) / 128.0
    elif code == np_wav.S16:
        ref = x.astype(np.float64) / 32768.0
    elif code in (np_wav.S24, np_wav.S32):
        ref = x.astype(np.float64) / 2147483648.0                     # scipy widens 24-bit to the top of an int32
    else:
        ref = x.astype(np.float64)
    assert np.array_equal(y, ref.T.astype(np.float32))


# ------------------------------------------------------------------ C ABI parser vs oracle
def _same_header(data, total=None):
    code, got = emu_util.wav_parse(data, total)
    assert code == 0, got
    ref = np_wav.parse(data, total)
    assert {k: got[k] for k in ref} == ref
    return got


@pytest.mark.parametrize("bits,tag,ch,code", CASES)
def test_parse_formats(bits, tag, ch, code):
    data, _ = _case_bytes(bits, tag, ch)
    got = _same_header(data)
    assert got["format"] == code and got["block_align"] == ch * bits // 8 and got["data_offset"] == 44


def test_parse_chunks_extensible_streamed_truncated_prefix():
    frames = _pcm(np.random.default_rng(1), 500, 2, 16)
    # LIST (odd size -> pad byte) and fact chunks before the data; extensible fmt
    data = _wav_bytes(frames, 2, 8000, 16, extra_chunks=[(b"LIST", b"INFOISFT\x05\0\0\0Lavf\0"[:17]), (b"fact", b"\0" * 4)],
                      extensible=True)
    got = _same_header(data)
    assert got["num_frames"] == 500 and got["data_offset"] == 12 + 8 + 40 + 8 + 18 + 8 + 4 + 8
    # ffmpeg writing to a pipe (the reference's .ogg path): sizes are placeholders
    for placeholder in (0xFFFFFFFF, 0):
        got = _same_header(_wav_bytes(frames, 2, 8000, 16, data_size=placeholder, riff_size=0xFFFFFFFF))
        assert got["num_frames"] == 500
    # truncated file: the data chunk promises more than there is; a partial last frame is dropped
    got = _same_header(_wav_bytes(frames[:-3], 2, 8000, 16, data_size=len(frames)))
    assert got["num_frames"] == 499 and got["data_bytes"] == 499 * 4
    # trailing chunk after the data is not counted
    got = _same_header(_wav_bytes(frames, 2, 8000, 16) + b"LIST" + struct.pack("<I", 4) + b"abcd")
    assert got["num_frames"] == 500
    # header-only prefix of a longer file (torchaudio.info reads no samples)
    full = _wav_bytes(frames, 2, 8000, 16)
    got = _same_header(full[:64], len(full))
    assert got["num_frames"] == 500
    # empty data chunk
    got = _same_header(_wav_bytes(b"", 2, 8000, 16))
    assert got["num_frames"] == 0
    assert not np_wav.validate(_wav_bytes(b"", 2, 8000, 16))           # _validate_file: empty -> False
    assert not np_wav.validate(_wav_bytes(frames, 1, 8000, 16))        # mono -> False (pipeline.py:99-101)
    assert np_wav.validate(full)


def test_parse_errors():
    frames = _pcm(np.random.default_rng(2), 10, 2, 16)
    good = _wav_bytes(frames, 2, 8000, 16)
    assert emu_util.wav_parse(b"RIFX" + good[4:])[0] == -1
    assert emu_util.wav_parse(good[:8] + b"AVI " + good[12:])[0] == -1
    assert emu_util.wav_parse(good[:30])[0] == -1                      # fmt chunk cut
    assert emu_util.wav_parse(good[:40])[0] == -1                      # no data chunk within the bytes
    assert emu_util.wav_parse(_wav_bytes(frames, 2, 8000, 16, tag=0x55))[0] == -4      # MP3 in a wav container
    assert emu_util.wav_parse(_wav_bytes(frames, 2, 8000, 12))[0] in (-1, -4)
    bad_align = bytearray(good); bad_align[32] = 3
    assert emu_util.wav_parse(bytes(bad_align))[0] == -1
    assert emu_util.wav_parse(b"")[0] == -1


# ------------------------------------------------------------------ kernels through the host emulation
@pytest.mark.parametrize("bits,tag,ch,code", CASES)
@pytest.mark.parametrize("misalign", [0, 1, 3])
def test_emu_wav_decode(bits, tag, ch, code, misalign):
    data, frames = _case_bytes(bits, tag, ch, n=4100 + misalign)
    got = emu_util.wav_decode(frames, code, ch, misalign)
    assert np.array_equal(got, np_wav.decode(frames, code, ch), equal_nan=True)


def test_emu_wav_decode_extremes():
    s16 = np.array([[-32768, 32767], [0, -1], [1, 12345]], "<i2").tobytes()
    assert np.array_equal(emu_util.wav_decode(s16, np_wav.S16, 2), np_wav.decode(s16, np_wav.S16, 2))
    s32 = np.array([[-2**31], [2**31 - 1], [2**31 - 64], [16777217]], "<i4").tobytes()     # float32 rounding of int32
    assert np.array_equal(emu_util.wav_decode(s32, np_wav.S32, 1), np_wav.decode(s32, np_wav.S32, 1))
    s24 = bytes([0, 0, 0x80, 0xFF, 0xFF, 0x7F, 0xFF, 0xFF, 0xFF])
    assert emu_util.wav_decode(s24, np_wav.S24, 1).tolist() == [[-1.0, (2**23 - 1) / 2**23, -1 / 2**23]]
    u8 = bytes([0, 128, 255])
    assert emu_util.wav_decode(u8, np_wav.U8, 1).tolist() == [[-1.0, 0.0, 127 / 128]]


def test_emu_resample_wav16_equals_float_path():
    """Interleaved 16-bit frames -> de-interleave + x/32768 + 1:2 resample + stereo mean in one pass
    == resampling torchaudio.load's float32 waveform, bit for bit."""
    rng = np.random.default_rng(17)
    lens = [3, 1024, 1029, 2500, 4100, 0, 9000]
    pcm = [rng.integers(-32768, 32768, (2, L), dtype=np.int16) for L in lens]
    outs, mixes = emu_util.resample_batch(pcm, 8000, 16000, want_mixed=True, wav=True)
    for p, o, m in zip(pcm, outs, mixes):
        if p.shape[1] == 0:
            continue
        ref = np_resample.resample(p.astype(np.float32) / np.float32(32768.0), 8000, 16000)
        assert np.array_equal(o, ref)
        assert np.array_equal(m[0], (ref[0] + ref[1]) * np.float32(0.5))
    # rows as separate items (no mixed output), 2 and 3 channels
    for ch in (2, 3):
        p = rng.integers(-32768, 32768, (ch, 2777), dtype=np.int16)
        one = emu_util.resample_batch([p], 8000, 16000, wav=True)[0]
        assert np.array_equal(one, np_resample.resample(p.astype(np.float32) / np.float32(32768.0), 8000, 16000))


# ------------------------------------------------------------------ product path (B200)
@pytest.fixture(scope="module")
def fe():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from audio_processor_b200 import runtime
    return runtime.get_frontend(0)


@pytest.mark.gpu
@pytest.mark.parametrize("bits,tag,ch,code", CASES)
def test_gpu_load_audio_matches_oracle(fe, tmp_path, bits, tag, ch, code):
    from audio_processor_b200 import wav
    data, _ = _case_bytes(bits, tag, ch, n=70001, extra_chunks=[(b"LIST", b"x" * 7)])   # data chunk at an odd-ish offset
    path = tmp_path / "a.wav"
    path.write_bytes(data)
    y, sr = wav.load_audio(path, fe.device)
    ref, ref_sr = np_wav.load(data)
    assert sr == ref_sr and tuple(y.shape) == ref.shape
    assert np.array_equal(y.cpu().numpy(), ref, equal_nan=True)
    meta = wav.info(path)
    assert (meta.num_frames, meta.num_channels, meta.sample_rate, meta.bits_per_sample) == (70001, ch, 8000, bits)
    assert wav.validate_file(path) == np_wav.validate(data) == (ch != 1)


@pytest.mark.gpu
def test_gpu_validate_and_info_edge_cases(fe, tmp_path):
    from audio_processor_b200 import wav
    frames = _pcm(np.random.default_rng(3), 100, 2, 16)
    (tmp_path / "empty.wav").write_bytes(_wav_bytes(b"", 2, 8000, 16))
    (tmp_path / "junk.wav").write_bytes(b"not a wav file at all")
    big_list = _wav_bytes(frames, 2, 8000, 16, extra_chunks=[(b"LIST", b"z" * 100_000)])       # header beyond the first 4 KB read
    (tmp_path / "biglist.wav").write_bytes(big_list)
    assert not wav.validate_file(tmp_path / "missing.wav")
    assert not wav.validate_file(tmp_path / "empty.wav")
    assert not wav.validate_file(tmp_path / "junk.wav")
    assert wav.validate_file(tmp_path / "biglist.wav") and wav.info(tmp_path / "biglist.wav").num_frames == 100
    with pytest.raises(ValueError):
        wav.load_audio(tmp_path / "junk.wav", fe.device)            # load_audio re-raises (audio_processor.py:91-93)


@pytest.mark.gpu
def test_gpu_wav_route_equals_float_route(fe, tmp_path):
    """load_pcm -> split_audio and process_calls on the file's own frames == load_audio -> the float route."""
    import torch
    from audio_processor_b200 import AudioProcessor, CallFrontend
    rng = np.random.default_rng(23)
    ap = AudioProcessor({"chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000, "batch_size": 16, "gpu_index": 0})
    paths, lens = [], [240_000, 480_001, 30_000]
    for i, L in enumerate(lens):
        pcm = (rng.standard_normal((L, 2)) * 6000).clip(-32768, 32767).astype("<i2")
        p = tmp_path / f"call{i}.wav"
        p.write_bytes(_wav_bytes(pcm.tobytes(), 2, 8000, 16, extra_chunks=[(b"LIST", b"q" * (2 * i + 2))]))
        paths.append(p)
    floats = [ap.load_audio(p) for p in paths]
    pcms = [ap.load_pcm(p) for p in paths]
    for (xf, sr_f), (xp, sr_p), p in zip(floats, pcms, paths):
        assert sr_f == sr_p == 8000 and tuple(xf.shape) == tuple(xp.shape)
        assert fe.pcm_kind(xp) == "wav" and xp.is_pinned()
        assert np.array_equal(xf.cpu().numpy(), np_wav.load(p.read_bytes())[0])
        ca, a1, c1 = ap.split_audio(xf, 8000, p.name)
        cb, a2, c2 = ap.split_audio(xp, 8000, p.name)
        assert torch.equal(a1, a2) and torch.equal(c1, c2) and len(ca) == len(cb)
        for u, v in zip(ca, cb):
            assert torch.equal(u["stereo_waveform"], v["stereo_waveform"]) and u["end_time"] == v["end_time"]
    cf = CallFrontend(fe.device)
    ra = cf.process_calls([x for x, _ in floats], 8000)
    rb = cf.process_calls([x for x, _ in pcms], 8000)
    assert ra.chunks == rb.chunks and ra.order == rb.order
    for key in ("stereo", "agent", "client"):
        for (fa, ma), (fb, mb) in zip(ra.features[key], rb.features[key]):
            assert torch.equal(fa, fb) and torch.equal(ma, mb)
    assert torch.equal(ra.acoustic_rows, rb.acoustic_rows)
    # 16 kHz wav frames need no resampling: decoded, mixed and chunked directly
    pcm16k = (rng.standard_normal((50_000, 2)) * 6000).astype("<i2")
    p = tmp_path / "wide.wav"
    p.write_bytes(_wav_bytes(pcm16k.tobytes(), 2, 16000, 16))
    xp, sr = ap.load_pcm(p)
    chunks, agent, client = ap.split_audio(xp, sr, "wide.wav")
    ref = np_wav.load(p.read_bytes())[0]
    assert np.array_equal(agent.cpu().numpy()[0], ref[0]) and np.array_equal(client.cpu().numpy()[0], ref[1])
    assert np.array_equal(chunks[0]["stereo_waveform"].cpu().numpy()[0], (ref[0] + ref[1]) * np.float32(0.5))
