"""Kernel logic vs oracle / golden, run through the C ABI of the host emulation (CPU)."""
import numpy as np
import pytest

from oracle import np_fbank, np_resample
from tests import emu_util


def _split(flat, lengths):
    out, o = [], 0
    for n in lengths:
        out.append(flat[o:o + int(n)]); o += int(n)
    return out


def _cmp_nan(a, b, tol):
    assert a.shape == b.shape
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = ~np.isnan(b)
    return float(np.max(np.abs(a[ok] - b[ok]), initial=0.0)) <= tol


# ------------------------------------------------------------------ resample: bit-exact
@pytest.mark.parametrize("src,pre", [("x8", ""), ("xw8", "w_")])
def test_emu_resample_split_bit_exact(golden, src, pre):
    g = golden("resample")
    out, mixed = emu_util.resample(g[src], 8000, 16000, want_mixed=True)
    assert np.array_equal(out[0:1], g[pre + "agent16"])
    assert np.array_equal(out[1:2], g[pre + "client16"])
    assert np.array_equal(mixed, g[pre + "mixed16"])


def test_emu_resample_mono_and_44k(golden):
    g = golden("resample")
    assert np.array_equal(emu_util.resample(g["xm8"], 8000, 16000), g["m_agent16"])
    assert np.array_equal(emu_util.resample(g["x44_long"], 44100, 16000), g["y44to16_long"])
    assert np.max(np.abs(emu_util.resample(g["x44"], 44100, 16000) - g["y44to16"])) <= 1.2e-7


@pytest.mark.parametrize("L", [1, 3, 7, 1023, 1024, 1025, 2049])
def test_emu_resample_ragged_lengths(L):
    x = (np.random.default_rng(L).uniform(-0.5, 0.5, (2, L))).astype(np.float32)
    out, mixed = emu_util.resample(x, 8000, 16000, want_mixed=True)
    ref = np_resample.resample(x, 8000, 16000)
    assert np.array_equal(out, ref)
    assert np.array_equal(mixed[0], (ref[0] + ref[1]) * np.float32(0.5))


def test_emu_resample_batch_matches_per_call():
    rng = np.random.default_rng(11)
    lens = [1, 2, 5, 1023, 1024, 1026, 2500, 4097, 0, 3001]
    calls = [rng.uniform(-0.5, 0.5, (2, L)).astype(np.float32) for L in lens]
    outs, mixes = emu_util.resample_batch(calls, 8000, 16000, want_mixed=True)
    for c, o, m in zip(calls, outs, mixes):
        ref = np_resample.resample(c, 8000, 16000) if c.shape[1] else np.zeros((2, 0), np.float32)
        assert np.array_equal(o, ref)
        assert np.array_equal(m[0], (ref[0] + ref[1]) * np.float32(0.5))
    outs = emu_util.resample_batch(calls, 8000, 16000)          # no mixed output: rows are separate items
    for c, o in zip(calls, outs):
        if c.shape[1]:
            assert np.array_equal(o, np_resample.resample(c, 8000, 16000))
    mono = [c[:1] for c in calls if c.shape[1]]
    for c, o in zip(mono, emu_util.resample_batch(mono, 8000, 16000)):
        assert np.array_equal(o, np_resample.resample(c, 8000, 16000))
    x = rng.uniform(-0.5, 0.5, (1, 3000)).astype(np.float32)    # other rate pairs loop over the generic kernel
    pair = [x, x[:, :1777]]
    for c, o in zip(pair, emu_util.resample_batch(pair, 16000, 8000)):
        assert np.array_equal(o, np_resample.resample(c, 16000, 8000))


def test_emu_resample_pcm16_equals_float_path():
    """int16 PCM staged as x / 32768 == resampling torchaudio.load's float32 waveform."""
    rng = np.random.default_rng(13)
    lens = [3, 1024, 1029, 2500, 4100]
    pcm = [rng.integers(-32768, 32768, (2, L), dtype=np.int16) for L in lens]
    outs, mixes = emu_util.resample_batch(pcm, 8000, 16000, want_mixed=True, pcm16=True)
    for p, o, m in zip(pcm, outs, mixes):
        ref = np_resample.resample(p.astype(np.float32) / np.float32(32768.0), 8000, 16000)
        assert np.array_equal(o, ref)
        assert np.array_equal(m[0], (ref[0] + ref[1]) * np.float32(0.5))
    one = emu_util.resample_batch(pcm[3:4], 8000, 16000, pcm16=True)[0]       # single call, rows as items
    assert np.array_equal(one, np_resample.resample(pcm[3].astype(np.float32) / np.float32(32768.0), 8000, 16000))


def test_emu_resample_generic_rates():
    x = (np.random.default_rng(5).uniform(-0.5, 0.5, (1, 3000))).astype(np.float32)
    for orig, new in ((16000, 8000), (48000, 16000), (22050, 16000)):
        assert np.array_equal(emu_util.resample(x, orig, new), np_resample.resample(x, orig, new)), (orig, new)


# ------------------------------------------------------------------ fbank
def test_emu_fbank_ragged_batch(golden):
    g = golden("fbank")
    waves = _split(g["ragged_waves"], g["ragged_lengths"])
    feat, mask = emu_util.fbank(waves)
    assert np.array_equal(mask, g["ragged_mask"])
    assert _cmp_nan(feat, g["ragged_features"], 1e-3)


@pytest.mark.parametrize("n", [399, 400, 559, 560, 720])
def test_emu_fbank_degenerate(golden, n):
    g = golden("fbank")
    feat, mask = emu_util.fbank([g[f"single_{n}_wave"]])
    assert np.array_equal(mask, g[f"single_{n}_mask"])
    assert _cmp_nan(feat, g[f"single_{n}_features"], 1e-3)


@pytest.mark.parametrize("name", ["up8k", "white", "tone1k", "zeros"])
def test_emu_fbank_signals(golden, name):
    g = golden("fbank")
    w = g[f"{name}_wave"]
    raw, _ = emu_util.fbank([w], raw=True)
    T = np_fbank.num_frames(w.size)
    raw = raw.reshape(1, -1, 80)[0, :T]
    assert np.max(np.abs(raw - g[f"{name}_raw"])) <= (2e-4 if name == "tone1k" else 2e-5), name
    if name in ("up8k", "white"):
        feat, mask = emu_util.fbank([w])
        assert np.array_equal(mask, g[f"{name}_mask"])
        assert _cmp_nan(feat, g[f"{name}_features"], 1e-3 if name == "up8k" else 2e-4)


def test_emu_fbank_padding_value_and_f32(golden):
    g = golden("fbank")
    waves = _split(g["ragged_waves"], g["ragged_lengths"])
    feat, mask = emu_util.fbank(waves, padding_value=1.0)
    ref = g["ragged_features"].reshape(len(waves), -1, 80)
    got = feat.reshape(len(waves), -1, 80)
    for b, n in enumerate(g["ragged_lengths"]):
        T = np_fbank.num_frames(int(n))
        assert np.all(got[b, T:] == 1.0)
        assert _cmp_nan(got[b, :T], ref[b, :T], 1e-3)
    feat32, mask32 = emu_util.fbank(waves, precision=1)
    assert np.array_equal(mask32, mask)
    assert _cmp_nan(feat32, g["ragged_features"], 2e-2)      # opt-in fast path: looser by design


def test_emu_fbank_grouped_matches_per_group_calls(golden):
    g = golden("fbank")
    waves = _split(g["ragged_waves"], g["ragged_lengths"])
    order = [6, 0, 5, 3, 1, 2, 4]
    waves = [waves[i] for i in order]
    groups = emu_util.fbank_grouped(waves, 3)
    assert len(groups) == 3
    for gi, (feat, mask) in enumerate(groups):
        ref_f, ref_m = np_fbank.featurise(waves[3 * gi:3 * gi + 3])
        assert np.array_equal(mask, ref_m)
        assert _cmp_nan(feat, ref_f, 1e-3)


# ------------------------------------------------------------------ acoustic
def _tempo_from_flag(flag, L):
    F2 = 1 + L // 160
    lag = (F2 - 2) if flag > 0 else 0
    period = lag * 160 / 16000 if lag > 0 else 0.0
    return 60.0 / period if period > 0 else 0.0


def test_emu_acoustic_golden(golden):
    g = golden("acoustic")
    names = [str(n) for n in g["names"]]
    waves = [g[f"{n}_wave"] for n in names] + [np.zeros(399, np.float32)]
    out, status = emu_util.acoustic(waves)
    assert status.tolist() == [0] * len(names) + [1]
    assert np.all(out[-1] == 0)
    for i, name in enumerate(names):
        ref = g[f"{name}_values"]
        got = out[i].astype(np.float64)
        got[11] = _tempo_from_flag(out[i, 11], waves[i].size)
        assert got[11] == ref[11], name                                     # tempo exact
        assert np.array_equal(np.isnan(got), np.isnan(ref)), name
        ok = ~np.isnan(ref)
        rel = 5e-3 if name.split("_")[0] in ("tone", "dc", "square") else 1e-4
        tol = np.maximum(1e-3, rel * np.abs(ref))
        bad = [(k, got[k], ref[k]) for k in range(38) if ok[k] and abs(got[k] - ref[k]) > tol[k]]
        assert not bad, (name, bad)


def test_emu_fbank_fused_and_fallback_normalisation_agree():
    """An utterance normalised by the fbank kernel itself (all its tiles in one block) and the
    same utterance finished by the stats / norm kernels (tiles split over blocks) give the same
    bits: alone, each waveform is split over the emulation's 3 blocks; in the batch most are whole."""
    rng = np.random.default_rng(77)
    lens = [400, 5000, 16000, 40000, 33333, 90000, 12345, 70001]
    waves = [rng.uniform(-0.5, 0.5, n).astype(np.float32) for n in lens]
    batch = emu_util.fbank_grouped(waves, 1)
    for w, (fb, mb) in zip(waves, batch):
        fa, ma = emu_util.fbank_grouped([w], 1)[0]
        assert np.array_equal(ma, mb)
        assert np.array_equal(fa.view(np.uint32), fb.view(np.uint32))
