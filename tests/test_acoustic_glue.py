"""Acoustic classifier glue: ``analyze_batch_sentiment`` with the StandardScaler step fused
into the last acoustic kernel (services/sentiment_analysis.py:696-805).

CPU: the epilogue through the host emulation against numpy / sklearn (bit-exact float64), and the whole
method against the UNMODIFIED reference method with real fitted sklearn models (container only).
GPU: the product path against sklearn on its own feature lists (bit-exact) and on the oracle's."""
import numpy as np
import pytest
import torch

from oracle import np_acoustic, reference_calls as rc
from tests import emu_util


def _waves(seed=0, lens=(16000, 4000, 399, 48000, 560, 25000)):
    rng = np.random.default_rng(seed)
    out = []
    for n in lens:
        t = np.arange(n) / 16000.0
        x = 0.3 * np.sin(2 * np.pi * rng.uniform(100, 3000) * t) * (0.5 + 0.5 * np.sin(2 * np.pi * 3 * t)) + 0.01 * rng.standard_normal(n)
        out.append(x.astype(np.float32))
    return out


def _tempo(flux, n):
    from audio_processor_b200.acoustic import _tempo as t
    return t(flux, n, 16000)


def _lists(rows, status, waves, n_features):
    """The feature lists the reference hands to scaler.transform (:736-748)."""
    out = []
    for r, st, w in zip(rows, status, waves):
        vals = [0.0] * 38 if st != 0 else [float(v) for v in r]
        if st == 0:
            vals[11] = _tempo(vals[11], w.size)
        vals = vals[:n_features] + [0.0] * max(0, n_features - 38)
        out.append(vals)
    return out


@pytest.mark.parametrize("n_features", [38, 41, 30])
def test_emu_scaler_epilogue_is_numpy_float64(n_features):
    from sklearn.preprocessing import StandardScaler
    waves = _waves(lens=(16000, 4000, 399, 48000, 400, 25000))
    rng = np.random.default_rng(1)
    sc = StandardScaler().fit(rng.standard_normal((50, n_features)) * rng.uniform(0.1, 2000, n_features) + rng.uniform(-50, 50, n_features))
    rows, status, scaled = emu_util.acoustic_scaled(waves, sc.mean_, sc.scale_)
    rows0, status0 = emu_util.acoustic(waves)
    assert np.array_equal(rows, rows0, equal_nan=True) and np.array_equal(status, status0)     # float32 rows untouched
    assert status.tolist() == [0, 0, 1, 0, 0, 0]
    ref = sc.transform(_lists(rows, status, waves, n_features))
    assert scaled.dtype == np.float64 and np.array_equal(scaled, ref, equal_nan=True)    # (L = 400: one frame, std = NaN)
    # the failed waveform is the scaled ZERO feature vector, which is a valid row for the classifier
    assert np.array_equal(scaled[2], (0.0 - sc.mean_) / sc.scale_)
    assert rows[0, 11] > 0 and ref[0, 11] != rows[0, 11] if n_features > 11 else True


class _EmuGlue:
    """The product glue with its single device step replaced by the host emulation of the same kernels."""
    def __new__(cls, **kw):
        from audio_processor_b200.acoustic import AcousticSentimentAnalyzer

        class G(AcousticSentimentAnalyzer):
            def _scaled_features(self, waveforms, sample_rate, scaler):
                mean = scaler.mean_ if scaler.with_mean else np.zeros(scaler.n_features_in_)
                scale = scaler.scale_ if scaler.with_std else np.ones(scaler.n_features_in_)
                return emu_util.acoustic_scaled([w.numpy().reshape(-1) for w in waveforms], mean, scale, sample_rate)[2]
        return G(**kw)


def _fit(seed, waves):
    """A StandardScaler + probabilistic classifier fitted on real feature vectors of perturbed waveforms."""
    from sklearn.linear_model import LogisticRegression
    from sklearn.preprocessing import StandardScaler
    rng = np.random.default_rng(seed)
    X, y = [], []
    for k in range(24):
        w = waves[k % len(waves)]
        if w.size < 400:
            continue
        f = np_acoustic.acoustic_features(np.roll(w, 37 * k) * np.float32(0.5 + 0.05 * k))
        X.append(list(f.values())); y.append(["neg", "neu", "pos"][k % 3])
    sc = StandardScaler().fit(np.asarray(X))
    clf = LogisticRegression(max_iter=500).fit(sc.transform(np.asarray(X)), y)
    return clf, sc


def test_glue_contract_without_reference():
    waves = _waves(3)
    clf, sc = _fit(0, waves)
    g = _EmuGlue(agent_acoustic_model=clf, agent_acoustic_scaler=sc, client_acoustic_model=clf, client_acoustic_scaler=sc)
    batch = [torch.from_numpy(w)[None] for w in waves]
    batch.insert(2, None); batch.insert(4, torch.zeros((1, 0)))
    res = g.analyze_batch_sentiment(batch, 16000, "agent")
    assert len(res) == len(batch)
    assert res[2] == {"prediction": "", "confidence": 0.0, "probabilities": []} == res[4]
    good = [r for i, r in enumerate(res) if i not in (2, 4)]
    assert all(r["prediction"] in ("neg", "neu", "pos") and abs(sum(r["probabilities"]) - 1) < 1e-9 for r in good)
    assert all(r["confidence"] == max(r["probabilities"]) for r in good)
    assert g.analyze_batch_sentiment([], 16000, "agent") == []
    none = _EmuGlue()
    assert none.analyze_batch_sentiment(batch[:2], 16000, "client") == [{"prediction": "", "confidence": 0.0, "probabilities": []}] * 2
    half = _EmuGlue(agent_acoustic_model=clf, agent_acoustic_scaler=sc, client_acoustic_model=clf, client_acoustic_scaler=sc)
    half.client_acoustic_scaler = None
    assert half.analyze_batch_sentiment(batch[:2], 16000, "client")[0]["prediction"] == ""
    assert half.analyze_batch_sentiment(batch[:2], 16000, "agent")[0]["prediction"] != ""


@pytest.mark.skipif(not rc.available(), reason="/root/reference not mounted")
@pytest.mark.parametrize("speaker", ["agent", "client"])
def test_glue_matches_reference_method(speaker):
    waves = _waves(5)
    clf_a, sc_a = _fit(1, waves)
    clf_c, sc_c = _fit(2, waves)
    ref = rc.acoustic_analyzer()
    names = dict(agent_acoustic_model=clf_a, agent_acoustic_scaler=sc_a, agent_acoustic_id2label=None,
                 client_acoustic_model=clf_c, client_acoustic_scaler=sc_c, client_acoustic_id2label=None)
    for k, v in names.items():
        setattr(ref, k, v)
    ref.models_available = True
    batch = [torch.from_numpy(w)[None] for w in waves]
    batch.insert(1, None)
    try:
        a = ref.analyze_batch_sentiment(batch, 16000, speaker)
    finally:
        ref.models_available = False
    from audio_processor_b200.acoustic import AcousticSentimentAnalyzer
    g = _EmuGlue(**names)
    assert isinstance(g, AcousticSentimentAnalyzer)
    b = g.analyze_batch_sentiment(batch, 16000, speaker)
    assert len(a) == len(b) == len(batch)
    assert sum(1 for r in a if r["prediction"] != "") == len(waves)
    for ra, rb in zip(a, b):
        assert ra["prediction"] == rb["prediction"]
        assert abs(ra["confidence"] - rb["confidence"]) < 2e-3          # float32 reference features vs float64 kernels
        assert np.allclose(ra["probabilities"], rb["probabilities"], atol=2e-3)
    # a one-frame waveform (L = 400) has NaN standard deviations: sklearn refuses the WHOLE matrix, the
    # reference's outer except turns the batch into empty results -- and so does the glue
    nan_batch = batch + [torch.from_numpy(_waves(9, lens=(400,))[0])[None]]
    ref.models_available = True
    try:
        a = ref.analyze_batch_sentiment(nan_batch, 16000, speaker)
    finally:
        ref.models_available = False
    assert a == g.analyze_batch_sentiment(nan_batch, 16000, speaker) == [{"prediction": "", "confidence": 0.0, "probabilities": []}] * len(nan_batch)


# ------------------------------------------------------------------ product path (B200)
@pytest.mark.gpu
def test_gpu_glue_scaled_matrix_and_predictions():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from audio_processor_b200.acoustic import AcousticSentimentAnalyzer
    waves = _waves(7, lens=(16000, 4000, 399, 400_000, 560, 25000, 8001))
    clf, sc = _fit(4, waves)
    g = AcousticSentimentAnalyzer(0, agent_acoustic_model=clf, agent_acoustic_scaler=sc, client_acoustic_model=clf, client_acoustic_scaler=sc)
    tens = [torch.from_numpy(w)[None] for w in waves]
    scaled = g._scaled_features(tens, 16000, sc)
    lists = g.extract_batch(tens, 16000)                              # the reference's per-waveform lists (zeros on failure)
    assert np.array_equal(scaled, sc.transform(lists), equal_nan=True)   # same float64 operations as sklearn
    res = g.analyze_batch_sentiment(tens + [None], 16000, "client")
    ref_lists = [list(np_acoustic.acoustic_features(w).values()) if w.size >= 400 else [0.0] * 38 for w in waves]
    ref_proba = clf.predict_proba(sc.transform(ref_lists))
    ref_pred = clf.predict(sc.transform(ref_lists))
    for r, p, q in zip(res, ref_pred, ref_proba):
        assert r["prediction"] == p and np.allclose(r["probabilities"], q, atol=2e-3)
    assert res[-1]["prediction"] == ""
