"""Parity of the CUDA path (through the C ABI / runtime) against golden vectors and the
oracle.  Needs a B200; run with ``pytest -m gpu``."""
import numpy as np
import pytest

from oracle import np_fbank, np_resample, np_tables

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from audio_processor_b200 import runtime
    return runtime.get_frontend(0)


def _t(a, fe):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(fe.device)


def _split(flat, lengths):
    out, o = [], 0
    for n in lengths:
        out.append(flat[o:o + int(n)]); o += int(n)
    return out


def _cmp_nan(a, b, tol):
    assert a.shape == b.shape
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = ~np.isnan(b)
    err = float(np.max(np.abs(a[ok] - b[ok]), initial=0.0))
    assert err <= tol, err


def test_native_library_is_cuda_build(fe):
    assert fe.lib.backend == 1 and fe.num_sms >= 100


# ------------------------------------------------------------------ resample (bit-exact)
@pytest.mark.parametrize("src,pre", [("x8", ""), ("xw8", "w_")])
def test_resample_split_bit_exact(fe, golden, src, pre):
    g = golden("resample")
    out, mixed = fe.resample(_t(g[src], fe), 8000, 16000, want_mixed=True)
    out, mixed = out.cpu().numpy(), mixed.cpu().numpy()
    assert np.array_equal(out[0:1], g[pre + "agent16"])
    assert np.array_equal(out[1:2], g[pre + "client16"])
    assert np.array_equal(mixed, g[pre + "mixed16"])


def test_resample_mono_44k_generic(fe, golden):
    g = golden("resample")
    assert np.array_equal(fe.resample(_t(g["xm8"], fe), 8000, 16000).cpu().numpy(), g["m_agent16"])
    assert np.array_equal(fe.resample(_t(g["x44_long"], fe), 44100, 16000).cpu().numpy(), g["y44to16_long"])
    x = (np.random.default_rng(5).uniform(-0.5, 0.5, (1, 3000))).astype(np.float32)
    for orig, new in ((16000, 8000), (48000, 16000), (22050, 16000)):
        assert np.array_equal(fe.resample(_t(x, fe), orig, new).cpu().numpy(), np_resample.resample(x, orig, new))


@pytest.mark.parametrize("L", [1, 3, 7, 1023, 1024, 1025, 2049, 100001])
def test_resample_ragged_lengths(fe, L):
    x = (np.random.default_rng(L).uniform(-0.5, 0.5, (2, L))).astype(np.float32)
    out, mixed = fe.resample(_t(x, fe), 8000, 16000, want_mixed=True)
    ref = np_resample.resample(x, 8000, 16000)
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(mixed.cpu().numpy()[0], (ref[0] + ref[1]) * np.float32(0.5))


# ------------------------------------------------------------------ fbank
def test_fbank_ragged_batch(fe, golden):
    g = golden("fbank")
    waves = _split(g["ragged_waves"], g["ragged_lengths"])
    feat, mask = fe.fbank(waves)
    assert np.array_equal(mask.cpu().numpy(), g["ragged_mask"])
    _cmp_nan(feat.cpu().numpy(), g["ragged_features"], 1e-3)      # tolerance of the north star


@pytest.mark.parametrize("n", [399, 400, 559, 560, 720])
def test_fbank_degenerate(fe, golden, n):
    g = golden("fbank")
    feat, mask = fe.fbank([g[f"single_{n}_wave"]])
    assert np.array_equal(mask.cpu().numpy(), g[f"single_{n}_mask"])
    _cmp_nan(feat.cpu().numpy(), g[f"single_{n}_features"], 1e-3)


@pytest.mark.parametrize("name", ["up8k", "white", "zeros", "tone1k"])
def test_fbank_signals(fe, golden, name):
    g = golden("fbank")
    w = g[f"{name}_wave"]
    feat, mask = fe.fbank([w])
    assert np.array_equal(mask.cpu().numpy(), g[f"{name}_mask"])
    # zeros: every bin sits at the log floor in every frame; the reference's output there is its own float32
    #   rounding noise / sqrt(1e-7) (3.0e-3 for these 23 frames) and fbank_mean_den reproduces numpy's pairwise
    #   mean / var bit for bit, so the result is EXACT.
    # tone1k: 1 kHz at 16 kHz repeats every 16 samples and the hop is 160, so all frames are identical and every
    #   bin is constant: the reference's normalised output is again pure rounding noise, now of raw values that
    #   the two implementations agree on to 1 ulp only (raw parity 2e-4 below) -- each ulp becomes 1.5e-3, the
    #   worst bin 6.0e-3.  Ill-conditioned in the reference itself; bounded, not matched.
    tol = {"zeros": 0.0, "tone1k": 7e-3}.get(name, 1e-3)
    _cmp_nan(feat.cpu().numpy(), g[f"{name}_features"], tol)
    raw, _ = fe.fbank([w], raw=True)
    T = np_fbank.num_frames(w.size)
    raw_tol = 2e-4 if name == "tone1k" else 5e-5
    assert np.max(np.abs(raw.cpu().numpy().reshape(1, -1, 80)[0, :T] - g[f"{name}_raw"])) <= raw_tol


@pytest.mark.parametrize("group", [0, 2])
def test_fbank_digital_silence_equals_reference_noise(fe, group):
    """Whole utterances of digital silence (a muted channel), alone and next to speech, fused and split across
    blocks: the normalised features equal the oracle's (numpy's own float32 mean / var order) to 1e-6, where
    round 1 returned 0 against up to 6e-3."""
    rng = np.random.default_rng(5)
    lens = [4000, 16000, 400_000, 48_000, 123_456, 30_000]
    waves = [np.zeros(n, np.float32) for n in lens]
    waves[3] = (rng.random(48_000, dtype=np.float32) - 0.5)
    out = fe.fbank(waves, group_size=group)
    groups = out if group else [out]
    step = group if group else len(waves)
    noisy = 0.0
    for gi, (feat, mask) in enumerate(groups):
        ref_f, ref_m = np_fbank.featurise(waves[gi * step:(gi + 1) * step])
        assert np.array_equal(mask.cpu().numpy(), ref_m)
        got = feat.cpu().numpy()
        for b in range(got.shape[0]):
            tol = 1e-3 if gi * step + b == 3 else 1e-6
            _cmp_nan(got[b], ref_f[b], tol)
        noisy = max(noisy, float(np.nanmax(np.abs(ref_f))))
    assert noisy > 1e-3                      # the reference really is non-zero on some of these


def test_fbank_grouped_and_device_inputs(fe, golden):
    g = golden("fbank")
    waves = _split(g["ragged_waves"], g["ragged_lengths"])
    order = [6, 0, 5, 3, 1, 2, 4]
    waves = [waves[i] for i in order]
    groups = fe.fbank([_t(w, fe) for w in waves], group_size=3, padding_value=1.0)
    assert len(groups) == 3
    for gi, (feat, mask) in enumerate(groups):
        ref_f, ref_m = np_fbank.featurise(waves[3 * gi:3 * gi + 3], padding_value=1.0)
        assert np.array_equal(mask.cpu().numpy(), ref_m)
        _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)


def test_fbank_f32_fast_path_is_looser(fe, golden):
    g = golden("fbank")
    from audio_processor_b200 import _native
    feat, mask = fe.fbank([g["up8k_wave"]], precision=_native.PRECISION_F32)
    assert np.array_equal(mask.cpu().numpy(), g["up8k_mask"])
    _cmp_nan(feat.cpu().numpy(), g["up8k_features"], 2e-2)


def test_chain_cfg1_resample_chunk_fbank(fe):
    """cfg1: 60 s stereo 8 kHz call -> resample/split -> 25 s chunks -> fbank of each stream."""
    from audio_processor_b200 import synth
    call, sr = synth.cfg1_call()
    out, mixed = fe.resample(call.to(fe.device), sr, 16000, want_mixed=True)
    ref = np_resample.resample(call.numpy(), sr, 16000)
    assert np.array_equal(out.cpu().numpy(), ref)                 # bit-exact resample
    table = np_tables.chunk_table(out.shape[1], 16000, 25, 1)
    assert [(s, e) for s, e, _, _ in table] == [(0, 400000), (384000, 784000), (768000, 960000)]
    agent = out[0]
    chunks = [agent[s:e] for s, e, _, _ in table]
    feat, mask = fe.fbank(chunks)
    ref_f, ref_m = np_fbank.featurise([ref[0, s:e] for s, e, _, _ in table])
    assert tuple(feat.shape) == (3, 1249, 160) and mask.sum(1).tolist() == [1249, 1249, 599]
    assert np.array_equal(mask.cpu().numpy(), ref_m)
    _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)


# ------------------------------------------------------------------ acoustic
def test_acoustic_golden(fe, golden):
    from audio_processor_b200 import acoustic
    g = golden("acoustic")
    names = [str(n) for n in g["names"]]
    waves = [g[f"{n}_wave"] for n in names] + [np.zeros(399, np.float32)]
    feats = acoustic.extract_acoustic_features_batch(waves, 16000, fe.device)
    assert feats[-1] is None
    for name, f in zip(names, feats):
        ref = g[f"{name}_values"]
        assert list(f.keys()) == [str(k) for k in g[f"{name}_keys"]]
        got = np.asarray(list(f.values()), np.float64)
        assert got[11] == ref[11], name
        assert np.array_equal(np.isnan(got), np.isnan(ref)), name
        ok = ~np.isnan(ref)
        rel = 5e-3 if name.split("_")[0] in ("tone", "dc", "square") else 1e-4
        tol = np.maximum(1e-3, rel * np.abs(ref))
        bad = [(k, got[k], ref[k]) for k in range(38) if ok[k] and abs(got[k] - ref[k]) > tol[k]]
        assert not bad, (name, bad)


def test_acoustic_drop_in_object(fe, golden):
    import torch
    from audio_processor_b200 import acoustic
    g = golden("acoustic")
    ex = acoustic.AcousticFeatureExtractor(fe.device)
    f = ex._extract_acoustic_features(torch.from_numpy(g["white_32000_wave"])[None, :], 16000)
    assert len(f) == 38 and abs(f["mfcc_0_mean"] - g["white_32000_values"][12]) <= 1e-3
    assert ex._extract_acoustic_features(torch.zeros(1, 399), 16000) is None
    assert list(ex._get_zero_features().keys()) == acoustic.FEATURE_KEYS


# ------------------------------------------------------------------ drop-in classes
def test_audio_processor_drop_in(fe):
    import torch
    from audio_processor_b200 import AudioProcessor, synth
    cfg = {"chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000, "batch_size": 16, "gpu_index": 0}
    ap = AudioProcessor(cfg)
    call, sr = synth.cfg1_call()
    chunks, agent, client = ap.split_audio(call, sr, "f.wav")
    assert [c["chunk_idx"] for c in chunks] == [0, 1, 2]
    assert [(c["start_time"], c["end_time"]) for c in chunks] == [(0.0, 25.0), (24.0, 49.0), (48.0, 60.0)]
    assert list(chunks[0].keys()) == ["file_name", "stereo_waveform", "agent_waveform", "client_waveform",
                                      "chunk_idx", "start_time", "end_time"]
    ref = np_resample.resample(call.numpy(), sr, 16000)
    assert np.array_equal(agent.cpu().numpy(), ref[0:1]) and np.array_equal(client.cpu().numpy(), ref[1:2])
    assert np.array_equal(chunks[2]["stereo_waveform"].cpu().numpy()[0], ((ref[0] + ref[1]) * np.float32(0.5))[768000:])
    # the processor() call of _run_inference on the length-sorted batch
    batches = ap.bucket_chunks(chunks)
    assert [c["chunk_idx"] for c in batches[0]] == [2, 0, 1]
    out = ap.featurise_batch(batches[0])
    ref_f, ref_m = np_fbank.featurise([ref[1, 768000:], ref[1, :400000], ref[1, 384000:784000]])
    assert np.array_equal(out["client_mask"].cpu().numpy(), ref_m)
    _cmp_nan(out["client"].cpu().numpy(), ref_f, 1e-3)
    allb = ap.featurise_all(chunks)
    assert torch.equal(allb[0][1]["client"][0], out["client"])
    # resampler object and the ValueError of a wrong sampling rate
    assert ap.get_resampler(8000, 16000) is ap.get_resampler(8000, 16000)
    with pytest.raises(ValueError):
        ap.processor([np.zeros(1000, np.float32)], sampling_rate=8000, return_tensors="pt", padding=True)


def test_speech_segment_drop_in(fe, golden):
    import torch
    from audio_processor_b200 import SpeechSegment
    g = golden("tables")
    cfg = {"gpu_index": 0, "batch_size": 4, "vad_max_segment_sec": 25.0, "vad_window_overlap_sec": 0.5,
           "vad_gap_threshold": 0.8, "vad_padding": 0.5}
    n = int(g["seg_num_samples"][0])
    wf = (torch.rand(1, n, generator=torch.Generator().manual_seed(3)) - 0.5).to(fe.device)
    seg = SpeechSegment(cfg)
    ex = seg._extract_segments(wf, 16000, [tuple(r) for r in g["seg_intervals"]], "agent", 0.5)
    exp = g["seg_cfg"]
    assert len(ex) == len(exp)
    for e, r in zip(ex, exp):
        assert e["segment_waveform"].shape[1] == int(r[0]) and e["start"] == r[1] and e["end"] == r[2]
        assert e["speaker"] == "agent"
    assert np.array_equal(np.asarray(seg._merge_segments([tuple(r) for r in g["merge_in"]], 0.8)), g["merge_out"])
    groups = seg.featurise_segments(ex[:5], 16000)
    lens = sorted((e["segment_waveform"].shape[1], i) for i, e in enumerate(ex[:5]))
    assert [i for idxs, _, _ in groups for i in idxs] == [i for _, i in lens]
    idxs, feat, mask = groups[0]
    ref_f, ref_m = np_fbank.featurise([ex[i]["segment_waveform"][0].cpu().numpy() for i in idxs])
    assert np.array_equal(mask.cpu().numpy(), ref_m)
    _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)


def test_call_frontend_whole_path(fe):
    from audio_processor_b200 import CallFrontend, synth
    from oracle import np_acoustic
    calls = [synth.make_call(30.0, 8000, 77, 2), synth.make_call(12.5, 8000, 78, 2)]
    cf = CallFrontend(fe.device, batch_size=2)
    res = cf.process_calls(calls, 8000)
    assert [(c, i) for c, i, _, _ in res.chunks] == [(0, 0), (0, 1), (1, 0)]
    assert res.order == [1, 2, 0] and len(res.features["agent"]) == 2
    ref0 = np_resample.resample(calls[0].numpy(), 8000, 16000)
    ref1 = np_resample.resample(calls[1].numpy(), 8000, 16000)
    ref_f, ref_m = np_fbank.featurise([ref0[0, 384000:], ref1[0]])
    feat, mask = res.features["agent"][0]
    assert np.array_equal(mask.cpu().numpy(), ref_m)
    _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)
    agent_d, client_d = cf.acoustic_dicts(res)
    ref_a = np_acoustic.acoustic_features(ref1[1])
    got = np.asarray(list(client_d[2].values())); exp = np.asarray(list(ref_a.values()))
    assert got[11] == exp[11]                                      # tempo: exact
    assert np.all(np.abs(got - exp) <= np.maximum(1e-3, 1e-4 * np.abs(exp)))     # the 8d gate


def test_resample_batch_one_launch(fe):
    rng = np.random.default_rng(21)
    lens = [1, 2, 5, 1023, 1024, 1026, 2500, 4097, 0, 100001]
    calls = [rng.uniform(-0.5, 0.5, (2, L)).astype(np.float32) for L in lens]
    outs, mixes = fe.resample_batch([_t(c, fe) for c in calls], 8000, 16000, want_mixed=True)
    for c, o, m in zip(calls, outs, mixes):
        ref = np_resample.resample(c, 8000, 16000) if c.shape[1] else np.zeros((2, 0), np.float32)
        assert np.array_equal(o.cpu().numpy(), ref)
        assert np.array_equal(m.cpu().numpy()[0], (ref[0] + ref[1]) * np.float32(0.5))
    mono = [c[:1] for c in calls if c.shape[1]]
    for c, o in zip(mono, fe.resample_batch([_t(c, fe) for c in mono], 8000, 16000)):
        assert np.array_equal(o.cpu().numpy(), np_resample.resample(c, 8000, 16000))


def test_call_frontend_pipelined_equals_single_shot(fe):
    """Host (pinned) inputs in sub-batches over two streams == device inputs in one shot, and both
    equal the per-call drop-in path; covers mono calls and sample_rate == target."""
    import torch
    from audio_processor_b200 import CallFrontend, synth
    calls = [synth.make_call(d, 8000, 300 + i, 2) for i, d in enumerate((30.0, 12.5, 51.0, 0.02, 26.0))]
    calls.append(synth.make_call(27.0, 8000, 399, 1))
    cf = CallFrontend(fe.device, batch_size=4)
    a = cf.process_calls([c.pin_memory() for c in calls], 8000, sub_batch=2)
    b = cf.process_calls([c.to(fe.device) for c in calls], 8000)
    assert a.chunks == b.chunks and a.order == b.order
    for key in ("stereo", "agent", "client"):
        assert len(a.features[key]) == len(b.features[key])
        for (fa, ma), (fb, mb) in zip(a.features[key], b.features[key]):
            assert torch.equal(fa, fb) and torch.equal(ma, mb)
    assert torch.equal(a.acoustic_rows, b.acoustic_rows.cpu()) and torch.equal(a.acoustic_status, b.acoustic_status.cpu())
    # against the oracle: the sorted batches of the mixed stream
    refs = [np_resample.resample(c.numpy(), 8000, 16000) for c in calls]
    mixed = [((r[0] + r[1]) * np.float32(0.5)) if r.shape[0] == 2 else r[0] for r in refs]
    waves = [mixed[ci][s:e] for ci, _, s, e in a.chunks]
    for g, (feat, mask) in enumerate(a.features["stereo"]):
        idx = a.order[g * 4:(g + 1) * 4]
        ref_f, ref_m = np_fbank.featurise([waves[i] for i in idx])
        assert np.array_equal(mask.cpu().numpy(), ref_m)
        _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)
    # 16 kHz input: no resampling, mixed = mean of the rows
    c16 = synth.make_call(3.0, 16000, 5, 2)
    r16 = cf.process_calls([c16], 16000)
    ref_f, ref_m = np_fbank.featurise([((c16[0] + c16[1]) * 0.5).numpy()])
    _cmp_nan(r16.features["stereo"][0][0].cpu().numpy(), ref_f, 1e-3)


def test_call_frontend_pcm16_ingest_is_bit_identical(fe):
    """int16 PCM calls (2 bytes/sample over PCIe, x / 32768 on the device) give exactly the
    tensors of the float32 waveforms torchaudio.load would have produced."""
    import torch
    from audio_processor_b200 import CallFrontend
    g = torch.Generator().manual_seed(5)
    pcm = [torch.randint(-20000, 20000, (2, n), generator=g, dtype=torch.int16) for n in (240000, 100003, 1000)]
    flt = [p.to(torch.float32) / 32768.0 for p in pcm]
    cf = CallFrontend(fe.device, batch_size=4)
    a = cf.process_calls([p.pin_memory() for p in pcm], 8000, sub_batch=2)
    b = cf.process_calls([f.pin_memory() for f in flt], 8000, sub_batch=2)
    assert a.chunks == b.chunks
    for (ya, ma), (yb, mb) in zip(a.waveforms, b.waveforms):
        assert torch.equal(ya, yb) and torch.equal(ma, mb)
    for key in ("stereo", "agent", "client"):
        for (fa, ka), (fb, kb) in zip(a.features[key], b.features[key]):
            assert torch.equal(fa, fb) and torch.equal(ka, kb)
    assert torch.equal(a.acoustic_rows, b.acoustic_rows)
    # 16 kHz PCM: no resampling, plain conversion
    p16 = torch.randint(-20000, 20000, (2, 48000), generator=g, dtype=torch.int16)
    c = cf.process_calls([p16], 16000)
    d = cf.process_calls([p16.to(torch.float32) / 32768.0], 16000)
    assert torch.equal(c.features["stereo"][0][0], d.features["stereo"][0][0])


def test_repeated_runs_are_bit_identical(fe):
    """Race / ordering check in place of compute-sanitizer (closed on this pool): every kernel's reductions are
    fixed trees or in-order sums and no shared word is written by two threads inside a phase, so the same batch --
    ragged segments through fbank and the four acoustic kernels, interleaved with other work on the device --
    must give the same bits every time."""
    import torch
    from audio_processor_b200 import acoustic, synth
    segs = [synth.speech_like(n, 16000, 40 + i).to(fe.device)
            for i, n in enumerate((8000, 16001, 47999, 131072, 400, 5359, 5360, 5361, 400000, 123457))]
    noise = torch.empty(64 << 20, dtype=torch.uint8, device=fe.device)
    ref = None
    for rep in range(6):
        noise.random_(0, 255)                                   # different L2 / scheduling state each time
        feat, mask = fe.fbank(segs)
        rows = acoustic.extract_acoustic_features_batch([s.unsqueeze(0) for s in segs], 16000, fe.device)
        got = (feat.clone(), mask.clone(), [tuple(d.values()) for d in rows])
        torch.cuda.synchronize()
        if ref is None:
            ref = got
            continue
        assert torch.equal(ref[0].view(torch.int32), got[0].view(torch.int32))     # bit patterns (NaN-safe)
        assert torch.equal(ref[1], got[1])
        assert all(np.array_equal(np.asarray(a, np.float64), np.asarray(b, np.float64), equal_nan=True)
                   for a, b in zip(ref[2], got[2]))


def test_two_threads_on_one_frontend_do_not_share_descriptor_blocks(fe):
    """staging_acquire handed a block that thread A had reserved but not yet submitted to
    thread B (its event was still the previous, completed one).  Two threads, each on its own stream, hammer
    fbank / acoustic with DIFFERENT small batches; every result must equal the single-threaded one."""
    import threading
    import torch
    rng = np.random.default_rng(11)
    jobs = []
    for t in range(2):
        lens = [int(n) for n in rng.integers(2_000, 40_000, size=5)]
        waves = [torch.from_numpy(rng.random(n, dtype=np.float32) - 0.5).to(fe.device) for n in lens]
        jobs.append(waves)
    want = []
    for waves in jobs:
        f, m = fe.fbank(waves)
        a, s = fe.acoustic(waves)
        want.append((f.clone(), m.clone(), a.clone(), s.clone()))
    torch.cuda.synchronize(fe.device)
    errors = []

    def worker(t):
        try:
            stream = torch.cuda.Stream(device=fe.device)
            with torch.cuda.stream(stream):
                for _ in range(150):
                    f, m = fe.fbank(jobs[t])
                    a, s = fe.acoustic(jobs[t])
                    stream.synchronize()
                    if not (torch.equal(f, want[t][0]) and torch.equal(m, want[t][1]) and
                            torch.equal(a, want[t][2]) and torch.equal(s, want[t][3])):
                        errors.append(f"thread {t}: result differs from the single-threaded one")
                        return
        except Exception as e:                                   # noqa: BLE001
            errors.append(f"thread {t}: {e!r}")
    threads = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors


def test_fbank_aligned_and_unaligned_tiles_alternate_within_a_block(fe):
    """an unaligned tile is staged with ordinary shared-memory stores, the next aligned one by a
    TMA bulk copy into the same buffer two iterations later (now behind a fence.proxy.async).  300 two-tile
    utterances (so that every block of the launch owns tiles of several utterances) cut from one buffer at
    offsets that alternate between 16-byte aligned and odd -- VAD segment slices look like this -- must give
    exactly what aligned copies of the same samples give, and match the oracle."""
    import torch
    rng = np.random.default_rng(13)
    n_utt, L = 300, 6_640                       # 40 frames = two 32-frame tiles
    base = torch.from_numpy(rng.random(n_utt * (L + 8) + 16, dtype=np.float32) - 0.5).to(fe.device)
    views, copies = [], []
    for i in range(n_utt):
        off = i * (L + 8) + (0 if i % 2 == 0 else (1, 3, 2)[(i // 2) % 3])
        v = base[off:off + L]
        assert (v.data_ptr() % 16 == 0) == (i % 2 == 0)
        views.append(v)
        copies.append(v.clone())
    f1, m1 = fe.fbank(views)
    f2, m2 = fe.fbank(copies)
    assert torch.equal(m1, m2) and torch.equal(f1, f2)
    for i in (0, 1, 2, 151, 299):
        ref_f, _ = np_fbank.featurise([views[i].cpu().numpy()])
        _cmp_nan(f1[i:i + 1].cpu().numpy(), ref_f, 1e-3)


def test_acoustic_shape_rule_is_the_reference_squeeze0(fe, golden):
    """``waveform.squeeze(0)`` + ``unfold(0, ...)`` of the reference accepts (L,) and (1, L); a (2, L) or (1, 1, L)
    tensor makes it raise -> None -> zero features.  Same here, on both entry paths."""
    import torch
    from audio_processor_b200 import acoustic
    w = torch.from_numpy(golden("acoustic")["speech_8000_wave"]).to(fe.device)
    stereo = torch.stack([w, w.flip(0)])
    feats = acoustic.extract_acoustic_features_batch([w, w[None], stereo, w[None, None]], 16000, fe.device)
    assert feats[0] is not None and feats[0] == feats[1] and feats[2] is None and feats[3] is None
    ex = acoustic.AcousticFeatureExtractor(fe.device)
    assert ex._extract_acoustic_features(stereo, 16000) is None
    assert ex.extract_batch([w[None], stereo], 16000)[1] == [0.0] * 38


def test_fbank_pipeline_kernel_is_bit_identical_to_the_serial_one(fe, tmp_path):
    """AFE_FBANK_PIPE=1 selects the three-group tile pipeline (csrc/afe_fbank_pipe.h, 16-frame tiles, named
    barriers); the switch is read once per process, so the pipeline runs in a child process on the same seeded
    batch (ragged, grouped, silence, unaligned views, an utterance spanning several blocks) and must reproduce
    the serial kernel's features and masks bit for bit."""
    import os
    import subprocess
    import sys
    script = tmp_path / "pipe_case.py"
    script.write_text('''
import sys, numpy as np, torch
sys.path.insert(0, sys.argv[1])
from audio_processor_b200 import runtime
fe = runtime.get_frontend(0)
g = torch.Generator().manual_seed(99)
base = (torch.rand(3_000_000, generator=g) - 0.5).to(fe.device)
lens = [400_000, 399, 400, 16_000, 123_457, 720, 48_000, 1_500_001, 30_000, 250_001]
waves, off = [], 0
for i, n in enumerate(lens):
    waves.append(base[off + (i % 3):off + (i % 3) + n]); off += n + 8
waves[4] = torch.zeros(123_457, device=fe.device)
outs = fe.fbank(waves, group_size=4)
f = torch.cat([a.reshape(-1) for a, _ in outs]).cpu().numpy(); m = torch.cat([b.reshape(-1) for _, b in outs]).cpu().numpy()
np.savez(sys.argv[2], f=f, m=m)
''')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = {}
    for mode in ("0", "1"):
        out = tmp_path / f"mode{mode}.npz"
        env = {**os.environ, "AFE_FBANK_PIPE": mode}
        subprocess.run([sys.executable, str(script), root, str(out)], check=True, env=env, timeout=300)
        res[mode] = np.load(out)
    assert np.array_equal(res["0"]["m"], res["1"]["m"])
    assert np.array_equal(res["0"]["f"], res["1"]["f"], equal_nan=True)
    assert np.isfinite(res["1"]["f"][np.repeat(res["1"]["m"], 160) == 1]).all()
