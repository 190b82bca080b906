"""BASELINE.json configs cfg2 / cfg3 / cfg5 on the CUDA path: oracle parity at sizes the oracle finishes
in seconds, and size-independent properties at the full sizes.  Needs a B200 (``pytest -m gpu``)."""
import numpy as np
import pytest

from oracle import np_acoustic, np_fbank, np_resample, np_tables

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from audio_processor_b200 import runtime
    return runtime.get_frontend(0)


def _cmp_nan(a, b, tol):
    assert a.shape == b.shape
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = ~np.isnan(b)
    err = float(np.max(np.abs(a[ok] - b[ok]), initial=0.0))
    assert err <= tol, err


def _acoustic_close(got, ref, rel=1e-4, floor=1e-3):
    got = np.asarray(got, np.float64); ref = np.asarray(ref, np.float64)
    assert got[11] == ref[11]                                     # tempo: exact
    bad = [(k, got[k], ref[k]) for k in range(38) if abs(got[k] - ref[k]) > max(floor, rel * abs(ref[k]))]
    assert not bad, bad


# ------------------------------------------------------------------ cfg3: VAD-style ragged segments
def test_cfg3_ragged_segments_fbank_and_acoustic(fe):
    """0.5-30 s segments incl. the forced edge lengths, featurised in length-sorted batches of 16 exactly like
    SpeechSegment.transcribe_segments_batched (speech_segment.py:199-232), and their 38 acoustic features."""
    from audio_processor_b200 import SpeechSegment, acoustic, synth
    segs, sr = synth.cfg3_segments(37)
    lens = [int(s.shape[1]) for s in segs]
    assert lens[:5] == list(synth.CFG3_EDGE_LENGTHS)
    seg = SpeechSegment({"gpu_index": 0, "batch_size": 16})
    items = [{"segment_waveform": s.to(fe.device), "start": 0.0, "end": n / sr, "speaker": "client"} for s, n in zip(segs, lens)]
    groups = seg.featurise_segments(items, sr)
    order = sorted(range(len(lens)), key=lens.__getitem__)
    assert [i for idxs, _, _ in groups for i in idxs] == order     # stable sort by length, batches of 16
    assert [len(idxs) for idxs, _, _ in groups] == [16, 16, 5]
    for idxs, feat, mask in groups:
        ref_f, ref_m = np_fbank.featurise([segs[i][0].numpy() for i in idxs])
        assert tuple(feat.shape) == ref_f.shape
        assert np.array_equal(mask.cpu().numpy(), ref_m)            # bit-exact masks / padded lengths
        _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)
    # acoustic features of the edge lengths and a few random ones
    pick = [0, 1, 2, 3, 4, 9, 17, 30]
    feats = acoustic.extract_acoustic_features_batch([segs[i][0].numpy() for i in pick], sr, fe.device)
    for i, f in zip(pick, feats):
        ref = np_acoustic.acoustic_features(segs[i][0].numpy())
        assert list(f.keys()) == list(ref.keys())
        _acoustic_close(list(f.values()), list(ref.values()))


def test_cfg3_full_size_4096_segments(fe):
    """cfg3 at BASELINE size: 4 096 mono segments of 0.5-30 s (the lengths of synth.cfg3_segments, seed 7, incl. the
    forced edge lengths 8 000 / 8 001 / 16 000 / 479 999 / 480 000).  Properties on everything -- batch order, padded
    lengths, masks, finiteness, tempo's closed form -- and oracle parity on the first, a middle and the last batch of
    16 and on eight sampled acoustic rows.  The waveforms are speech-like for the sampled ones and seeded device noise
    for the rest (synthesising 17 hours of speech-like audio on one host core would take minutes)."""
    import torch
    from audio_processor_b200 import SpeechSegment, synth
    from audio_processor_b200.acoustic import TEMPO_SLOT, _tempo
    n_seg, sr = 4096, 16000
    u = torch.rand(n_seg, generator=torch.Generator().manual_seed(7), dtype=torch.float64)
    lens = list(synth.CFG3_EDGE_LENGTHS) + [int(sr * (0.5 + 29.5 * float(v))) for v in u][:n_seg - len(synth.CFG3_EDGE_LENGTHS)]
    order = sorted(range(n_seg), key=lens.__getitem__)
    sampled_groups = (0, 128, 255)
    sampled = {i for g in sampled_groups for i in order[16 * g:16 * g + 16]}
    gen = torch.Generator(device=fe.device).manual_seed(77)
    segs = []
    for i, n in enumerate(lens):
        if i in sampled:
            segs.append(synth.speech_like(n, sr, 1000 + i).to(fe.device).unsqueeze(0))
        else:
            segs.append(((torch.rand(n, generator=gen, device=fe.device) - 0.5) * 0.6).unsqueeze(0))
    seg = SpeechSegment({"gpu_index": 0, "batch_size": 16})
    items = [{"segment_waveform": s, "start": 0.0, "end": n / sr, "speaker": "client"} for s, n in zip(segs, lens)]
    groups = seg.featurise_segments(items, sr)
    assert len(groups) == 256 and [i for idxs, _, _ in groups for i in idxs] == order
    for g, (idxs, feat, mask) in enumerate(groups):
        T = [1 + (lens[i] - 400) // 160 for i in idxs]
        assert tuple(feat.shape) == (16, (max(T) + 1) // 2, 160)
        assert mask.sum(1).tolist() == [t // 2 for t in T]
        assert torch.isfinite(feat).all()
    for g in sampled_groups:
        idxs, feat, mask = groups[g]
        ref_f, ref_m = np_fbank.featurise([segs[i][0].cpu().numpy() for i in idxs])
        assert np.array_equal(mask.cpu().numpy(), ref_m)
        _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)
    rows, status = fe.acoustic([s[0] for s in segs], sr)
    assert int(status.sum()) == 0 and torch.isfinite(rows).all()
    host = rows.cpu()
    for i in (0, 1, 4, 777, 4095):                                  # tempo: the reference's closed form, exact
        assert _tempo(float(host[i, TEMPO_SLOT]), lens[i], sr) == 60.0 / (((1 + lens[i] // 160) - 2) * 160 / sr)
    for i in sorted(sampled)[::6]:
        ref = np_acoustic.acoustic_features(segs[i][0].cpu().numpy())
        vals = host[i].tolist()
        vals[TEMPO_SLOT] = _tempo(vals[TEMPO_SLOT], lens[i], sr)
        _acoustic_close(vals, list(ref.values()))


# ------------------------------------------------------------------ cfg5: 44.1 kHz long form
def test_cfg5_44k_resample_chunks_and_long_fbank(fe):
    import torch
    from audio_processor_b200 import AudioProcessor, synth
    # (a) generic 441 -> 160 polyphase resampler, bit-exact against the oracle on 12 s
    x, sr = synth.cfg5_longform(12.0)
    assert sr == 44100
    y = fe.resample(x.to(fe.device), sr, 16000)
    ref = np_resample.resample(x.numpy(), sr, 16000)
    assert y.shape[1] == int(np.ceil(x.shape[1] * 16000 / 44100)) == ref.shape[1]
    assert np.array_equal(y.cpu().numpy(), ref)
    # (b) the integer contract at the full 1-hour size (host only): 150 chunks, 149 full + 384000
    L16 = fe.resampled_length(158_760_000, 44100, 16000)
    assert L16 == 57_600_000
    table = np_tables.chunk_table(L16, 16000, 25, 1)
    ap = AudioProcessor({"chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000, "batch_size": 16, "gpu_index": 0})
    from audio_processor_b200.audio_processor import chunk_table
    mine = chunk_table(L16, 400_000, 384_000)
    assert [(s, e) for s, e, _, _ in table] == [(s, e) for s, e in mine]
    assert len(mine) == 150 and mine[-1] == (57_216_000, 57_600_000) and all(e - s == 400_000 for s, e in mine[:-1])
    # (c) one unchunked processor([x]) call on a long waveform: 10 min = 59 998 frames against the oracle,
    #     and a mono call goes through split_audio with all three streams aliasing one tensor
    x10, _ = synth.cfg5_longform(600.0 * 16000 / 44100)            # 600 s worth of 16 kHz samples at 44.1 kHz pace
    y10 = fe.resample(x10.to(fe.device), 44100, 16000)[0]
    T = 1 + (y10.numel() - 400) // 160
    feat, mask = fe.fbank([y10])
    assert tuple(feat.shape) == (1, (T + 1) // 2, 160) and int(mask.sum()) == T // 2
    ref_f, ref_m = np_fbank.featurise([y10.cpu().numpy()])
    assert np.array_equal(mask.cpu().numpy(), ref_m)
    _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)
    chunks, agent, client = ap.split_audio(x, 44100, "long.wav")
    assert agent.data_ptr() == client.data_ptr() and len(chunks) == 1
    assert torch.equal(chunks[0]["stereo_waveform"], chunks[0]["agent_waveform"])


def test_cfg5_long_sequence_properties(fe):
    """The 1-hour stress shape: 359 998 frames in ONE utterance (per-utterance statistics span 11 250 tiles
    and every block of the launch).  Properties: frame count, mask, zero mean / unit variance per mel bin."""
    import torch
    n = 57_600_000
    g = torch.Generator(device=fe.device).manual_seed(9)
    y = torch.rand(n, generator=g, device=fe.device) - 0.5
    y[1_000_000:1_500_000] = 0.0                                    # a stretch of digital silence
    feat, mask = fe.fbank([y])
    T = 1 + (n - 400) // 160
    assert T == 359_998 and tuple(feat.shape) == (1, T // 2, 160) and int(mask.sum()) == T // 2
    f = feat.view(-1, 80)[:T].double()
    assert torch.isfinite(f).all()
    assert float(f.mean(0).abs().max()) < 1e-4
    assert float((f.var(0, unbiased=True) - 1.0).abs().max()) < 1e-3


# ------------------------------------------------------------------ cfg2 at full size: properties + spot parity
def test_cfg2_full_size_properties_and_spot_parity(fe):
    import torch
    from audio_processor_b200 import CallFrontend, synth
    calls, sr = synth.cfg2_calls(64, 300.0, distinct=3)
    cf = CallFrontend(fe.device)
    res = cf.process_calls([c.pin_memory() for c in calls], sr)
    n = len(res.chunks)
    assert n == 64 * 13 and abs(res.call_seconds - 64 * 300.0) < 1e-6
    # chunk table of every call: 12 x 400000 + 192000, starts 384000 apart (bit-exact integers)
    per_call = [(s, e) for c, _, s, e in res.chunks if c == 5]
    assert per_call == [(384000 * i, min(384000 * i + 400000, 4_800_000)) for i in range(13)]
    # length-sorted batches of 16: 4 batches of the 64 short chunks, then full chunks
    lens = [e - s for _, _, s, e in res.chunks]
    assert res.order == sorted(range(n), key=lens.__getitem__)
    for key in ("stereo", "agent", "client"):
        views = res.features[key]
        assert len(views) == 52
        for g, (feat, mask) in enumerate(views):
            exp_T = 1 + (max(lens[i] for i in res.order[16 * g:16 * g + 16]) - 400) // 160
            assert tuple(feat.shape) == (16, (exp_T + 1) // 2, 160)
            assert mask.sum(1).tolist() == [(1 + (lens[i] - 400) // 160) // 2 for i in res.order[16 * g:16 * g + 16]]
        assert all(torch.isfinite(f).all() for f, _ in views)
    # identical calls give identical outputs wherever they land (calls 0, 3, 6, ... are the same waveform)
    a = res.order.index(0 * 13 + 2); b = res.order.index(3 * 13 + 2)
    fa = res.features["client"][a // 16][0][a % 16]; fb = res.features["client"][b // 16][0][b % 16]
    assert torch.equal(fa, fb)
    assert torch.equal(res.acoustic_rows[0 * 13 + 2], res.acoustic_rows[3 * 13 + 2])
    assert int(res.acoustic_status.sum()) == 0
    # oracle parity on EVERY chunk of one call: 13 chunks x 3 streams (fbank) and both speakers (acoustic)
    ref = np_resample.resample(calls[1].numpy(), sr, 16000)
    assert np.array_equal(res.waveforms[1][0].cpu().numpy(), ref)
    ref_mixed = (ref[0] + ref[1]) * np.float32(0.5)
    assert np.array_equal(res.waveforms[1][1].cpu().numpy()[0], ref_mixed)
    streams = {"stereo": ref_mixed, "agent": ref[0], "client": ref[1]}
    from audio_processor_b200.acoustic import TEMPO_SLOT, _tempo
    worst = 0.0
    for chunk_idx in range(13):
        j = 1 * 13 + chunk_idx
        pos = res.order.index(j)
        s, e = res.chunks[j][2], res.chunks[j][3]
        T = 1 + (e - s - 400) // 160
        for key, wave in streams.items():
            feat, mask = res.features[key][pos // 16]
            Tp = feat.shape[1] * 2
            ref_f, _ = np_fbank.featurise([wave[s:e]])
            got = feat[pos % 16].reshape(Tp, 80)[:T - (T % 2)].cpu().numpy()
            err = float(np.max(np.abs(got - ref_f.reshape(-1, 80)[:T - (T % 2)])))
            worst = max(worst, err)
            assert err <= 1e-3, (chunk_idx, key, err)
            assert int(mask[pos % 16].sum()) == T // 2
        for speaker, wave in ((0, ref[0]), (1, ref[1])):
            ref_a = np_acoustic.acoustic_features(wave[s:e])
            vals = res.acoustic_rows[speaker * n + j].tolist()
            vals[TEMPO_SLOT] = _tempo(vals[TEMPO_SLOT], e - s, 16000)
            _acoustic_close(vals, list(ref_a.values()))
    print(f"cfg2 call 1: 39 chunk-streams, worst fbank error {worst:.2e}")


def test_cfg5_full_chunk_at_the_hour_offset(fe):
    """cfg5's 1-hour call has 150 chunks; chunk 149 starts 57 216 000 samples in.  A waveform of that length is
    built on the device, chunk 148 (full) and 149 (the 384 000-sample tail) are featurised through split_audio's
    views at their true offsets (beyond 2^25 samples) and checked against the oracle: fbank, mask and the 38
    acoustic features."""
    import torch
    from audio_processor_b200 import AudioProcessor
    n = 57_600_000
    g = torch.Generator(device=fe.device).manual_seed(15)
    y = (torch.rand(n, generator=g, device=fe.device) - 0.5).unsqueeze(0)
    ap = AudioProcessor({"chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000, "batch_size": 16, "gpu_index": 0})
    chunks, agent, client = ap.split_audio(y, 16000, "hour.wav")
    assert len(chunks) == 150 and agent.data_ptr() == client.data_ptr() == y.data_ptr()
    assert (chunks[149]["start_time"], chunks[149]["end_time"]) == (3576.0, 3600.0)
    for idx in (148, 149):
        w = chunks[idx]["agent_waveform"]
        assert w.data_ptr() == y.data_ptr() + 4 * 384_000 * idx
        feat, mask = fe.fbank([w[0]])
        ref_f, ref_m = np_fbank.featurise([w[0].cpu().numpy()])
        assert np.array_equal(mask.cpu().numpy(), ref_m)
        _cmp_nan(feat.cpu().numpy(), ref_f, 1e-3)
        a, st = fe.acoustic([w[0]])
        from audio_processor_b200.acoustic import TEMPO_SLOT, _tempo
        vals = a[0].tolist()
        vals[TEMPO_SLOT] = _tempo(vals[TEMPO_SLOT], w.shape[1], 16000)
        _acoustic_close(vals, list(np_acoustic.acoustic_features(w[0].cpu().numpy()).values()))


# ------------------------------------------------------------------ threading contract
def test_concurrent_threads_share_one_processor(fe):
    """The reference calls split_audio from 8-32 io_executor threads and process_file_batch from up to 16
    threads on ONE AudioProcessor (audio_processor.py:46-47, 334; pipeline.py:397).  Eight threads, each on its
    own CUDA stream, run the whole per-file path concurrently; every result must equal the serial one bit for bit."""
    import threading
    import torch
    from audio_processor_b200 import AudioProcessor, acoustic, synth
    ap = AudioProcessor({"chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000, "batch_size": 16, "gpu_index": 0})
    calls = [synth.make_call(40.0 + 7.0 * i, 8000, 500 + i, 2) for i in range(8)]

    def one(call, name):
        chunks, agent, client = ap.split_audio(call, 8000, name)
        feats = ap.featurise_batch(chunks)
        rows = acoustic.extract_acoustic_features_batch([c["agent_waveform"] for c in chunks], 16000, fe.device)
        return agent.clone(), {k: v.clone() for k, v in feats.items()}, rows

    serial = [one(c, f"s{i}") for i, c in enumerate(calls)]
    torch.cuda.synchronize()
    out, errs = [None] * len(calls), []

    def work(i):
        try:
            with torch.cuda.stream(torch.cuda.Stream(fe.device)):
                for _ in range(3):                                    # a few rounds to interleave the threads
                    out[i] = one(calls[i], f"t{i}")
                torch.cuda.current_stream().synchronize()
        except BaseException as e:                                    # noqa: BLE001
            errs.append(e)
    threads = [threading.Thread(target=work, args=(i,)) for i in range(len(calls))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errs, errs
    torch.cuda.synchronize()
    for (a0, f0, r0), (a1, f1, r1) in zip(serial, out):
        assert torch.equal(a0, a1)
        assert f0.keys() == f1.keys() and all(torch.equal(f0[k], f1[k]) for k in f0)
        assert [list(d.values()) for d in r0] == [list(d.values()) for d in r1]


# ------------------------------------------------------------------ cfg4: calls sharded by file over the visible GPUs
def test_cfg4_queue_shards_files_over_all_visible_gpus(fe, tmp_path):
    """cfg4 at test size: wav files pulled from the host queue by one worker per visible GPU
    (multi_gpu_manager.py:67-139 is the reference's static deal; here any GPU count, dynamic pull).
    Whatever device a file lands on, its outputs equal the single-device ones bit for bit; `.done`
    markers make a second run a no-op.  With one visible GPU the queue has one worker."""
    import torch
    from audio_processor_b200 import CallFrontend, synth, wav, work_queue
    n_dev = torch.cuda.device_count()
    files = []
    for i in range(3 * max(n_dev, 2)):
        call = synth.make_call(26.0 + 3.0 * (i % 4), 8000, 900 + i, 2)
        p = tmp_path / f"call_{i:02d}.wav"
        p.write_bytes(synth.wav_image(call, 8000).numpy().tobytes())
        files.append(p)
    fronts = {d: CallFrontend(torch.device("cuda", d)) for d in range(n_dev)}
    used = []

    def one(dev, path):
        pcm, sr = wav.load_pcm16(wav.read_pinned(path))
        r = fronts[dev].process_calls([pcm], sr)
        torch.cuda.synchronize(dev)
        used.append(dev)
        feats = {k: [(f.cpu(), m.cpu()) for f, m in v] for k, v in r.features.items()}
        return r.chunks, feats, r.acoustic_rows.clone(), r.waveforms[0][0].cpu()

    serial = [one(0, p) for p in files]
    used.clear()
    markers = work_queue.ProcessedMarkers(tmp_path / "markers")
    out = work_queue.MultiGpuQueue().map(one, files, markers=markers, max_retries=1)
    assert sorted(set(used)) == list(range(n_dev)) or len(files) < n_dev
    for (c0, f0, a0, w0), got in zip(serial, out):
        assert got is not None
        c1, f1, a1, w1 = got
        assert c0 == c1 and torch.equal(w0, w1) and torch.equal(a0, a1)
        assert f0.keys() == f1.keys()
        for k in f0:
            assert all(torch.equal(x0, x1) and torch.equal(m0, m1) for (x0, m0), (x1, m1) in zip(f0[k], f1[k]))
    assert all(markers.is_already_processed(p) for p in files)
    assert work_queue.MultiGpuQueue().map(one, files, markers=markers) == [None] * len(files)   # resumed run: nothing left
