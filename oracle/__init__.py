"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's waveform front end.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import anything from this package, and only as the checker
or the timed CPU baseline.  The product package ``audio_processor_b200`` never imports it
and fails loudly when its CUDA library is missing.

Parity pinning: the reference ships no tests or golden vectors and
its arithmetic lives in third-party wheels (torchaudio / transformers / torch / numpy).
The oracle is therefore pinned against outputs of the UNMODIFIED reference call sites run
in the build container (``oracle/make_golden.py`` -> ``tests/golden/*.npz``), installed
versions torch 2.11.0, torchaudio 2.11.0, transformers 5.5.0, numpy 2.3.5 (the reference
pins 2.6.0 / 2.6.0 / 4.48.0 / 1.26.4, which cannot be fetched offline).

Modules
  np_resample   torchaudio sinc_interp_hann resampler (taps in float64, in-order FMA apply)
  np_fbank      SeamlessM4T / Kaldi fbank + per-utterance normalisation + pad/stack/mask
  np_acoustic   the 38 acoustic features of AcousticSentimentAnalyzer
  np_tables     chunk table, VAD-segment table, interval merge (integer / float64 contract)
  libref        the reference's call sites restated over the same third-party libraries
                (the timed CPU baseline; cannot import /root/reference on the GPU box)
  reference_calls  container-only import of the unmodified reference (fixture generation)
"""
