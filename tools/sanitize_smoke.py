"""Small end-to-end pass for compute-sanitizer (memcheck / racecheck / initcheck):
    compute-sanitizer --tool racecheck python tools/sanitize_smoke.py     (closed on the round-1 pool; also a plain smoke)
Every kernel family runs once on small ragged inputs: wav decode, 1:2 stereo resample (float and 16-bit),
44.1 k -> 16 k polyphase, fbank (fused-tail and split-utterance normalisation), the four acoustic kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import CallFrontend, acoustic, runtime, synth, wav

fe = runtime.get_frontend(0)
dev = fe.device
calls = [synth.make_call(27.0 + 1.3 * i, 8000, 10 + i, 2) for i in range(3)]
cf = CallFrontend(dev)
r = cf.process_calls([c.pin_memory() for c in calls], 8000)
pcm = [wav.load_pcm16(synth.wav_image(c, 8000).pin_memory())[0] for c in calls]
r2 = cf.process_calls(pcm, 8000)
x44, _ = synth.cfg5_longform(3.0)
y = fe.resample(x44.to(dev), 44100, 16000)
segs = [synth.speech_like(n, 16000, 3 + n).to(dev) for n in (8000, 8001, 16000, 47999, 400, 399)]
feat, mask = fe.fbank(segs[:5])
rows = acoustic.extract_acoustic_features_batch([s.unsqueeze(0) for s in segs], 16000, dev)
w, sr = wav.load_audio(synth.wav_image(calls[0], 8000))
torch.cuda.synchronize()
print("sanitize smoke ok", float(r.acoustic_rows.sum()), float(feat.float().nan_to_num().sum()), len(rows), tuple(w.shape), tuple(y.shape))
