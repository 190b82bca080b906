"""Drop-in for the hot-path methods of the reference's ``AudioProcessor``
(services/audio_processor.py): ``get_resampler`` (:99-102), ``split_audio`` (:104-160),
the length bucketing of ``transcribe_batch`` (:170-181) and the featurise step of
``_run_inference`` (:253-266).  Model loading, decoding and the database are out of scope
 and stay with the reference.

Output placement: waveforms and chunk views live on the CUDA device (they feed the fbank
and acoustic kernels directly, without the reference's ``.numpy()`` hop).  Everything
integer -- chunk boundaries, ``start_time`` / ``end_time`` floats, ``chunk_idx`` -- is
bit-identical to the reference.
"""
from __future__ import annotations

import ctypes as C
import logging
from functools import lru_cache
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

from . import _native, runtime
from .feature_extractor import FbankFeatureExtractor, FbankProcessor


_CLEAN = str.maketrans("", "", ",?.!-;:\"%'\u00bb\u00ab\u061f()\u060c")


def remove_special_characters(text) -> str:
    """Lower-cases and drops punctuation (Latin and Arabic) like the reference's helper of the same name
    (utils/utils.py:74-79); ``None`` -> ""."""
    return "" if text is None else text.lower().translate(_CLEAN)


_EMPTY = {"transcription_chunk": "", "agent_transcription": "", "client_transcription": ""}


def chunk_table(total_samples: int, chunk_samples: int, step_samples: int) -> List[Tuple[int, int]]:
    """(start, end) sample pairs of ``split_audio``'s while-loop, from the C ABI."""
    lib = _native.load()
    n = C.c_int(0)
    lib.check(lib.afe_chunk_table(int(total_samples), int(chunk_samples), int(step_samples), None, None, 0, C.byref(n)))
    starts = (C.c_int64 * max(1, n.value))()
    ends = (C.c_int64 * max(1, n.value))()
    lib.check(lib.afe_chunk_table(int(total_samples), int(chunk_samples), int(step_samples), starts, ends, n.value,
                                  C.byref(n)))
    return [(int(starts[i]), int(ends[i])) for i in range(n.value)]


class Resampler:
    """Callable with the behaviour of ``torchaudio.transforms.Resample(orig, new)``."""

    def __init__(self, orig_freq: int, new_freq: int, device=None):
        self.orig_freq, self.new_freq = int(orig_freq), int(new_freq)
        self._device = device

    def __call__(self, waveform: torch.Tensor) -> torch.Tensor:
        if self.orig_freq == self.new_freq:
            return waveform
        fe = runtime.get_frontend(self._device)
        return fe.resample(waveform.to(fe.device, non_blocking=True), self.orig_freq, self.new_freq)

    forward = __call__


class AudioProcessor:
    """``AudioProcessor(config)`` with the reference's config keys
    (``chunk_duration_sec``, ``overlap_sec``, ``target_sample_rate``, ``batch_size`` /
    ``chunk_batch_size``, ``enable_length_bucketing``, ``gpu_index``)."""

    def __init__(self, config: dict, db_manager=None, processor=None, model=None, frontend=None):
        self.config = config
        gpu_index = int(config.get("gpu_index", 0))
        self.device = torch.device(f"cuda:{gpu_index}")
        self._frontend = frontend
        self.db_manager = db_manager
        self.chunk_duration_sec = config["chunk_duration_sec"]
        self.overlap_sec = config.get("overlap_sec", 1)
        self.target_sample_rate = config.get("target_sample_rate", 16000)
        self.processor = processor or FbankProcessor(FbankFeatureExtractor(device=self.device))
        self.model = model                     # callable(input_features=...) -> object with .logits (Wav2Vec2BertForCTC)
        self.text_cleaner = remove_special_characters
        self.logger = logging.getLogger(__name__)

    @property
    def frontend(self) -> "runtime.Frontend":
        if self._frontend is None:
            self._frontend = runtime.get_frontend(self.device)
        return self._frontend

    # -------------------------------------------------------------- reference :72-93
    def load_audio(self, audio_path):
        """``(waveform float32 (C, L), sample_rate)`` like ``torchaudio.load``; the file's bytes go to the
        device as they are and are decoded there (wav.py).  ``.ogg`` goes through ffmpeg as in the
        reference.  Errors propagate (the reference logs and re-raises)."""
        from . import wav
        if str(audio_path).lower().endswith(".ogg"):
            return wav.load_audio(wav.ffmpeg_wav_bytes(audio_path), self.device)
        return wav.load_audio(audio_path, self.device)

    def load_pcm(self, audio_path):
        """The fast route for 16-bit wav files: ``(pcm, sample_rate)`` where ``pcm`` is an int16 (C, L) view
        of the frames in pinned memory; ``split_audio`` / ``CallFrontend.process_calls`` take it in place
        of the float waveform (2 bytes per sample over PCIe, decode fused into the resampler)."""
        from . import wav
        if str(audio_path).lower().endswith(".ogg"):
            return wav.load_pcm16(wav.ffmpeg_wav_bytes(audio_path))
        return wav.load_pcm16(audio_path)

    # -------------------------------------------------------------- reference :99-102
    @lru_cache(maxsize=100)
    def get_resampler(self, orig_freq: int, new_freq: int) -> Resampler:
        return Resampler(orig_freq, new_freq, self.device)

    # -------------------------------------------------------------- reference :104-160
    def split_audio(self, waveform: torch.Tensor, sample_rate: int, file_name: str):
        fe = self.frontend
        waveform = waveform.to(fe.device, non_blocking=True)
        mixed = None
        if waveform.dtype == torch.int16 and sample_rate != self.target_sample_rate:
            # 16-bit PCM from load_pcm: decode fused into the (batched, here one-call) resampler
            if waveform.shape[0] == 2:
                (waveform,), (mixed,) = fe.resample_batch([waveform], sample_rate, self.target_sample_rate, want_mixed=True)
            else:
                (waveform,) = fe.resample_batch([waveform], sample_rate, self.target_sample_rate)
            sample_rate = self.target_sample_rate
        elif waveform.dtype == torch.int16:
            waveform = fe.to_float(waveform)
        if sample_rate != self.target_sample_rate:
            if waveform.shape[0] == 2:
                waveform, mixed = fe.resample(waveform, sample_rate, self.target_sample_rate, want_mixed=True)
            else:
                waveform = fe.resample(waveform, sample_rate, self.target_sample_rate)
            sample_rate = self.target_sample_rate
        # integer contract of the reference (:115-118): int() truncation of seconds x rate, step = window - overlap
        window = int(self.chunk_duration_sec * sample_rate)
        hop = window - int(self.overlap_sec * sample_rate)
        if waveform.shape[0] == 2:
            agent, client = waveform[0:1], waveform[1:2]            # views of the resampled call
            both = mixed if mixed is not None else fe.mix_stereo(waveform)
        else:
            agent = client = both = waveform                          # mono: the three streams alias one tensor
        name = f"{file_name}"
        chunks: List[Dict[str, Any]] = [
            {"file_name": name, "stereo_waveform": both[:, lo:hi], "agent_waveform": agent[:, lo:hi],
             "client_waveform": client[:, lo:hi], "chunk_idx": n, "start_time": lo / sample_rate, "end_time": hi / sample_rate}
            for n, (lo, hi) in enumerate(chunk_table(waveform.shape[1], window, hop))]
        return chunks, agent, client

    # -------------------------------------------------------------- reference :170-181
    def bucket_chunks(self, chunks: Sequence[Dict]) -> List[List[Dict]]:
        """The stable length sort and ``batch_size`` slicing of ``transcribe_batch``."""
        batch_size = int(self.config.get("batch_size", self.config.get("chunk_batch_size", 16)))
        chunks = list(chunks)
        if bool(self.config.get("enable_length_bucketing", True)):
            chunks = sorted(chunks, key=lambda c: int(c.get("stereo_waveform", c.get("agent_waveform")).shape[1]))
        return [chunks[i:i + batch_size] for i in range(0, len(chunks), batch_size)]

    # -------------------------------------------------------------- reference :253-266
    def featurise_batch(self, batch_chunks: Sequence[Dict]) -> Dict[str, torch.Tensor]:
        """The three ``processor(...)`` calls of ``_run_inference`` for one batch of chunks:
        ``{"stereo"|"agent"|"client": input_features (B, T/2, 160)}`` (+ ``*_mask``)."""
        out: Dict[str, torch.Tensor] = {}
        for key, field in (("stereo", "stereo_waveform"), ("agent", "agent_waveform"), ("client", "client_waveform")):
            arrays = [c[field].squeeze(0) for c in batch_chunks]
            inputs = self.processor(arrays, sampling_rate=16000, return_tensors="pt", padding=True)
            out[key] = inputs["input_features"]
            out[key + "_mask"] = inputs["attention_mask"]
        return out

    def featurise_all(self, chunks: Sequence[Dict]):
        """All batches of ``transcribe_batch`` in three launches (one per stream): a list of
        ``(batch_chunks, {"stereo": (feat, mask), "agent": .., "client": ..})``."""
        batches = self.bucket_chunks(chunks)
        if not batches:
            return []
        flat = [c for b in batches for c in b]
        gs = int(self.config.get("batch_size", self.config.get("chunk_batch_size", 16)))
        fx = self.processor.feature_extractor
        per_stream = {}
        for key, field in (("stereo", "stereo_waveform"), ("agent", "agent_waveform"), ("client", "client_waveform")):
            per_stream[key] = self.frontend.fbank([c[field].squeeze(0) for c in flat], padding_value=fx.padding_value,
                                                  precision=fx.precision, group_size=gs)
        return [(b, {k: per_stream[k][i] for k in per_stream}) for i, b in enumerate(batches)]

    # -------------------------------------------------------------- reference :162-212
    def transcribe_batch(self, chunks: List[Dict]) -> List[Dict]:
        """Length-bucketed batches of ``batch_size`` chunks through ``_process_single_batch``, with the
        reference's error contract: an exception whose text contains "out of memory" halves the batch size
        and retries the same slice; any other marks the slice's chunks with empty transcriptions + ``error``.

        Batch assembly stays on the device: the features of the slices ahead are produced by ONE fbank launch
        per stream (``group_size = batch_size``: every slice padded to its own longest chunk, exactly as one
        ``processor(...)`` call per slice would) and handed to the model as views -- no numpy hop."""
        if not chunks:
            return []
        size = int(self.config["batch_size"])
        done: List[Dict] = []
        if bool(self.config.get("enable_length_bucketing", True)):
            try:
                chunks = sorted(chunks, key=lambda c: int(c.get("stereo_waveform", c.get("agent_waveform")).shape[1]))
            except Exception:
                pass
        ahead_n = max(1, int(self.config.get("featurise_ahead_batches", 64)))
        ahead = None                  # (first chunk, batch size, [features per slice])
        at = 0
        while at < len(chunks):
            stop = min(at + size, len(chunks))
            members = chunks[at:stop]
            try:
                if ahead is None or ahead[1] != size or at >= ahead[0] + size * len(ahead[2]):
                    ahead = (at, size, self._featurise_ahead(chunks[at:at + ahead_n * size], size))
                done.extend(self._process_single_batch(members, ahead[2][(at - ahead[0]) // size]))
                self._cleanup_memory()
            except Exception as e:
                self.logger.error(f"slice {at // size + 1} (chunks {at}..{stop - 1}) failed: {e}")
                if "out of memory" in str(e).lower() and size > 1:
                    self.logger.warning(f"device memory exhausted: batch size {size} -> {max(1, size // 2)}, same slice again")
                    size, ahead = max(1, size // 2), None
                    self._release_cache()
                    continue                                   # same `at`, smaller slice
                for chunk in members:
                    chunk.update({**_EMPTY, "error": str(e)})
                done.extend(members)
            at = stop
        return done

    def _featurise_ahead(self, chunks: Sequence[Dict], batch_size: int):
        """Features of consecutive slices of ``batch_size`` chunks: a list of ``{"stereo"|"agent"|"client": input_features}``
        per slice, or a list of ``None`` when the processor is not the CUDA one (then each slice calls it itself)."""
        n_slices = (len(chunks) + batch_size - 1) // batch_size
        fx = getattr(self.processor, "feature_extractor", None)
        if not isinstance(fx, FbankFeatureExtractor):
            return [None] * n_slices
        per_stream = {}
        for key, field in (("stereo", "stereo_waveform"), ("agent", "agent_waveform"), ("client", "client_waveform")):
            per_stream[key] = self.frontend.fbank([c[field].squeeze(0) for c in chunks], padding_value=fx.padding_value,
                                                  precision=fx.precision, group_size=batch_size)
        return [{k: per_stream[k][g][0] for k in per_stream} for g in range(n_slices)]

    # -------------------------------------------------------------- reference :214-245
    def _process_single_batch(self, batch_chunks: List[Dict], features: Optional[Dict[str, torch.Tensor]] = None) -> List[Dict]:
        agent = [chunk["agent_waveform"].squeeze(0) for chunk in batch_chunks]
        client = [chunk["client_waveform"].squeeze(0) for chunk in batch_chunks]
        stereo = [chunk["stereo_waveform"].squeeze(0) for chunk in batch_chunks]
        if torch.cuda.is_available():
            with torch.autocast("cuda"):
                results = self._run_inference(stereo, agent, client, features)
        else:
            results = self._run_inference(stereo, agent, client, features)
        for i, chunk in enumerate(batch_chunks):
            if i < len(results):
                chunk.update(results[i])
            else:
                chunk.update({**_EMPTY, "error": "Index out of range"})
        return batch_chunks

    # -------------------------------------------------------------- reference :247-309
    def _run_inference(self, stereo_arrays, agent_arrays, client_arrays,
                       features: Optional[Dict[str, torch.Tensor]] = None) -> List[Dict]:
        """Three ``processor(...)`` calls (skipped when ``features`` were produced ahead), the CTC model on each
        ``input_features`` -- WITHOUT the attention mask, as the reference (:270-272) --, argmax, ``batch_decode``,
        ``remove_special_characters``.  Any exception becomes a list of empty results carrying its text."""
        try:
            if features is None:
                features = {}
                for key, arrays in (("stereo", stereo_arrays), ("agent", agent_arrays), ("client", client_arrays)):
                    inputs = self.processor(arrays, sampling_rate=16000, return_tensors="pt", padding=True)
                    feat = inputs["input_features"]
                    features[key] = feat.to(self.device) if torch.cuda.is_available() else feat
            with torch.no_grad():
                logits = {k: self.model(input_features=features[k]).logits for k in ("stereo", "agent", "client")}
            texts = {k: self.processor.batch_decode(torch.argmax(v, dim=-1)) for k, v in logits.items()}
            cleaned = {k: [self.text_cleaner(t) for t in v] for k, v in texts.items()}
            return [{"transcription_chunk": cleaned["stereo"][i],
                     "agent_transcription": cleaned["agent"][i] if i < len(cleaned["agent"]) else "",
                     "client_transcription": cleaned["client"][i] if i < len(cleaned["client"]) else "",
                     "error": ""} for i in range(len(cleaned["stereo"]))]
        except Exception as e:
            self.logger.error(f"Inference error: {e}")
            return [{**_EMPTY, "error": str(e)}] * len(stereo_arrays)

    # -------------------------------------------------------------- reference :311-322
    def _cleanup_memory(self):
        """The reference collects garbage and empties the CUDA cache after EVERY batch; ``cudaFree`` is a device
        synchronisation, so here the caching allocator keeps its blocks unless ``release_cache_every_batch``."""
        if bool(self.config.get("release_cache_every_batch", False)):
            self._release_cache()

    @staticmethod
    def _release_cache():
        if torch.cuda.is_available():
            torch.cuda.empty_cache()
