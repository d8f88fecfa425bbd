"""Request-batching shell over the batched decode path (reference: server.py:60-161).

The reference serves `/tts` strictly one request at a time: `prepare_conditioning -> generate(batch_size=1) -> autoencoder.decode ->
numpy clip / * 32767 / astype(int16) -> scipy WAV` inside the Flask handler.  What is accelerated here is everything behind
the HTTP framing:

* `BatchingEngine` - a worker thread that collects concurrently submitted requests into batches (requests whose conditioning
  has the same length and whose sampling parameters are equal share a batch; at most `max_batch`, waiting at most `max_wait_ms`
  for company) and runs ONE `generate` + ONE DAC decode per batch on the batched tcgen05 path.  Batch semantics are those of
  `Zonos.generate(batch_size=B)`: B independent utterances (greedy streams are identical to batch-1 runs; sampled streams draw
  from a Philox stream keyed by the utterance's index in the batch).  This is bucketed static batching: a batch runs to its
  longest utterance; requests that arrive meanwhile wait for the next batch (joining a running batch is not implemented).
* `pcm16` / `wav_bytes` - the handler's float -> 16-bit PCM conversion on the device (`zb_wav_to_pcm16`, bit-exact with the
  numpy expression of server.py:142-147) and the 44-byte RIFF header scipy.io.wavfile.write produces for int16 mono.
* `tts_request` - the handler's payload validation and defaults (server.py:72-134) as a plain function returning
  `(status, body)`, so any HTTP framework can be the shell.  Speaker embeddings and phonemisation stay the caller's
  (SURVEY §8 f4 / host side of f1): they are passed in as callables.
"""
from __future__ import annotations

import queue
import struct
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import Any, Callable

import torch

from . import _lib

DEFAULT_CFG_SCALE = 2.0                  # server.py:18
DEFAULT_MAX_NEW_TOKENS = 86 * 30         # server.py:126
DEFAULT_SAMPLING = dict(top_p=0, top_k=0, min_p=0, linear=0.5, conf=0.4, quad=0.0)     # server.py:129
UNCONDITIONAL_KEYS = ["emotion", "vqscore_8", "fmax", "pitch_std", "dnsmos_ovrl", "speaker_noised"]   # server.py:117


def pcm16(wav: torch.Tensor) -> torch.Tensor:
    """fp32 waveform (any shape, CUDA) -> int16 PCM of the same shape: clip to [-1, 1], * 32767, truncate (server.py:142-147)."""
    if not wav.is_cuda:
        raise _lib.ZbError("pcm16 runs on a CUDA (sm_100) device; there is no CPU fallback")
    w = wav.to(torch.float32).contiguous()
    out = torch.empty(w.shape, dtype=torch.int16, device=w.device)
    with torch.cuda.device(w.device):
        _lib.check(_lib.load().zb_wav_to_pcm16(w.data_ptr(), out.data_ptr(), w.numel(), torch.cuda.current_stream(w.device).cuda_stream))
    return out


def wav_bytes(pcm: torch.Tensor, sample_rate: int) -> bytes:
    """Mono int16 samples -> a WAV file in memory, byte-identical to scipy.io.wavfile.write(buf, sr, int16_array) (server.py:149-152)."""
    data = pcm.detach().reshape(-1).to("cpu", torch.int16).numpy().tobytes()
    hdr = b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE"
    hdr += b"fmt " + struct.pack("<IHHIIHH", 16, 1, 1, sample_rate, sample_rate * 2, 2, 16)
    hdr += b"data" + struct.pack("<I", len(data))
    return hdr + data


@dataclass
class _Request:
    prefix: torch.Tensor                 # [2, Lc, d] conditioning of ONE utterance (row 0 conditional, row 1 unconditional)
    max_new_tokens: int
    cfg_scale: float
    sampling: dict
    future: Future = field(default_factory=Future)
    t_submit: float = field(default_factory=time.perf_counter)

    def key(self):
        return (self.prefix.shape[1], self.max_new_tokens, float(self.cfg_scale), tuple(sorted(self.sampling.items())))


@dataclass
class TtsResult:
    codes: torch.Tensor                  # [9, T] int64 (device)
    pcm: torch.Tensor | None             # [512 * T] int16 (device), None when the engine has no autoencoder
    sample_rate: int
    batch_size: int                      # how many requests shared the batch
    queue_ms: float
    run_ms: float

    def wav(self) -> bytes:
        return wav_bytes(self.pcm, self.sample_rate)


class BatchingEngine:
    """Collects requests into batches for `model.generate(batch_size=B)` + `model.autoencoder.decode`.

    `model` needs the interface of zonos_vibes_b200.Zonos: `generate(prefix[2B, Lc, d], max_new_tokens=, cfg_scale=, batch_size=,
    sampling_params=, progress_bar=False, seed=)`, `last_lengths`, and (optionally) `autoencoder.decode / sampling_rate`.
    One engine per model (a context is not re-entrant): the worker thread is the only caller of the model."""

    def __init__(self, model: Any, max_batch: int = 64, max_wait_ms: float = 5.0, seed: int = 420):
        self.model, self.max_batch, self.max_wait_ms, self.seed = model, int(max_batch), float(max_wait_ms), seed
        self._q: "queue.Queue[_Request | None]" = queue.Queue()
        self._held: list[_Request] = []          # requests taken off the queue that did not fit the batch being formed
        self.batches: list[int] = []             # sizes of the batches run so far (observability / tests)
        self._thread = threading.Thread(target=self._loop, name="zb-batching-engine", daemon=True)
        self._thread.start()

    # ---- client side --------------------------------------------------------------------------------------------
    def submit(self, prefix: torch.Tensor, max_new_tokens: int = DEFAULT_MAX_NEW_TOKENS, cfg_scale: float = DEFAULT_CFG_SCALE,
               sampling_params: dict | None = None) -> Future:
        """prefix: the output of `prepare_conditioning` for one utterance, [2, Lc, d].  Returns a Future of TtsResult."""
        if prefix.dim() != 3 or prefix.shape[0] != 2:
            raise ValueError(f"prefix must be [2, Lc, d] (conditional, unconditional); got {tuple(prefix.shape)}")
        req = _Request(prefix, int(max_new_tokens), float(cfg_scale), dict(DEFAULT_SAMPLING if sampling_params is None else sampling_params))
        self._q.put(req)
        return req.future

    def close(self) -> None:
        self._q.put(None)
        self._thread.join()

    # ---- worker -------------------------------------------------------------------------------------------------
    def _next_batch(self) -> list[_Request] | None:
        """Blocks for the first request, then gathers compatible ones until the batch is full or max_wait_ms has passed."""
        first = self._held.pop(0) if self._held else self._q.get()
        if first is None:
            return None
        batch, key = [first], first.key()
        keep = []
        for r in self._held:                      # earlier leftovers first (arrival order)
            (batch if r.key() == key and len(batch) < self.max_batch else keep).append(r)
        self._held = keep
        deadline = time.perf_counter() + self.max_wait_ms / 1e3
        while len(batch) < self.max_batch:
            left = deadline - time.perf_counter()
            try:
                r = self._q.get(timeout=left) if left > 0 else self._q.get_nowait()
            except queue.Empty:
                break
            if r is None:
                self._q.put(None)                 # shut down after this batch
                break
            (batch if r.key() == key else self._held).append(r)
        return batch

    def _run(self, batch: list[_Request]) -> None:
        t0 = time.perf_counter()
        b = len(batch)
        # generate() wants rows [0, B) conditional and [B, 2B) unconditional (model.py:204-212)
        prefix = torch.cat([r.prefix[0:1] for r in batch] + [r.prefix[1:2] for r in batch], dim=0)
        r0 = batch[0]
        codes = self.model.generate(prefix, max_new_tokens=r0.max_new_tokens, cfg_scale=r0.cfg_scale, batch_size=b,
                                    sampling_params=dict(r0.sampling), progress_bar=False, seed=self.seed)
        lengths = list(getattr(self.model, "last_lengths", None) or [codes.shape[-1]] * b)
        auto = getattr(self.model, "autoencoder", None)
        pcm_all = pcm16(auto.decode(codes)) if auto is not None else None          # [B, 1, 512 * T] int16
        if codes.is_cuda:
            torch.cuda.synchronize(codes.device)
        run_ms = (time.perf_counter() - t0) * 1e3
        self.batches.append(b)
        for i, r in enumerate(batch):
            n = int(lengths[i])
            res = TtsResult(codes[i, :, :n], pcm_all[i, 0, : 512 * n] if pcm_all is not None else None,
                            int(getattr(auto, "sampling_rate", 44100)), b, (t0 - r.t_submit) * 1e3, run_ms)
            r.future.set_result(res)

    def _loop(self) -> None:
        while True:
            batch = self._next_batch()
            if batch is None:
                return
            try:
                self._run(batch)
            except BaseException as e:  # noqa: BLE001 - the error belongs to the requests of this batch
                for r in batch:
                    if not r.future.done():
                        r.future.set_exception(e)


def tts_request(engine: BatchingEngine, data: Any, speaker_embedding: Callable[[str], torch.Tensor],
                default_language: str = "en-us", timeout_s: float | None = None) -> tuple[int, bytes | dict]:
    """The `/tts` handler of server.py:60-161 without the web framework: `data` is the decoded JSON payload
    (`text`, `speaker_audio_path`, `speaking_rate`, optional `language`); returns `(http_status, wav_bytes | {"error": ...})` with
    the reference's validation order and messages.  `speaker_embedding(path)` is the caller's (cached) speaker-embedding
    function (server.py:38-57); FileNotFoundError from it maps to 404 like the reference."""
    from .conditioning import make_cond_dict
    if not isinstance(data, dict):
        return 400, {"error": "Request must be JSON"}
    text, path, rate = data.get("text"), data.get("speaker_audio_path"), data.get("speaking_rate")
    language = data.get("language", default_language)
    if not text:
        return 400, {"error": "Missing 'text' parameter"}
    if not path:
        return 400, {"error": "Missing 'speaker_audio_path' parameter"}
    if rate is None:
        return 400, {"error": "Missing 'speaking_rate' parameter"}
    try:
        rate = float(rate)
    except (TypeError, ValueError):
        return 400, {"error": "'speaking_rate' must be a number"}
    try:
        model = engine.model
        cond = make_cond_dict(text=text, language=language, speaker=speaker_embedding(path), speaking_rate=rate,
                              device=model.device, unconditional_keys=UNCONDITIONAL_KEYS)
        prefix = model.prepare_conditioning(cond)
        res = engine.submit(prefix, DEFAULT_MAX_NEW_TOKENS, DEFAULT_CFG_SCALE, DEFAULT_SAMPLING).result(timeout=timeout_s)
        return 200, res.wav()
    except FileNotFoundError as e:
        return 404, {"error": str(e)}
    except Exception as e:  # noqa: BLE001 - server.py:167-171
        return 500, {"error": f"An internal server error occurred: {str(e)}"}
