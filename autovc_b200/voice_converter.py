"""Batched conversion around the generator: the mel-level part of the reference's `VoiceConverter.convert` /
`convert_multiple` (reference `autovc/voice_converter.py:88-211`, `:346-398`), re-designed for throughput:

* the reference converts one (source, target) pair per `forward` call in a Python loop (`:391-392`, B = 1) and re-embeds
  the target for every pair unless it is a cached key (`:150-155`); here every speaker is embedded ONCE and cached like
  `SpeakerEncoder.speakers` (`speaker_encoder/model.py:101,446-460`), utterances are bucketed by frame count and each
  bucket is one batched `AutoEncoder.forward`;
* utterances are never padded across different lengths - the reference has no length masking, so padding would change
  the BiLSTM result;
* under `torch.distributed` every rank converts a contiguous block of the work list with replicated weights and no
  collective on the data path; an optional final tensor all-gather (NCCL over NVLink, one per length bucket) returns
  the complete list on every rank.

`VoiceConverter` works on mels: inputs are normalised mels `(80, T)`, outputs are `mel_outputs_postnet` `(80, T)` exactly
as `convert` hands them to the vocoder (`:170,:175`).  `WavVoiceConverter` is the waveform-level `convert`: mel front-end
(`audio.py`), d-vector encoder, generator and WaveRNN vocoder chained on the device.  File I/O,
the `Audio` pre/out-processing pipelines and wandb logging stay outside.
"""
from __future__ import annotations

from dataclasses import dataclass
from itertools import product
from typing import Callable, Dict, Hashable, Iterable, List, Optional, Sequence, Tuple, Union

import torch

from .model import overlap_average

SpeakerRef = Union[Hashable, torch.Tensor]


@dataclass
class Utterance:
    """A source (or target) utterance: its mel `(80, T)` and the key of its speaker."""
    mel: torch.Tensor
    speaker: Hashable
    speaker_data: object = None      # whatever the embedder needs to embed this speaker (e.g. d-vector partials)


def shard_range(n: int, rank: int, world: int) -> range:
    """Contiguous block of ceil(n/world) items per rank (SURVEY.md 8e 'contiguous blocks of ceil(B/G) rows per GPU')."""
    per = (n + world - 1) // world
    return range(min(n, rank * per), min(n, (rank + 1) * per))


def bucket_by_length(lengths: Sequence[int]) -> Dict[int, List[int]]:
    """frame count -> indices, in first-seen order; equal-length utterances batch together, nothing is padded."""
    buckets: Dict[int, List[int]] = {}
    for i, t in enumerate(lengths):
        buckets.setdefault(int(t), []).append(i)
    return buckets


def chunk_mel(mel: torch.Tensor, n_frames: int, overlap: float = 0.5) -> torch.Tensor:
    """Cut `(80, T)` into `(M, 80, n_frames)` windows with hop int(n_frames*(1-overlap)) - the wire format
    `batch_forward` expects (reference `model.py:395-437`; the reference's slicer is `spectrogram.py:248-311`).
    Trailing frames that do not fill a whole window are left out (the reference pads there, outside this path)."""
    T = mel.shape[-1]
    if T < n_frames:
        raise ValueError(f"utterance has {T} frames, fewer than one chunk of {n_frames}")
    hop = int(n_frames * (1 - overlap))
    starts = list(range(0, T - n_frames + 1, hop))
    return torch.stack([mel[:, s:s + n_frames] for s in starts])


def host_rows(rows: Sequence[torch.Tensor]) -> torch.Tensor:
    """Equal-shaped CPU tensors -> one pinned float32 tensor (len, *shape) for an asynchronous host->device copy.
    Rows that already are consecutive slices of one pinned float32 tensor are returned as a view of it (no copy);
    anything else is stacked into pinned staging memory (torch's caching host allocator recycles it)."""
    first = rows[0]
    n, shape = len(rows), tuple(first.shape)
    if first.dtype == torch.float32 and first.is_contiguous() and first.is_pinned():
        step = first.numel() * 4
        base, store = first.data_ptr(), first.untyped_storage().data_ptr()
        if all(r.dtype == torch.float32 and tuple(r.shape) == shape and r.is_contiguous() and r.data_ptr() == base + j * step
               and r.untyped_storage().data_ptr() == store for j, r in enumerate(rows)):
            strides = (first.numel(),) + tuple(first.stride())
            return torch.as_strided(first, (n,) + shape, strides)
    staged = torch.empty((n,) + shape, dtype=torch.float32, pin_memory=torch.cuda.is_available())
    torch.stack([r.to(torch.float32) for r in rows], out=staged)
    return staged


class VoiceConverter:
    """Mel-level drop-in for the conversion loop of the reference `VoiceConverter`.

    `auto_encoder`  : an `autovc_b200.AutoEncoder` (or any callable with the same `forward` contract - tests inject the oracle)
    `embed_speaker` : callable(speaker_data) -> (dim_emb,) tensor, the d-vector encoder; called once per speaker key
    `max_batch`     : utterances per `forward` (config 2 uses 256 per GPU)
    """

    def __init__(self, auto_encoder, embed_speaker: Optional[Callable[[object], torch.Tensor]] = None,
                 device: Optional[torch.device] = None, max_batch: int = 256):
        self.AE = auto_encoder
        self.embed_speaker = embed_speaker
        self.device = torch.device(device) if device is not None else getattr(auto_encoder, "device", torch.device("cpu"))
        self.max_batch = int(max_batch)
        self.speakers: Dict[Hashable, torch.Tensor] = {}
        self.embed_calls = 0
        self._pipe: Optional["HostPipeline"] = None

    # -- speakers ----------------------------------------------------------------------------------
    def learn_speaker(self, key: Hashable, embedding: torch.Tensor) -> None:
        """Register a (mean) speaker embedding under a key, like `SpeakerEncoder.learn_speaker` fills `.speakers`."""
        self.speakers[key] = embedding.detach().to(torch.float32).reshape(-1).cpu()

    def learn_speakers(self, keys: Sequence[Hashable], datas: Sequence[object], embed_many: Optional[Callable] = None) -> int:
        """Embed every not-yet-known speaker of `keys` ONCE, all of them in one batched call of the d-vector encoder
        (`embed_many(list of speaker data) -> (n, dim_emb)`, e.g. `SpeakerEncoder.embed_speakers`; default: the
        converter's per-speaker `embed_speaker`), and cache them like `SpeakerEncoder.speakers` (reference
        `speaker_encoder/model.py:101,446-460`).  Returns the number of speakers embedded."""
        todo, seen = [], set()
        for k, d in zip(keys, datas):
            if k not in self.speakers and k not in seen and d is not None:
                seen.add(k); todo.append((k, d))
        if not todo:
            return 0
        if embed_many is not None:
            embeds = embed_many([d for _, d in todo])
            self.embed_calls += len(todo)
            for (k, _), e in zip(todo, embeds.detach().to(torch.float32).cpu()):
                self.speakers[k] = e.reshape(-1)
        else:
            for k, d in todo:
                self.embedding(k, d)
        return len(todo)

    def embedding(self, ref: SpeakerRef, data: object = None) -> torch.Tensor:
        """Embedding of a speaker: a tensor is used as is, a known key hits the cache (reference `:150-151`), an
        unknown key is embedded once from `data` and cached."""
        if isinstance(ref, torch.Tensor):
            return ref.detach().to(torch.float32).reshape(-1).cpu()
        if ref in self.speakers:
            return self.speakers[ref]
        if self.embed_speaker is None or data is None:
            raise KeyError(f"unknown speaker {ref!r} and nothing to embed it from")
        self.embed_calls += 1
        self.learn_speaker(ref, self.embed_speaker(data))
        return self.speakers[ref]

    # -- core: batched forward over a work list ----------------------------------------------------
    def _plan_batches(self, mels: Sequence[torch.Tensor]) -> List[List[int]]:
        """Index lists, one per forward: (length bucket) x (chunks of max_batch), in first-seen order."""
        plan = []
        for _T, idx in bucket_by_length([m.shape[-1] for m in mels]).items():
            for s in range(0, len(idx), self.max_batch):
                plan.append(idx[s:s + self.max_batch])
        return plan

    def convert_batch(self, mels: Sequence[torch.Tensor], c_src: Sequence[torch.Tensor], c_trg: Sequence[torch.Tensor],
                      output: str = "device", host_buffers: Optional[Dict[Tuple[int, ...], torch.Tensor]] = None):
        """[(80,T_i)], [(E,)], [(E,)] -> [mel_outputs_postnet (80,T_i)] in input order.  One forward per
        (length bucket, chunk of max_batch).

        Host-resident mels on a CUDA converter go through `HostPipeline`: batch i+1 is staged in pinned memory and copied
        in while batch i computes and batch i-1's result is copied out.  Mels that are consecutive rows of one pinned
        tensor (e.g. `Utterance(mel=slab[i])`) are copied straight from it, without staging.
        `output`: "device" - results stay on the converter's device (stream-ordered, like the reference's tensors);
        "host" - results are rows of pinned CPU tensors, complete when the call returns.
        `host_buffers` (with "host"): {mel shape: pinned float32 tensor (rows, *shape)} owned by the caller and reused from
        call to call - a serving loop should pass them: a fresh pinned block of this size costs a cudaHostAlloc (~100 ms for
        the 168 MB of 4096 utterances) whenever torch's host allocator has none cached.  Results of one shape fill the
        buffer's rows in input order."""
        if output not in ("device", "host"):
            raise ValueError("output must be 'device' or 'host'")
        out: List[Optional[torch.Tensor]] = [None] * len(mels)
        plan = self._plan_batches(mels)
        if not plan:
            return out
        pipelined = self.device.type == "cuda" and all(not m.is_cuda for m in mels)
        if not pipelined:
            for sel in plan:
                x = torch.stack([mels[i] for i in sel]).to(self.device, torch.float32)
                co = torch.stack([c_src[i] for i in sel]).to(self.device, torch.float32)
                ct = torch.stack([c_trg[i] for i in sel]).to(self.device, torch.float32)
                _, post, _ = self.AE(x, co, ct)
                if output == "host":
                    post = post.cpu()
                for j, i in enumerate(sel):
                    out[i] = post[j]
            return out

        if self._pipe is None:
            self._pipe = HostPipeline(self.AE, self.device)

        # Speaker embeddings: one table row per DISTINCT tensor object (the cache hands out one tensor per speaker), gathered
        # per batch with one index_select into pinned staging instead of stacking 2 x max_batch small tensors.
        table_rows: Dict[int, int] = {}
        table: List[torch.Tensor] = []
        def row_of(t: torch.Tensor) -> int:
            r = table_rows.get(id(t))
            if r is None:
                r = table_rows[id(t)] = len(table)
                table.append(t.detach().to(torch.float32).reshape(-1))
            return r
        src_rows = [row_of(t) for t in c_src]
        trg_rows = [row_of(t) for t in c_trg]
        emb = torch.stack(table)
        pin = torch.cuda.is_available()

        def gather(rows: List[int]) -> torch.Tensor:
            out_ = torch.empty((len(rows), emb.shape[1]), dtype=torch.float32, pin_memory=pin)
            torch.index_select(emb, 0, torch.tensor(rows, dtype=torch.int64), out=out_)
            return out_

        class _Staged:      # batch i is staged when the pipeline enqueues its copy, i.e. while batch i-1 computes
            def __len__(_s):
                return len(plan)

            def __getitem__(_s, i):
                sel = plan[i]
                return (host_rows([mels[k] for k in sel]), gather([src_rows[k] for k in sel]), gather([trg_rows[k] for k in sel]))

        if output == "host":
            # one pinned result tensor per length bucket; the per-utterance views are cut (one unbind per batch) while the
            # pipeline is running, not after it
            slabs: Dict[Tuple[int, ...], torch.Tensor] = {}
            fill: Dict[Tuple[int, ...], int] = {}
            for sel in plan:
                shp = tuple(mels[sel[0]].shape)
                fill[shp] = fill.get(shp, 0) + len(sel)
            for shp, n_rows in fill.items():
                buf = (host_buffers or {}).get(shp)
                if buf is not None and (buf.shape[0] < n_rows or tuple(buf.shape[1:]) != shp or buf.dtype != torch.float32 or not buf.is_contiguous()):
                    raise ValueError(f"host buffer for shape {shp} must be a contiguous float32 tensor with at least {n_rows} rows")
                slabs[shp] = buf if buf is not None else torch.empty((n_rows,) + shp, dtype=torch.float32, pin_memory=pin)
                fill[shp] = 0
            outs = []
            for sel in plan:
                shp = tuple(mels[sel[0]].shape)
                part = slabs[shp][fill[shp]:fill[shp] + len(sel)]
                fill[shp] += len(sel)
                outs.append((None, part, None))
            staged = _Staged()

            class _Outs:      # indexing batch i's outputs (the pipeline does so right after launching batch i) also cuts its views
                def __getitem__(_s, i):
                    for k, v in zip(plan[i], outs[i][1].unbind(0)):
                        out[k] = v
                    return outs[i]

            self._pipe.run(staged, _Outs())
            self._pipe.d2h_stream.synchronize()
        else:
            def keep(bi, res):
                for k, v in zip(plan[bi], res[1].unbind(0)):
                    out[k] = v
            self._pipe.run(_Staged(), None, on_result=keep)
        return out

    # -- reference-shaped API -----------------------------------------------------------------------
    def convert(self, source: Utterance, target: SpeakerRef, target_data: object = None, cut: bool = False,
                n_frames: int = 400, overlap: float = 0.5) -> torch.Tensor:
        """One conversion (reference `convert`, `:88-211`, mel in / mel out).  With `cut` the utterance is chunked and
        overlap-averaged through `batch_forward` semantics (`:163-169`), otherwise it is one sequence (`:172-174`).
        The source embedding is the SOURCE speaker's (the reference hard-codes `speakers['hilde']`, `:147`, with the
        intended call commented out at `:145`)."""
        c_s = self.embedding(source.speaker, source.speaker_data)
        c_t = self.embedding(target, target_data)
        if not cut:
            return self.convert_batch([source.mel], [c_s], [c_t])[0]
        chunks = chunk_mel(source.mel, n_frames, overlap)
        posts = self.convert_batch(list(chunks), [c_s] * len(chunks), [c_t] * len(chunks))
        return overlap_average(torch.stack(posts), overlap)

    def match(self, sources: Sequence[Utterance], targets: Sequence[Union[Utterance, Hashable]],
              match_method: str = "all_combinations") -> List[Tuple[int, int]]:
        """(source index, target index) pairs in the reference's order (`:383-388`)."""
        if match_method == "align":
            if len(sources) != len(targets):
                raise AssertionError("align needs as many targets as sources")
            return list(zip(range(len(sources)), range(len(targets))))
        if match_method == "all_combinations":
            return list(product(range(len(sources)), range(len(targets))))
        raise ValueError(f"unknown match_method {match_method!r}")

    def convert_multiple(self, sources: Sequence[Utterance], targets: Sequence[Union[Utterance, Hashable]],
                         match_method: str = "all_combinations", bidirectional: bool = False,
                         distributed: bool = False, gather: bool = True, output: str = "device",
                         host_buffers: Optional[Dict[Tuple[int, ...], torch.Tensor]] = None) -> List[Optional[torch.Tensor]]:
        """reference `convert_multiple` (`:346-398`) as batched work: same pairing and output order, but speakers are
        embedded once each, pairs are bucketed by length, and with `distributed` the pair list is split in contiguous
        blocks over the ranks (results of other ranks are `None` unless `gather`).  `output` / `host_buffers` as in
        `convert_batch`."""
        pairs: List[Tuple[Utterance, SpeakerRef, object]] = []
        for si, ti in self.match(sources, targets, match_method):
            tgt = targets[ti]
            if isinstance(tgt, Utterance):
                pairs.append((sources[si], tgt.speaker, tgt.speaker_data))
            else:
                if bidirectional:
                    raise AssertionError("Cannot convert in both ways if a mean speaker embedding is to be used")
                pairs.append((sources[si], tgt, None))
        if bidirectional:   # targets become sources (`:394-396`)
            for ti, si in self.match(targets, sources, match_method):
                pairs.append((targets[ti], sources[si].speaker, sources[si].speaker_data))

        rank, world = 0, 1
        if distributed:
            import torch.distributed as dist
            rank, world = dist.get_rank(), dist.get_world_size()
        mine = shard_range(len(pairs), rank, world)
        # one embedding per distinct speaker this rank touches (a 256-float vector; cheaper to recompute than to exchange)
        mels = [pairs[i][0].mel for i in mine]
        c_s = [self.embedding(pairs[i][0].speaker, pairs[i][0].speaker_data) for i in mine]
        c_t = [self.embedding(pairs[i][1], pairs[i][2]) for i in mine]
        local = self.convert_batch(mels, c_s, c_t, output=output, host_buffers=host_buffers)
        results: List[Optional[torch.Tensor]] = [None] * len(pairs)
        for i, r in zip(mine, local):
            results[i] = r
        if distributed and gather and world > 1:
            self._gather(results, [p[0].mel.shape for p in pairs], local, world, output)
        return results

    def _gather(self, results: List[Optional[torch.Tensor]], shapes: Sequence[Tuple[int, ...]], local: Sequence[torch.Tensor],
                world: int, output: str) -> None:
        """Optional final output gather (off the hot path): one `all_gather_into_tensor` per length bucket
        over the job's backend - NCCL over NVLink for CUDA tensors.  Every rank knows the whole pair list, so the
        per-rank row counts of each bucket are computed locally and no size exchange is needed; blocks are padded to the
        largest rank's count (the contiguous sharding leaves them within one row of each other)."""
        import torch.distributed as dist
        n = len(shapes)
        on_cuda = dist.get_backend() == "nccl"
        dev = self.device if on_cuda else torch.device("cpu")
        by_rank = [list(shard_range(n, rk, world)) for rk in range(world)]
        me = dist.get_rank()
        local_of = {i: r for i, r in zip(by_rank[me], local)}
        for shape in sorted({tuple(sh) for sh in shapes}):          # same bucket order on every rank
            rows = [[i for i in blk if tuple(shapes[i]) == shape] for blk in by_rank]
            cap = max(len(r) for r in rows)
            if cap == 0:
                continue
            send = torch.zeros((cap,) + shape, dtype=torch.float32, device=dev)
            if rows[me]:
                send[:len(rows[me])] = torch.stack([local_of[i].to(dev) for i in rows[me]])
            recv = torch.empty((world * cap,) + shape, dtype=torch.float32, device=dev)     # rank blocks concatenated along dim 0
            dist.all_gather_into_tensor(recv, send)
            if output == "host" and recv.is_cuda:
                recv = recv.cpu()
            recv = recv.view((world, cap) + shape)
            for rk, idx in enumerate(rows):
                if rk == me:
                    continue
                for j, i in enumerate(idx):
                    results[i] = recv[rk, j]


class WavVoiceConverter:
    """Waveform-level conversion: reference `VoiceConverter.convert` (`voice_converter.py:88-211`) from the point where the
    wavs are loaded (`:140-155`) to the vocoder output (`:170,:175`), every stage on the device:

        target wav -> mel_spec_speaker_encoder(cut) -> SpeakerEncoder partials -> mean + renorm      (`:153`, SE `:136-190`)
        source wav -> mel_spec_auto_encoder (-> chunks when `cut`)                                    (`:159`)
        AutoEncoder.forward / batch_forward                                                           (`:163-174`)
        WaveRNN.generate                                                                              (`:170,:175`)

    `convert_multiple` does that for a work list with the batching of `VoiceConverter` (one embedding per distinct wav,
    one generator call per length bucket, one vocoder loop over the folds of all results)."""

    def __init__(self, auto_encoder, speaker_encoder, vocoder=None, device: Optional[torch.device] = None, max_batch: int = 256):
        self.AE, self.SE, self.vocoder = auto_encoder, speaker_encoder, vocoder
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        self.mel_level = VoiceConverter(auto_encoder, device=self.device, max_batch=max_batch)

    def _embed(self, speakers: Sequence) -> List[torch.Tensor]:
        """speaker refs (a key of `SE.speakers`, an embedding tensor, or a wav) -> embeddings; all wavs in one SE batch."""
        out: List[Optional[torch.Tensor]] = [None] * len(speakers)
        wav_idx = []
        for i, sp in enumerate(speakers):
            if isinstance(sp, torch.Tensor) and sp.dim() == 1 and sp.numel() == self.AE.dims[1]:
                out[i] = sp.detach().to(torch.float32).cpu()
            elif isinstance(sp, (str, int)) and sp in getattr(self.SE, "speakers", {}):
                out[i] = self.SE.speakers[sp]                                      # reference `:150-151`
            else:
                wav_idx.append(i)
        if wav_idx:
            embeds = self.SE.embed_utterances([speakers[i] for i in wav_idx]).cpu()
            for j, i in enumerate(wav_idx):
                out[i] = embeds[j]
        return out

    def convert_mels(self, source_wavs: Sequence, source_speakers: Sequence, targets: Sequence, cut: bool = False,
                     overlap: float = 0.5, **mel_kwargs) -> List[torch.Tensor]:
        """[source wav], [source speaker ref], [target speaker ref] -> [mel_outputs_postnet (80, T_i)] on the device."""
        from . import audio
        n = len(source_wavs)
        embeds = self._embed(list(source_speakers) + list(targets))
        c_s, c_t = embeds[:n], embeds[n:]
        if not cut:
            mels = [audio.mel_spec_auto_encoder(w, device=self.device, **mel_kwargs) for w in source_wavs]
            return self.mel_level.convert_batch(mels, c_s, c_t)
        # chunked: every chunk of every utterance is a row of the same batched forward; overlap-average per utterance
        chunks, owner = [], []
        for i, w in enumerate(source_wavs):
            for c in audio.mel_spec_auto_encoder(w, cut=True, device=self.device, overlap=overlap, **mel_kwargs):
                chunks.append(c); owner.append(i)
        posts = self.mel_level.convert_batch(chunks, [c_s[i] for i in owner], [c_t[i] for i in owner])
        return [overlap_average(torch.stack([p for p, o in zip(posts, owner) if o == i]), overlap) for i in range(n)]

    def convert(self, source_wav, target, source_speaker=None, cut: bool = False, overlap: float = 0.5, vocode: bool = True,
                vocoder_args: Optional[dict] = None, **mel_kwargs):
        """One conversion.  `source_speaker` defaults to the source wav itself (the intended call the reference has
        commented out at `:145`).  Returns the waveform (numpy float32) or, with `vocode=False`, the converted mel."""
        post = self.convert_mels([source_wav], [source_wav if source_speaker is None else source_speaker], [target],
                                 cut=cut, overlap=overlap, **mel_kwargs)[0]
        if not vocode:
            return post
        if self.vocoder is None:
            raise RuntimeError("no vocoder was given to WavVoiceConverter")
        return self.vocoder.generate(post.unsqueeze(0), **(vocoder_args or {}))

    def convert_multiple(self, source_wavs: Sequence, target_wavs: Sequence, match_method: str = "all_combinations",
                         cut: bool = False, overlap: float = 0.5, vocode: bool = True, vocoder_args: Optional[dict] = None,
                         **mel_kwargs) -> list:
        """reference `convert_multiple` (`:346-398`) on waveforms, pairing and order as the reference; each distinct wav
        is embedded once, and the vocoder runs one loop over the folds of all converted utterances of a length."""
        pairs = self.mel_level.match(source_wavs, target_wavs, match_method)
        distinct = list(source_wavs) + list(target_wavs)
        embeds = self._embed(distinct)
        ns = len(source_wavs)
        posts = self.convert_mels([source_wavs[si] for si, _ in pairs], [embeds[si] for si, _ in pairs],
                                  [embeds[ns + ti] for _, ti in pairs], cut=cut, overlap=overlap, **mel_kwargs)
        if not vocode:
            return posts
        if self.vocoder is None:
            raise RuntimeError("no vocoder was given to WavVoiceConverter")
        waves: list = [None] * len(posts)
        for _, idx in bucket_by_length([p.shape[-1] for p in posts]).items():   # one vocoder loop per length bucket
            for i, w in zip(idx, self.vocoder.generate_batch([posts[i] for i in idx], **(vocoder_args or {}))):
                waves[i] = w
        return waves


class HostPipeline:
    """Throughput path for host-resident batches: step i+1's host->device copy and step i-1's device->host copy run on
    two copy streams (one per direction: the B200's copy engines work independently) while step i computes, with
    double-buffered device inputs/outputs and pinned host outputs.  The generator call itself is the unchanged public
    `AutoEncoder.forward`.

    One pipeline drives one `AutoEncoder` from one host thread; the engine behind the module owns one workspace, so two
    pipelines (or two streams) must not run the same module concurrently."""

    def __init__(self, auto_encoder, device: torch.device):
        self.AE = auto_encoder
        self.device = torch.device(device)
        self.h2d_stream = torch.cuda.Stream(self.device)
        self.d2h_stream = torch.cuda.Stream(self.device)
        self._dev_in = [None, None]
        self._dev_out = [None, None]
        import inspect
        try:
            self._supports_out = "out" in inspect.signature(auto_encoder.forward).parameters
        except (TypeError, ValueError):
            self._supports_out = False

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]],
            outputs: Optional[Sequence[Tuple[Optional[torch.Tensor], ...]]] = None,
            on_result: Optional[Callable[[int, Tuple[torch.Tensor, ...]], None]] = None):
        """`batches`: pinned host (x, c_org, c_trg); `outputs`: per batch pinned host (mel, mel_post, codes) to fill - an
        entry that is None is not copied back.  `on_result(i, (mel, mel_post, codes))` (device tensors, enqueued on the
        compute stream) is for callers that keep results on the device: the step then gets fresh output tensors instead
        of the pipeline's reusable slots.  Returns after everything is enqueued; the caller synchronises."""
        compute = torch.cuda.current_stream(self.device)
        if not hasattr(batches, "__getitem__"):
            batches = list(batches)     # a sequence is indexed lazily: batch i is first touched when its copy is enqueued
        n = len(batches)
        if n == 0:
            return
        dev_in, dev_out = self._dev_in, self._dev_out      # device slots persist across runs (stream order makes reuse safe)
        in_ready = [torch.cuda.Event(), torch.cuda.Event()]
        out_done = [torch.cuda.Event(), torch.cuda.Event()]
        comp_done = [torch.cuda.Event(), torch.cuda.Event()]
        reuse_out = on_result is None and self._supports_out

        def h2d(i):
            with torch.cuda.stream(self.h2d_stream):
                if i >= 2:
                    self.h2d_stream.wait_event(comp_done[i % 2])     # the slot's previous inputs have been consumed
                slot, host = dev_in[i % 2], batches[i]
                if slot is not None and all(d.shape == t.shape for d, t in zip(slot, host)):
                    for d, t in zip(slot, host):                       # steady state: reuse the slot's device buffers
                        d.copy_(t, non_blocking=True)
                else:
                    dev_in[i % 2] = tuple(t.to(self.device, non_blocking=True) for t in host)
                    for t in dev_in[i % 2]:
                        t.record_stream(compute)
                in_ready[i % 2].record(self.h2d_stream)

        h2d(0)
        for i in range(n):
            if i + 1 < n:
                h2d(i + 1)
            compute.wait_event(in_ready[i % 2])
            slot_out = dev_out[i % 2]
            if reuse_out and slot_out is not None and slot_out[0].shape == dev_in[i % 2][0].shape:
                if i >= 2:
                    compute.wait_event(out_done[i % 2])                # the slot's previous results have left the device
                res = self.AE(*dev_in[i % 2], out=slot_out)           # steady state: no allocation on any stream
            else:
                res = tuple(self.AE(*dev_in[i % 2]))
                if reuse_out:
                    dev_out[i % 2] = res
            comp_done[i % 2].record(compute)
            if on_result is not None:
                on_result(i, res)
            if outputs is not None:
                with torch.cuda.stream(self.d2h_stream):
                    self.d2h_stream.wait_event(comp_done[i % 2])
                    for dst, src in zip(outputs[i], res):
                        if dst is not None:
                            dst.copy_(src, non_blocking=True)
                    out_done[i % 2].record(self.d2h_stream)
                for t in res:
                    t.record_stream(self.d2h_stream)
        if outputs is not None:
            compute.wait_event(out_done[(n - 1) % 2])
            if n > 1:
                compute.wait_event(out_done[(n - 2) % 2])
