"""Drop-in for the inference side of the reference d-vector `SpeakerEncoder`
(reference `autovc/speaker_encoder/model.py:21-58`): the producer of the `c_org` / `c_trg` rows the generator consumes.

Same constructor defaults (`utils/hparams.py:60-63`: 40 mel channels, hidden 256, embedding 256, 3 layers,
batch_first), same state-dict key names (`lstm.weight_ih_l0 ... linear.bias`), same `forward(utterances)` contract -
but `forward` is ONE call into the sm_100a CUDA path (`torch.ops.autovc_b200.speaker_embed` -> C ABI
`avc_se_forward`): the generator's tensor-core GEMM for the hoisted input projections and the persistent recurrence
kernel for the three LSTM layers.  There is no PyTorch/CPU fallback.

The boundary is mel-level, as everywhere in this package: wav loading and the 40-channel filterbank
(`autovc/audio/spectrogram.py`, librosa) stay on the caller's side.  `embed_partials` is `embed_utterance` from the
point where the reference has `frames_batch` in hand (`model.py:181-186`); `embed_speakers` does that for many
utterances in one launch sequence, which the reference leaves as a TODO (`model.py:141`).

Training (GE2E loss, `model.py:188-370`) is out of scope.
"""
from __future__ import annotations

from typing import Dict, Hashable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

from ._weights import PackedWeightsMixin
from .engine import Engine


class SpeakerEncoder(PackedWeightsMixin, nn.Module):
    def __init__(self, input_size: int = 40, hidden_size: int = 256, embedding_size: int = 256, num_layers: int = 3,
                 batch_first: bool = True, device=None, verbose: bool = True, precision: str = "bf16"):
        super().__init__()
        if (input_size, hidden_size, embedding_size, num_layers, batch_first) != (40, 256, 256, 3, True):
            raise ValueError("the CUDA path is built for the reference configuration: LSTM(40, 256, 3, batch_first) + Linear(256, 256)")
        if device is None:
            device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.verbose = verbose
        self.device = device if not isinstance(device, str) else torch.device(device)
        self.precision = precision
        self.speakers: Dict[Hashable, torch.Tensor] = {}
        self.logging = {}
        # parameter containers under the reference's names; the forward is the fused CUDA op
        self.lstm = nn.LSTM(input_size=input_size, hidden_size=hidden_size, num_layers=num_layers, batch_first=batch_first)
        self.linear = nn.Linear(in_features=hidden_size, out_features=embedding_size)
        self.train(False)
        self._init_tracking()
        self._engines: Dict[Tuple[int, str], Engine] = {}
        self._engine_keys: Dict[Tuple[int, str], object] = {}

    # -- weights -----------------------------------------------------------------------------------
    def _engine(self, device: torch.device) -> Engine:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        key = (idx, self.precision)
        eng = self._engines.get(key)
        if eng is None:
            eng = Engine((32, 256, 512, 32), idx, self.precision)   # generator dims are unused by the speaker encoder
            self._engines[key] = eng
            self._engine_keys[key] = None
        wk = self._weights_key()[0]
        if self._engine_keys[key] != wk:
            eng.load_state_dict(self.state_dict(), speaker_encoder=True)
            self._engine_keys[key] = wk
        return eng

    def load(self, model_name: str, model_dir: str = "models/SpeakerEncoder", device=None):
        """reference `SpeakerEncoder.load` (`model.py:93-106`): checkpoint dict with "model_state" (non-strict: GE2E
        similarity_weight/bias are training-only) and an optional "speakers" table."""
        self.device = device if device is not None else self.device
        path = model_dir.strip("/") + "/" + model_name
        checkpoint = torch.load(path, map_location="cpu")
        self.load_state_dict(checkpoint["model_state"], strict=False)
        self.speakers.update(checkpoint.get("speakers", {}))
        if self.verbose:
            print('Loaded speaker encoder "%s" trained to step %s' % (path, checkpoint.get("step")))

    # -- the path ----------------------------------------------------------------------------------
    def forward(self, utterances: torch.Tensor, hidden_init=None) -> torch.Tensor:
        """reference `SpeakerEncoder.forward` (`model.py:70-90`): (batch, n_frames, 40) -> unit-norm (batch, 256)."""
        if hidden_init is not None:
            raise NotImplementedError("a non-zero initial LSTM state is a training-time feature of the reference and is not supported")
        if not utterances.is_cuda:
            raise RuntimeError("autovc_b200.SpeakerEncoder runs on CUDA (sm_100a) only; move the input to the GPU")
        eng = self._engine(utterances.device)
        return torch.ops.autovc_b200.speaker_embed(utterances, eng.id)

    def _dev(self) -> torch.device:
        return self.device if self.device.type == "cuda" else torch.device("cuda")

    def embed_frames_batch(self, frames_batch: np.ndarray) -> torch.Tensor:
        """reference `embed_frames_batch` (`model.py:124-136`): numpy (batch, n_frames, 40) -> CPU tensor (batch, 256)."""
        frames = torch.from_numpy(np.ascontiguousarray(frames_batch, dtype=np.float32)).to(self._dev())
        return self.forward(frames).detach().cpu()

    def embed_partials(self, frames_batch, return_partials: bool = False):
        """`embed_utterance` from its `frames_batch` on (`model.py:181-190`): the partial utterances' embeddings,
        averaged and renormalised.  Returns a CPU tensor (256,), plus the partial embeddings when asked."""
        embeds, partials = self._embed_groups([frames_batch])
        if return_partials:
            return embeds[0].cpu(), partials.cpu()
        return embeds[0].cpu()

    def embed_speakers(self, frames_batches: Sequence) -> torch.Tensor:
        """Many utterances at once: frames_batches[i] is (n_partials_i, n_frames, 40).  -> CUDA tensor (len, 256), row i
        = what `embed_utterance` returns for utterance i.  All partials of all utterances run as one batch."""
        return self._embed_groups(frames_batches)[0]

    def _embed_groups(self, groups: Sequence):
        dev = self._dev()
        ts: List[torch.Tensor] = []
        for g in groups:
            t = torch.as_tensor(np.asarray(g, dtype=np.float32) if not isinstance(g, torch.Tensor) else g, dtype=torch.float32)
            if t.dim() != 3 or t.shape[0] == 0:
                raise ValueError("each utterance must be a non-empty (n_partials, n_frames, 40) batch of partial utterances")
            ts.append(t)
        if len({tuple(t.shape[1:]) for t in ts}) > 1:
            raise ValueError("all partial utterances must have the same (n_frames, 40) shape")
        counts = [t.shape[0] for t in ts]
        flat = torch.cat([t.to(dev, non_blocking=True) for t in ts], dim=0)
        partial_embeds = self.forward(flat)
        offsets = torch.tensor(np.concatenate([[0], np.cumsum(counts)]), dtype=torch.int32).to(dev)
        eng = self._engine(flat.device)
        return eng.se_reduce(partial_embeds, offsets), partial_embeds

    def embed_utterance(self, wav, using_partials: bool = True, return_partials: bool = False, **kwargs):
        """reference `embed_utterance` (`model.py:136-190`) with the mel front-end on the device too: wav (1-D numpy
        array or tensor) -> CPU tensor (256,).  `kwargs` go to `compute_partial_slices` like the reference's."""
        from . import audio
        if not using_partials:
            frames = audio.mel_spec_speaker_encoder(wav, cut=False, device=self._dev())
            embed = self.forward(frames[None, ...].contiguous())[0].cpu()
            return (embed, None, None) if return_partials else embed
        frames_batch, wave_slices, _ = audio.mel_spec_speaker_encoder(wav, cut=True, return_slices=True, device=self._dev(), **kwargs)
        embeds, partial_embeds = self._embed_groups([torch.stack(frames_batch)])
        if return_partials:
            return embeds[0].cpu(), partial_embeds.cpu(), wave_slices
        return embeds[0].cpu()

    def embed_utterances(self, wavs: Sequence, **kwargs) -> torch.Tensor:
        """The batching the reference leaves as a TODO (`model.py:141`): the partial utterances of ALL wavs run through
        the LSTM stack as one batch.  -> CUDA tensor (len(wavs), 256), row i = `embed_utterance(wavs[i])`."""
        from . import audio
        groups = [torch.stack(audio.mel_spec_speaker_encoder(w, cut=True, device=self._dev(), **kwargs)) for w in wavs]
        return self._embed_groups(groups)[0]

    def learn_speaker(self, speaker: Hashable, frames_batches: Sequence) -> torch.Tensor:
        """reference `learn_speaker` (`model.py:446-460`) from mel partials instead of a folder of wavs: the mean (not
        renormalised, as in the reference) of the utterance embeddings, cached in `self.speakers`."""
        embeds = self.embed_speakers(frames_batches)
        self.speakers[speaker] = embeds.mean(dim=0).cpu()
        return self.speakers[speaker]

    def as_embed_fn(self):
        """Adapter for `VoiceConverter(embed_speaker=...)`: speaker data = one utterance's (n_partials, n_frames, 40)."""
        return lambda data: self.embed_partials(data)

    @property
    def last_launch_count(self) -> int:
        return max((e.last_launch_count for e in self._engines.values()), default=0)
