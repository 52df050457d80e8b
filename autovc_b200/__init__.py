"""autovc_b200 - B200-native (sm_100a) AutoVC generator forward pass behind the reference's Python API.

    from autovc_b200 import AutoEncoder           # drop-in for autovc.auto_encoder.model.AutoEncoder
    mel, mel_post, codes = AutoEncoder().cuda_forward...   (see model.py)

The compute path is hand-written CUDA in `csrc/` behind the C ABI in `include/autovc_b200.h`; importing this
package does not need a GPU, running it does (no CPU fallback).
"""
from .model import AutoEncoder, overlap_average  # noqa: F401
from .engine import Engine  # noqa: F401
from .voice_converter import HostPipeline, Utterance, VoiceConverter, WavVoiceConverter, bucket_by_length, chunk_mel, shard_range  # noqa: F401
from .speaker_encoder import SpeakerEncoder  # noqa: F401
from . import audio  # noqa: F401
from .audio import MelFrontend, compute_partial_slices, mel_spec_auto_encoder, mel_spec_speaker_encoder  # noqa: F401
from .vocoder import WaveRNN  # noqa: F401
from .synth import synth_inputs, synth_speaker_encoder, synth_state_dict, synth_wavernn_state_dict, state_dict_spec  # noqa: F401

__all__ = ["AutoEncoder", "SpeakerEncoder", "WaveRNN", "synth_wavernn_state_dict", "audio", "MelFrontend", "compute_partial_slices", "mel_spec_auto_encoder",
           "mel_spec_speaker_encoder", "Engine", "synth_speaker_encoder", "overlap_average", "VoiceConverter", "WavVoiceConverter", "Utterance", "HostPipeline", "shard_range",
           "bucket_by_length", "chunk_mel", "synth_inputs", "synth_state_dict", "state_dict_spec"]
