"""Import the UNMODIFIED reference modules (`autovc.auto_encoder.model.AutoEncoder`, `autovc.speaker_encoder.model.
SpeakerEncoder`) - from `/root/reference` in the build container, from the staged copy `oracle/_ref/` (see
`oracle/make_ref.py`) on the GPU box.  Test/measurement infrastructure: only `tests/`, `oracle/make_golden*.py`,
`__graft_entry__.smoke()` and `bench.py`'s reference / baseline legs may import this.

The reference imports five third-party modules that are not installed in this image (librosa, soundfile, matplotlib,
webrtcvad, noisereduce).  None of them is touched by the generator or speaker-encoder forward, so they are
replaced by empty stub modules; the reference source itself is imported as is.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref")
LIVE = "/root/reference"
_STUBS = ["matplotlib", "matplotlib.pyplot", "librosa", "soundfile", "webrtcvad", "noisereduce"]


def reference_root(prefer_staged: bool = False) -> str | None:
    """Directory holding the reference's `autovc/` package, or None."""
    order = [STAGED, LIVE] if prefer_staged else [LIVE, STAGED]
    for root in order:
        if os.path.isfile(os.path.join(root, "autovc", "auto_encoder", "model.py")):
            return root
    return None


def verify(root: str = STAGED) -> bool:
    """True when every staged file still has the SHA-256 recorded when it was copied from the reference."""
    mf = os.path.join(root, "MANIFEST.json")
    if not os.path.isfile(mf):
        return False
    with open(mf) as f:
        files = json.load(f)["files"]
    for rel, sha in files.items():
        p = os.path.join(root, rel)
        if not os.path.isfile(p):
            return False
        with open(p, "rb") as g:
            if hashlib.sha256(g.read()).hexdigest() != sha:
                return False
    return True


def import_reference(prefer_staged: bool = False):
    """-> (AutoEncoder, SpeakerEncoder, root) of the unmodified reference; raises ImportError when no copy is reachable."""
    root = reference_root(prefer_staged)
    if root is None:
        raise ImportError("the reference package is neither at /root/reference nor staged under oracle/_ref "
                          "(run `python oracle/make_ref.py` in the build container)")
    if root == STAGED and not verify(root):
        raise ImportError("oracle/_ref does not match its manifest: re-run `python oracle/make_ref.py`")
    for m in _STUBS:
        if m not in sys.modules:
            sys.modules[m] = types.ModuleType(m)
    sys.modules["noisereduce"].reduce_noise = lambda *a, **k: None     # introspected at import (audio/tools.py:348)
    os.environ.setdefault("WANDB_MODE", "disabled")
    if root not in sys.path:
        sys.path.insert(0, root)
    from autovc.auto_encoder.model import AutoEncoder  # noqa: E402
    from autovc.speaker_encoder.model import SpeakerEncoder  # noqa: E402
    return AutoEncoder, SpeakerEncoder, root
