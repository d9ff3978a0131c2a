"""Bare-name entry to the B200 per-frame front end (both BasicEncoder4 encoders + patch gathers)
for the unmodified reference code base.  With this directory on sys.path:

    import dpvo_b200_patchify
    self.network.patchify = dpvo_b200_patchify.Patchifier.from_reference(self.network.patchify)   # dpvo.py, after :112

Same call and return tuple as net.py:106-153 (inference, 'RANDOM' centroid selection).
"""
from _bridge import load as _load

Patchifier = _load("encoder").Patchifier
BasicEncoder4 = _load("encoder").BasicEncoder4
