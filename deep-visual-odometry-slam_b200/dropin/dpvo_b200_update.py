"""Bare-name entry to the B200 Update operator for the unmodified reference code base.  With this
directory on sys.path (as for cuda_corr / cuda_ba / lietorch_backends):

    import dpvo_b200_update
    self.network.update = dpvo_b200_update.Update.from_reference(self.network.update)   # dpvo.py, after :112

re-exports update.Update (net.py:23-88 on the tensor cores) and the layer wrapper linear384.
"""
from _bridge import load as _load

Update = _load("update").Update
linear384 = _load("update_ops").linear384
