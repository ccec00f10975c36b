"""intercession-b200: the physics tick and world-line record / replay / divergence scan of
meta-engineer/Intercession as hand-written sm_100a CUDA behind a C ABI (include/icx.h).

Python here is only the test / bench driver over that ABI; the host-side mirror of the
reference's C++ interface lives in intercession_b200/host/*.hpp.
"""
from . import abi  # noqa: F401

__version__ = "0.1.0"
