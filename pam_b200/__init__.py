"""pam_b200: B200-native AWFL dycore time step behind PAM's Dycore / PamCoupler interface.
Only what the hot path needs: csrc/ (CUDA kernels + C ABI), the host-side mirror of the reference interface
(coupler.py, dycore.py), the C++ shim (cxx/) and the synthetic-input generator shared by tests and bench."""
from .coupler import PamCoupler, DataManager, endrun   # noqa: F401
from .dycore import Dycore                             # noqa: F401
