"""kineticgas_b200 -- B200-native collision-integral hot path of KineticGas behind a C-ABI (include/kgcuda.h)."""
from .model import DeviceModel, OmegaPlan, OmegaTable, mie_desc, power_sum_desc, measure_fp64_peak  # noqa: F401
from .kingas import MieKinGas, QuantumMie, TCache, LJSpline, LJTS, ModTangToennis, HFD_B2, Patowski  # noqa: F401
from ._lib import pinned_empty, pinned_copy  # noqa: F401
from .eos import DeviceSaftVRMie  # noqa: F401
