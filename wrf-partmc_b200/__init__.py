"""B200-native PartMC per-grid-cell aerosol step (sort -> coagulation ->
condensation -> transport), the one hot path of WRF-PartMC.

The product is ``libpmcgpu.so`` (hand-written sm_100a CUDA behind the C ABI of
``include/pmcgpu.h``); ``host.pmcgpu`` is the thin host-side mirror of the
reference's entry points used by the tests and the benchmark.  There is no CPU
fallback: importing works anywhere, but every call needs the built library and
a CUDA device.
"""
from .host.pmcgpu import (AeroData, PartMCGPU, PmcGpuError, comm_unique_id, library_path, load_library,  # noqa: F401
                          KERNEL_ADDITIVE, KERNEL_BROWN, KERNEL_BROWN_CONT, KERNEL_BROWN_FREE, KERNEL_CONSTANT, KERNEL_SEDI, KERNEL_ZERO, WEIGHT_NONE, WEIGHT_POWER)
