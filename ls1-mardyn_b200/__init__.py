"""ls1-MarDyn short-range force + integrate path on B200 (sm_100a).

Host-side mirror of the reference's operator surface for this path, as a thin ctypes layer over the C ABI
(include/ls1g.h -> libls1g.so).  The product path is the CUDA library only: importing works without a GPU
(symbol checks, host-side geometry), every compute call fails loudly without one.  There is no CPU fallback.
"""
from .binding import (Container, Ls1gError, build, lib, lib_path, decomp_grid, decomp_box, fp64_peak_tflops, device_count,
                      exchange_group, run_group, step_group, comm_init_local, comm_unique_id, components_from_tables,
                      lj_component, FIELDS, EXPORTS)
from . import phasespace

__all__ = ["Container", "Ls1gError", "build", "lib", "lib_path", "decomp_grid", "decomp_box",
           "fp64_peak_tflops", "device_count", "exchange_group", "run_group", "step_group", "comm_init_local", "comm_unique_id", "components_from_tables",
           "lj_component", "FIELDS", "EXPORTS", "phasespace"]
