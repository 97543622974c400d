"""B200-native batched LUT-LDPC decoder (drop-in for the decode path of mmeidlinger/lut_ldpc).

The product is the C-ABI library liblutldpc_b200.so (include/lutldpc.h, sources in csrc/); this
package is the thin Python binding the tests and bench.py use, plus the host-side mirror of the
reference's LDPC_BER_Sim_LUT harness (ber_sim.py; the C++ mirror of LDPC_Code_LUT is host/ldpc_code_lut.hpp).  There is no CPU
fallback: importing works anywhere, creating a decoder needs the built library and a CUDA device.
"""
from .capi import (Decoder, LutLdpcError, SimCounters, INIT_CONT, INIT_QCHA, TRACE_INIT, TRACE_CN,  # noqa: F401
                   TRACE_VN, library_path, load_library, build_library)

__all__ = ["Decoder", "LutLdpcError", "SimCounters", "INIT_CONT", "INIT_QCHA", "library_path", "load_library",
           "build_library"]
