"""B200-native layer-dissemination engine — Python driver over the C ABI (include/dissem.h).

This package is test/bench plumbing around libdissem.so; the product is the shared library.  It mirrors
the reference's vocabulary (nodes, layers, assignment, modes) and never computes on the CPU: if the CUDA
library is missing or no GPU is visible, the compute entry points raise.

Import it as `import dissem_b200` (alias module at the repo root) or
`importlib.import_module("distributed-llm-dissemination_b200")`.
"""
from ._capi import (CHUNK_BYTES, FILL_PATTERN, FILL_POISON, FILL_ZERO, NOALLOC_MAX_CHUNKS, DissemError, Engine, HostBuffer,  # noqa: F401
                    Xfer, XFER_FILE, XFER_PULL, checksum_noalloc, lib, lib_path, load_library, plan_align, plan_directions, plan_localize, plan_optimize, plan_relay, xfer_round)
from .host import (HostError, parse_plan, plan, run_distribution)  # noqa: F401

__all__ = ["Engine", "Xfer", "DissemError", "HostError", "lib", "lib_path", "load_library", "plan", "parse_plan",
           "run_distribution", "CHUNK_BYTES", "FILL_ZERO", "FILL_PATTERN", "FILL_POISON"]
