"""bt2gpu: host-side mirror of the reference's SeedAligner / SwAligner call surface over the
C ABI in include/bt2gpu.h (libbt2gpu.so, hand-written sm_100a CUDA).  No CPU fallback."""
from . import structs, scoring, dpio, fmio, index  # noqa: F401
