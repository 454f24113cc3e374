"""Reserved batch slot (reference: config/constants.py:3-4).

Slot 0 of the page table / length tables is never handed to a sequence; CUDA
graph padding rows map to it and every store kernel skips it.
"""
RESERVED_BATCH = 0
# the reference wraps this in tl.constexpr for its Triton kernels; here it is a plain int
TRITON_RESERVED_BATCH = RESERVED_BATCH
