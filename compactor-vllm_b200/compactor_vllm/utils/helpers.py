"""Side-stream helper (reference: utils/helpers.py:6-28).

Same protocol: STORE_STREAM waits for the caller's stream, the function runs
inside `with STORE_STREAM`, inputs are recorded on STORE_STREAM and tensor
outputs on the caller's stream.  The reference names `torch.cuda.default_stream()`
for "the caller's stream"; every call site of both code bases runs on the default
stream, where the two are the same, and forking from the CURRENT stream is what
lets a whole step be captured in a CUDA graph (a capture cannot depend on the
legacy default stream).  The native wrappers always launch on
torch.cuda.current_stream(), so they follow the `with` block.
"""
from collections.abc import Callable

import torch


def maybe_execute_in_stream(fn: Callable, *args, STORE_STREAM: torch.cuda.Stream = None, **kwargs):
    if STORE_STREAM is None:
        return fn(*args, **kwargs)
    tensors = [a for a in args if isinstance(a, torch.Tensor)]
    tensors += [v for v in kwargs.values() if isinstance(v, torch.Tensor)]
    obj = getattr(fn, "__self__", None)
    if isinstance(obj, torch.Tensor):
        tensors.append(obj)
    caller = torch.cuda.current_stream(STORE_STREAM.device)
    STORE_STREAM.wait_stream(caller)
    with torch.cuda.stream(STORE_STREAM):
        output = fn(*args, **kwargs)
    for t in tensors:
        t.record_stream(STORE_STREAM)
    outs = output if isinstance(output, tuple) else (output,)
    for o in outs:
        if isinstance(o, torch.Tensor):
            o.record_stream(caller)
    return output
