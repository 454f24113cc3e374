"""Hand-off latencies of prefill_attn2_kernel from in-kernel clock64 stamps (CTA (0,0,0) = the heaviest query tile of
head pair 0; first 256 KV tiles).  Needs a library built with -DCVLLM_PREFILL_TRACE (tools/build_variant.sh), selected
with CVLLM_LIB.  Rows: 0 issuer woke on p_full0, 1 issuer done issuing PV0(j) + S0(j+1), 2 woke on p_full1, 3 done
issuing PV1(j) + S1(j+1), 4 / 6 softmax group 0 / 1 woke on s_full(j), 5 / 7 it arrived on p_full(j)."""
import ctypes as C
import math
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import numpy as np
import torch
from ref_compare import _inputs, _paged, HQ, HKV, D, PS
from compactor_vllm import _native as N
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
n = 32768
q, k, v, _ = _inputs(torch, n, 0, dev)
kc, vc, pt, bm = _paged(torch, 1, n + PS, 1, dev, fill=False)
zl = torch.zeros((1, HKV), dtype=torch.int32, device=dev)
cu = torch.tensor([0, n], dtype=torch.int32, device=dev)
for _ in range(3):
    causal_sparse_varlen_with_cache(q, k, v, kc, vc, zl, pt, bm, cu, n, 0, HKV, PS, 1 / math.sqrt(D))
torch.cuda.synchronize()
raw = C.CDLL(str(N.LIB_PATH))
buf = np.zeros((8, 256), dtype=np.int64)
rc = raw.cvllm_debug_prefill_trace_read(C.c_void_p(buf.ctypes.data))
assert rc == 0, rc
t = buf - buf[0, 0]
lo, hi = 32, 250  # steady state
def med(a):
    return float(np.median(a[lo:hi]))
period = med(np.diff(t[0]))
print(f"{os.environ.get('CVLLM_LIB', 'default').split('/')[-1]}: period per KV tile {period:.0f} clk")
print(f"  issuer: woke p_full0 -> issued PV0+S0' {med(t[1] - t[0]):.0f} | -> woke p_full1 {med(t[2] - t[1]):.0f} | "
      f"-> issued PV1+S1' {med(t[3] - t[2]):.0f} | -> woke p_full0(j+1) {med(t[0][1:] - t[3][:-1]):.0f}")
for g in (0, 1):
    w, a = t[4 + 2 * g], t[5 + 2 * g]
    issued = t[1 + 2 * g]  # S_g(j+1) issued at row 1/3 of iteration j
    print(f"  head {g}: S(j+1) issued -> softmax woke {med(w[1:] - issued[:-1]):.0f} | softmax woke -> arrived {med(a - w):.0f} | "
          f"arrived -> issuer woke {med(t[2 * g] - a):.0f}")
np.save(str(ROOT / "gpurun_out" / f"prefill_trace_{os.environ.get('CVLLM_LIB', 'default').split('/')[-1]}.npy"), buf)
