import ctypes, sys
lib = ctypes.CDLL("/root/repo/compactor-vllm_b200/lib/libcompactor_sm100.so")
out = (ctypes.c_int * 8)()
print("rc", lib.cvllm_debug_prefill2_attrs(out))
print(dict(zip(["numRegs", "maxThreadsPerBlock", "staticSmem", "maxDynSmem", "localBytes", "constBytes", "wantSmem", "optinMax"], list(out))))
