for i in 1 2 3; do
  CVLLM_LIB=compactor-vllm_b200/lib/variants/lib_p2old.so timeout 100 python tools/probe_prefill.py 2>&1 | grep "N=32768"
  timeout 100 python tools/probe_prefill.py 2>&1 | grep "N=32768"
done
