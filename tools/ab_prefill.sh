# same-box A/B of prefill library variants: bash tools/ab_prefill.sh <variant> ...   ("default" = the product library)
for i in 1 2; do
  for v in "$@"; do
    if [ "$v" = default ]; then timeout 100 python tools/probe_prefill.py 2>&1 | grep "N=32768"
    else CVLLM_LIB=compactor-vllm_b200/lib/variants/lib_$v.so timeout 100 python tools/probe_prefill.py 2>&1 | grep "N=32768"; fi
  done
done
