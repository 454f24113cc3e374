#!/bin/bash
# Per-kernel counts of the Blackwell-native (UTCHMMA / LDTM / STTM / UTMALDG / UTMASTG) and legacy (HMMA) instructions
# in the shipped library -> profiles/<prefix>_sass_summary.txt      usage: bash tools/sass_summary.sh r02
prefix=${1:-r02}
lib=compactor-vllm_b200/lib/libcompactor_sm100.so
out=profiles/${prefix}_sass_summary.txt
{
  echo "# cuobjdump -sass $lib : instruction counts per kernel (sm_100a)"
  printf "%-64s %8s %6s %6s %8s %8s %6s %6s %6s\n" kernel UTCHMMA LDTM STTM UTMALDG UTMASTG HMMA DFMA MEMBAR
  cuobjdump -sass "$lib" | awk '
    /Function :/ { if (name != "") flush(); name=$3; a=b=c=d=e=f=g=h=0; next }
    /UTCHMMA/ {a++} /LDTM/ {b++} /STTM/ {c++} /UTMALDG/ {d++} /UTMASTG/ {e++} / HMMA/ {f++} /DFMA/ {g++} /MEMBAR/ {h++}
    function flush() { cmd="c++filt " name; cmd | getline dem; close(cmd); sub(/\(.*/, "", dem); sub(/^void /, "", dem);
      printf "%-64s %8d %6d %6d %8d %8d %6d %6d %6d\n", substr(dem,1,64), a,b,c,d,e,f,g,h }
    END { if (name != "") flush() }' | sort
} > "$out"
echo "wrote $out"; head -40 "$out"
