#!/bin/bash
# Evidence that the MLP kernels are hand-written tcgen05 / TMEM / bulk-TMA code: SASS mnemonic counts of the built
# library and the ptxas register / spill report of the two kernel files.  Runs without a GPU.
#   bash scripts/sass_counts.sh > profiles/r02_sass_counts.txt
set -e
cd "$(dirname "$0")/.."
SO=nerf_rs_b200/libnerf_b200.so
echo "# cuobjdump -sass $SO | grep -c <mnemonic>   ($(date -u +%Y-%m-%d), nvcc $(nvcc --version | grep -o 'V[0-9][0-9.]*' | tail -1))"
cuobjdump -sass $SO > /tmp/_sass.txt
for m in UTCHMMA UTCHMMA.2CTA LDTM STTM UTCBAR UTCBAR.2CTA.MULTICAST UBLKCP.S.G UTMALDG SYNCS.ARRIVE.TRANS64 SYNCS.PHASECHK \
         UTCATOMSWS HMMA WGMMA STL LDL; do
  printf "%-24s %s\n" "$m" "$(grep -c "[ .]$m[ .;]" /tmp/_sass.txt || true)"
done
echo
echo "# kernels in the library (cuobjdump -sass | grep Function)"
grep "Function :" /tmp/_sass.txt | sed 's/.*Function : //' | c++filt | sed 's/(anonymous namespace):://; s/(nerf::.*//; s/(float.*//' | sort
echo
echo "# ptxas -v (registers are the launch bound of 640 threads; setmaxnreg gives the epilogue warps 112, the control warps 32)"
cd nerf_rs_b200/csrc
for f in mlp_tc mlp_tc_x3 mlp_tc_tf32 sampling; do
  extra=""; [ $f = sampling ] && extra="-fmad=false"
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
       -Xptxas -v $extra -c $f.cu -o /tmp/_$f.o 2>&1 | grep -E "Compiling entry|Used|spill" | sed 's/ptxas info    : //' | c++filt
done
