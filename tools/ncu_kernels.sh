#!/bin/bash
# ncu --set full captures of the GEMM / conv kernel at the UNet's config-2 shapes and of the HBM-bound kernels.
# Run under gpurun from the repo root; reports land in gpurun_out/ (summarise with tools/ncu_summary.py full <rep>).
set -u
mkdir -p gpurun_out
python tools/bench_gemm.py > gpurun_out/gemm_plain.log 2>&1 || exit 1
python tools/bench_norm.py > gpurun_out/norm_plain.log 2>&1 || exit 1
i=0
for only in "conv 160->160 @64 (+res)" "conv 320->320 @32" "conv 640->320 @32" "conv 320->320 @64" "geglu 160->1280" "qkv 160->768" "attn-out 160->160"; do
  i=$((i+1))
  ONLY="$only" ncu --set full --clock-control none --import-source on -k regex:umma_gemm -s 3 -c 1 -f \
    -o gpurun_out/gemm_$i python tools/bench_gemm.py > gpurun_out/ncu_gemm_$i.log 2>&1
done
ncu --set full --clock-control none -k regex:"gn_coop|add_layernorm" -s 3 -c 1 -f -o gpurun_out/gn_coop python tools/bench_norm.py > gpurun_out/ncu_gn.log 2>&1
ncu --set full --clock-control none -k regex:add_layernorm -s 3 -c 1 -f -o gpurun_out/ln python tools/bench_norm.py > gpurun_out/ncu_ln.log 2>&1
ls -la gpurun_out/*.ncu-rep
