#!/bin/bash
# ncu --set full of conv_gemm_kernel / conv_wgrad_mn_kernel on three Inception-v3 shapes (one GPU).  The raw pages are
# written under gpurun_out/; scripts/ncu_summarise.py turns them into profiles/ncu_inception_<case>_summary.txt.
mkdir -p gpurun_out
for c in inc35 inc17 inc8; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"conv_gemm_kernel|conv_wgrad_mn_kernel" \
    -c 2 -s 8 -f -o gpurun_out/ncu_inception_$c python scripts/ncu_conv_case.py $c > gpurun_out/ncu_inception_$c.log 2>&1
  ncu -i gpurun_out/ncu_inception_$c.ncu-rep --page raw --csv > gpurun_out/ncu_inception_${c}_raw.csv 2>/dev/null
  echo "$c rc=$? $(wc -c < gpurun_out/ncu_inception_${c}_raw.csv) bytes"
done
