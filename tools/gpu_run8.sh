mkdir -p gpurun_out
timeout 600 python tools/ncu_step.py painn > gpurun_out/painn_plain.log 2>&1; tail -1 gpurun_out/painn_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tn_grouped_kernel -c 2 -f -o gpurun_out/tn_grouped_painn python tools/ncu_step.py painn > gpurun_out/ncu_tn.log 2>&1; tail -2 gpurun_out/ncu_tn.log
