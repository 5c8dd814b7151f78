#!/bin/bash
# One eager training step of the dpp.py model under `ncu --set full` (every kernel), reduced ON THE BOX to a CSV and a
# roofline table; plus a source-level capture of the tcgen05 convolution kernel.  Run through gpurun (one GPU).
set -u
mkdir -p gpurun_out
python benchmarks/profile_step.py --steps 1 > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --profile-from-start off -o /tmp/step_full python benchmarks/profile_step.py --steps 1 \
    > gpurun_out/ncu_step.log 2>&1
python benchmarks/tools/ncu_extract.py /tmp/step_full.ncu-rep gpurun_out/ncu_step_full.csv gpurun_out/ncu_step_roofline.md \
    >> gpurun_out/ncu_step.log 2>&1
ls -la /tmp/step_full.ncu-rep >> gpurun_out/ncu_step.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_kernel -c 6 \
    -o gpurun_out/ncu_conv_tc python benchmarks/profile_step.py --steps 1 > gpurun_out/ncu_conv.log 2>&1
tail -3 gpurun_out/ncu_step.log
