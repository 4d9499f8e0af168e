#!/bin/bash
# One gpurun call that re-establishes the state of the build on a fresh B200 box (about four minutes):
#   /usr/local/graft/bin/gpurun --timeout 1200 -- 'bash tools/gpu_round_start.sh'
# = tools/r02_final_1gpu.sh: smoke under a short timeout (a hung kernel must cost seconds), the whole GPU suite, the default
# bench line and the reference arm, the other workloads' bench lines, the ncu launch list and one --set full capture each of
# the Burgers and heat stage kernels, summarised on the box (profiles/traffic.json is regenerated: copy
# gpurun_out/r02/final/traffic.json over it if a kernel changed).  Multi-GPU: tools/r02_run_4gpu.sh N (tests + bench + the
# reference's MultiGPU programs), tools/r02_run_8gpu.sh (bench only), tools/r02_ab_fused.sh N (launch-schedule A/B).
exec bash "$(dirname "$0")/r02_final_1gpu.sh"
