#!/bin/bash
O=gpurun_out; mkdir -p $O
L=raytracing-programs_b200/lib
bash tools/gpu_variants2.sh r02f $L/librt_b200.so $L/librt_b200_t11.so $L/librt_b200_t9.so $L/librt_b200_t3.so $L/librt_b200_t13_5.so
bash tools/gpu_ncu_kshade.sh r02f
