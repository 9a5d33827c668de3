#!/bin/bash
# narrow-pass experiment: 8 lights per pass (9 ray slots, two passes for 16 lights) at higher occupancy
L=raytracing-programs_b200/lib
bash tools/gpu_variants2.sh r02m_lb16 $L/librt_b200.so
export RT_LIGHT_BATCH=8
bash tools/gpu_variants2.sh r02m_lb8 $L/librt_b200.so $L/librt_b200_n5.so $L/librt_b200_n6.so $L/librt_b200_n8.so
