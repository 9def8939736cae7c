#!/bin/bash
# Builds the tcgen05 probe (tools/umma_probe.cu) for sm_100a. Output: build/umma_probe
set -e
cd "$(dirname "$0")/.."
mkdir -p build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o build/umma_probe tools/umma_probe.cu
