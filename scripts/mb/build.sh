#!/bin/bash
# builds the micro-benchmarks into scripts/mb/bin/ (git-ignored; travels to the GPU box with the snapshot)
set -e
cd "$(dirname "$0")/../.."
mkdir -p scripts/mb/bin
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I bayesbridge_b200/csrc scripts/mb/mb_round.cu -o scripts/mb/bin/mb_round
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/mb/mb_dmma.cu -o scripts/mb/bin/mb_dmma
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I bayesbridge_b200/csrc scripts/mb/mb_reduce.cu -o scripts/mb/bin/mb_reduce
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I bayesbridge_b200/csrc scripts/mb/mb_potrf.cu -o scripts/mb/bin/mb_potrf
