#!/bin/bash
N=$1
bash tools/gpu_tpN.sh $N llama2-7b-nf4-bs1
bash tools/gpu_tpN.sh $N llama2-70b-int4-bs1-ctx2048
