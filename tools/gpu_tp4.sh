#!/bin/bash
set -u
bash tools/gpu_tpN.sh 4 llama2-7b-nf4-bs1
bash tools/gpu_tpN.sh 4 llama2-70b-int4-bs1-ctx2048
