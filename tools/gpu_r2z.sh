#!/bin/bash
timeout 200 python tools/trace_step.py --layers 2 --batch 16 --model Llama-2-13b-hf --qt int4 --dump 5 2>&1 | grep -v "^DEBUG\|Warn\|warn" | grep -v "1793" | head -40
timeout 200 python tools/trace_step.py --layers 2 --batch 16 --model Llama-2-13b-hf --qt int4 --dump 7 2>&1 | grep -v "^DEBUG\|Warn\|warn" | grep -v "1793" | head -40
