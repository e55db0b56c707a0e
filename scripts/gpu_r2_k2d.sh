#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/k2_profile.py --ranks 16 --one-pass 0 > gpurun_out/r02m_k2_apply_profile.jsonl 2> gpurun_out/r02m.err; cat gpurun_out/r02m_k2_apply_profile.jsonl; tail -3 gpurun_out/r02m.err
