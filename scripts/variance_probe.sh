#!/bin/bash
# run-to-run variance of the blind-rotate launch: per-step times with nvidia-smi clocks/power sampled at 10 Hz
mkdir -p gpurun_out
nvidia-smi --query-gpu=timestamp,clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_throttle_reasons.active --format=csv -lms 100 > gpurun_out/smi_trace.csv &
SMI=$!
python scripts/ab_knob.py generic_gadget --rounds 8 > gpurun_out/variance.json 2> gpurun_out/variance.err
kill $SMI
cat gpurun_out/variance.json
awk -F, 'NR>1{c[$2" "$6]++} END{for(k in c) print c[k], k}' gpurun_out/smi_trace.csv | sort -rn | head
awk -F, 'NR>1{gsub(/ W/,"",$4); if($4>m)m=$4} END{print "max power", m}' gpurun_out/smi_trace.csv
