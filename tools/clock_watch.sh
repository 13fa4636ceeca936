#!/bin/bash
# Sample SM clock, power draw and throttle reasons every 50 ms while a command runs:  tools/clock_watch.sh <log> <command...>
log=$1; shift
nvidia-smi --query-gpu=timestamp,clocks.sm,clocks.mem,power.draw,power.limit,temperature.gpu,clocks_event_reasons.active,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown --format=csv -lms 50 -i 0 > "$log" 2>&1 &
smi=$!
"$@"
rc=$?
kill $smi
exit $rc
