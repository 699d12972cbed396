#!/bin/bash
# inst_count.sh [variant]: executed warp instructions and duration of one 1080p LDB-medium P frame of k_encode_frame (third
# launch) -- the deterministic yardstick for "did this change remove work", free of the +-5 % of the wall-clock numbers.
cd "$(dirname "$0")/../.."
[ -n "$1" ] && [ "$1" != base ] && export LD_LIBRARY_PATH=$PWD/build/variants/$1
export THORCUDA_ENC_CTAS=${CTAS:-1}   # one CTA walks every superblock: no wavefront waiting, duration = SM-seconds per frame
python profiles/r2/enc_cmd.py LDB_medium_complexity 1920 1080 4 32 > /tmp/enc_cmd.txt
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,sm__cycles_active.sum --clock-control none -k regex:k_encode_frame -s 2 -c 1 --csv $(cat /tmp/enc_cmd.txt) 2>/dev/null | grep -E "smsp__inst_executed|gpu__time_duration|sm__cycles_active" | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | tr -d '"'
