#!/bin/bash
# One `ncu --set full` capture of the step kernel for a bench configuration, summarised ON THE BOX (the report is too
# large to bring back): tools/ncu_capture.sh <tag> <env-us per launch> <bench args...>
# Writes gpurun_out/<tag>_ncu_summary.txt and gpurun_out/<tag>_sass_hist.txt.  Run the same bench command without
# ncu first; numbers printed under ncu are not bench values.
tag=$1; units=$2; shift 2
rep=/tmp/$tag.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 4 -c 1 -f -o /tmp/$tag \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${tag}_ncu.log 2>&1
python tools/ncu_report.py $rep $units > gpurun_out/${tag}_ncu_summary.txt
ncu -i $rep --page source --csv --print-source sass > /tmp/$tag.src.csv 2>/dev/null
python tools/sass_hist.py /tmp/$tag.src.csv $units > gpurun_out/${tag}_sass_hist.txt
gzip -c /tmp/$tag.src.csv > gpurun_out/${tag}_src.csv.gz
head -3 gpurun_out/${tag}_sass_hist.txt
