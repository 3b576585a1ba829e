for a in 0 1; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/ext_$a.csv python profiles/ext_case.py $a > gpurun_out/ncu1.log 2>&1
python profiles/launch_summary.py gpurun_out/ext_$a.csv | grep query
done
