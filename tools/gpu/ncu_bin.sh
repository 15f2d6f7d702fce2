# source-level ncu capture of the one-kernel binning at C3 (or $1); the hot-spot text is produced on the box
c=${1:-c3}
mkdir -p gpurun_out
timeout 200 python tools/prof_all.py $c > gpurun_out/plain_prof_$c.log 2>&1 || { tail -5 gpurun_out/plain_prof_$c.log; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:bin_small -c 1 -f -o gpurun_out/r01_bin_$c python tools/prof_all.py $c > gpurun_out/ncu_bin_$c.log 2>&1
tail -2 gpurun_out/ncu_bin_$c.log
python tools/ncu_source_hot.py gpurun_out/r01_bin_$c.ncu-rep > gpurun_out/bin_hot_$c.txt 2>&1
ncu -i gpurun_out/r01_bin_$c.ncu-rep --page source --csv > gpurun_out/bin_source_$c.csv 2>/dev/null
rm -f gpurun_out/r01_bin_$c.ncu-rep
cat gpurun_out/bin_hot_$c.txt
