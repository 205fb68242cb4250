for cfg in "1000 150" "3000 600" "3000 150" "500 50"; do
  ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/ctc_l_${cfg// /_}.csv python tools/profile_ctc.py $cfg > /dev/null 2>&1
done
