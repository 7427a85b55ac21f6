cd $GRAFT_REPO_ROOT
python scratch/prof_w16.py 16 28672 4096 || exit 1
ncu --set full --clock-control none --import-source on -k regex:gemm_w8a16 -s 3 -c 1 -o /tmp/w16sk -f python scratch/prof_w16.py 16 28672 4096 > gpurun_out/r22_ncu.log 2>&1
python tools/ncu_summary.py /tmp/w16sk.ncu-rep > gpurun_out/r22_w16sk.csv
ncu -i /tmp/w16sk.ncu-rep --page details > gpurun_out/r22_w16sk_details.txt
ncu -i /tmp/w16sk.ncu-rep --page source --csv > gpurun_out/r22_w16sk_source.csv 2>/dev/null
