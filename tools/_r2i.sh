mkdir -p gpurun_out/r2i; O=gpurun_out/r2i
timeout 300 python -m pytest tests/test_gpu_umma.py tests/test_gpu_fullsize.py -q -m gpu -x -k "umma2 or pair" > $O/pytest_pair.log 2>&1; echo "pytest_pair rc=$?" >> $O/rc.txt
timeout 600 python tools/pair_check.py > $O/pair_check.txt 2>&1; echo "pair_check rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -5 $O/pytest_pair.log; cat $O/pair_check.txt | tail -12
