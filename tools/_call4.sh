O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_dropin.py -m gpu -x -q -k "config3 or config5" -s ) > $O/c4_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 $O/c4_pytest.log
