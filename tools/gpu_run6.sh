mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_training.py -q -x -s > gpurun_out/train_test.log 2>&1; echo "training test rc=$?"; tail -15 gpurun_out/train_test.log
