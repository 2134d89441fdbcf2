N=$1
timeout 600 python -m pytest tests/test_gpu_frames.py -m gpu -x -q -k "instrumented or wide" 2>&1 | tail -3
bash tools/_mg.sh $N
