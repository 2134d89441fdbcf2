timeout 900 python -m pytest tests/test_gpu_reference_megakernel.py -m gpu -x -q -s 2>&1 | tail -25
timeout 600 python tools/g0_megakernel.py > gpurun_out/r1n_g0_megakernel.json 2> gpurun_out/g0.err; tail -2 gpurun_out/g0.err; cat gpurun_out/r1n_g0_megakernel.json
