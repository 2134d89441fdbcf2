timeout 300 python tools/probe_tex.py > gpurun_out/r1n_texture_unit.json 2> gpurun_out/probe.err; tail -2 gpurun_out/probe.err; python -c "
import json;d=json.load(open('gpurun_out/r1n_texture_unit.json'));print(json.dumps(d['two_dimensional_on_ltc_table'],indent=1))"
timeout 900 python -m pytest tests/test_gpu_reference_megakernel.py -m gpu -q -s --tb=line 2>&1 | cut -c1-1200 | tail -30
