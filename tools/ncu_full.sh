# usage: tools/ncu_full.sh <label> [target.py] [kernel-regex] [count]
# One `ncu --set full` capture with source of the named kernels, after the target ran clean without ncu.  The
# raw page and the per-instruction source pages are exported to CSV ON THE BOX (gpurun copies back at most 64 MiB;
# a report above 40 MB is dropped after the export).  Read here with tools/ncu_summary.py and tools/ncu_lines.py.
L=${1:-r2a}
T=${2:-tools/ncu_target2.py}
K=${3:-k_trace_stream_spec|k_path_persistent|k_ltc}
C=${4:-12}
mkdir -p gpurun_out
timeout 300 python $T || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$K" -c $C \
  -o gpurun_out/${L}_full -f python $T > gpurun_out/${L}_full.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/${L}_full.ncu-rep --page raw --csv > gpurun_out/${L}_raw.csv 2>/dev/null
for k in $(echo "$K" | tr '|' ' '); do
  ncu -i gpurun_out/${L}_full.ncu-rep --page source --csv --kernel-name regex:$k > gpurun_out/${L}_src_$k.csv 2>/dev/null
done
ls -la gpurun_out/${L}_*
sz=$(stat -c %s gpurun_out/${L}_full.ncu-rep)
if [ "$sz" -gt 40000000 ]; then rm gpurun_out/${L}_full.ncu-rep; echo "report dropped ($sz bytes); CSV pages kept"; fi
