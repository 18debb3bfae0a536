#!/bin/bash
# ncu captures of the timed kernels at HEAD, summarised ON the box (only the small files travel back)
SHA=$1
for w in c2 c3 c4 c5; do
  extra=""; [ $w = c5 ] && extra="--steps 6"
  bash tools/prof.sh r02_h_$w --workload $w
  python tools/ncu_summary.py --commit $SHA --tag r02h $w=gpurun_out/r02_h_$w.ncu-rep > /dev/null 2>&1
  if [ $w = c2 ] || [ $w = c4 ]; then
    ncu -i gpurun_out/r02_h_$w.ncu-rep --page source --csv --print-source sass,cuda > /tmp/src_$w.csv 2>/dev/null
    python profiles/ncu_func.py /tmp/src_$w.csv voxelraytracer_b200/csrc | head -30 >> profiles/r02h_ncu_$w.txt
  fi
  rm -f gpurun_out/r02_h_$w.ncu-rep
done
cp profiles/ncu_summary.json profiles/r02h_ncu_*.txt gpurun_out/
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02h_launches_c2.csv python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/r02_h_launches_c2.log 2>&1; echo launches rc=$?
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k ray_queue > gpurun_out/r02_pytest_queue.log 2>&1; echo "queue test rc=$?"; tail -3 gpurun_out/r02_pytest_queue.log
ls -la gpurun_out | head -30
