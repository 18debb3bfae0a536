#!/bin/bash
# On an N-GPU box: rebuild with each -D set and run the weak C2 and strong C4 benches at N ranks.  usage: N=8 tools/sweep_n.sh "<flags>" ...
cd "$(dirname "$0")/.."
N=${N:-8}
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
for f in "$@"; do
  make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc EXTRA="$f" > /dev/null 2>&1 || { echo "BUILD FAILED: $f"; continue; }
  for w in "--workload c2" "--workload c4 --scaling strong" ${SWEEP_EXTRA:+"$SWEEP_EXTRA"}; do
    timeout 300 $L bench.py --gpus $N --steps 200 --warmup 20 --no-cpu-baseline $w 2>/dev/null | tail -n 1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('%-28s %-34s %.0f Mrays/s  %.4f ms  check %s' % ('$f', '$w', d['value'], d['ms_per_step'], d['gather_check']))"
  done
done
make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc > /dev/null 2>&1
