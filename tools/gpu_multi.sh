#!/bin/bash
# multi-GPU pass (gpurun --gpus N): raw copy ceiling at 1/2/4/N GPUs, then bench.py under torchrun at 2, 4, N ranks
mkdir -p gpurun_out
N=${N:-8}
rm -f gpurun_out/copy_ceiling.json
[ -n "$SKIP_PROBE" ] || for g in 1 2 4 $N; do
  [ $g -le $N ] || continue
  echo "== probe gpus=$g"; timeout 300 python tools/pcie_probe.py --gpus $g --out gpurun_out/copy_ceiling.json 2>&1 | tail -3
done
[ -n "$SKIP_PROBE" ] || { echo "== probe gpus=$N affinity"; timeout 300 python tools/pcie_probe.py --gpus $N --affinity --out gpurun_out/copy_ceiling.json 2>&1 | tail -3
echo "== probe gpus=$N chunk 16 MB"; timeout 300 python tools/pcie_probe.py --gpus $N --chunk-mb 16 --out gpurun_out/copy_ceiling.json 2>&1 | tail -3; }
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
for g in 2 4 $N; do
  [ $g -le $N ] || continue
  extra="--no-variants"; [ $g -eq $N ] && extra=""
  echo "== bench n=$g $extra"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $((29500+g)) bench.py --gpus $g --steps 20 --warmup 3 $extra > gpurun_out/bench_n$g.json 2> gpurun_out/bench_n$g.err; echo "rc=$?"
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_n$g.json'))
    e=d['e2e']
    print('n', d['n_gpus'], 'value %.3g' % d['value'], 'ms %.4f' % d['ms_per_step'], 'e2e %.3g blocks/s' % e['value'], 'ms %.2f' % e['ms_per_step'], 'GB/s/gpu %.1f' % e['gbps_each_way'], 'ceiling', e['copy_ceiling']['gbps_each_way_per_gpu'], 'frac %.2f' % e['frac_of_copy_ceiling'])
    print(e.get('single_process_all_gpus'))
except Exception as ex:
    print('parse failed', ex)
PY
done
