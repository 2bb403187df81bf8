#!/bin/bash
# full GPU suite + MNIST parity runs through the DistriServer facade
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/b_gpu_tests.log 2>&1
echo "gpu tests exit $?" >> gpurun_out/b_gpu_tests.log
for m in mlp cnn; do
  timeout 300 python -m distriflow_b200.experiment.mnist.mnist_gpu --model $m --mode sync --parity --normalise --examples 4096 --lr 0.02 --out gpurun_out/b_parity_${m}_div255.json > /dev/null 2> gpurun_out/b_parity_${m}_div255.err
done
timeout 300 python -m distriflow_b200.experiment.mnist.mnist_gpu --model mlp --mode async --parity --examples 4096 --out gpurun_out/b_parity_mlp_raw255.json > /dev/null 2> gpurun_out/b_parity_mlp_raw255.err
tail -n 5 gpurun_out/b_gpu_tests.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/b_parity_*.json')):
    d=json.load(open(f)); p=d.get('parity',{})
    print(f, {k:p.get(k) for k in ('steps_compared','max_abs_loss_diff','max_abs_weight_diff','max_weight_movement','gpu_first','gpu_last')}, d['samples_per_sec'])
PY
tail -n 3 gpurun_out/b_parity_*.err
