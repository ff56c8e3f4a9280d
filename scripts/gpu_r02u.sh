# what makes the pipelined realisations differ: visibility of the global row copy (fence variant) / acquire scope of the wait
TAG=r02u
mkdir -p gpurun_out
cp maria_b200/libmaria_atm.so /tmp/lib_new.so
for v in FENCE_EXPERIMENT ACQ_CLUSTER; do
  cp maria_b200/variants/libmaria_atm_$v.so maria_b200/libmaria_atm.so
  for n in 1 2 3 4 5 6; do
    timeout 300 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "pipelined" > gpurun_out/pipe_${v}_${n}_$TAG.log 2>&1
    echo "$v run $n: $(tail -1 gpurun_out/pipe_${v}_${n}_$TAG.log) $(grep -o 'AssertionError: ([0-9], [0-9])' gpurun_out/pipe_${v}_${n}_$TAG.log | head -1)"
  done
  timeout 300 python scripts/chain_bench.py > gpurun_out/chain_${v}_$TAG.txt 2>&1; grep -E "process [47]" gpurun_out/chain_${v}_$TAG.txt
done
cp /tmp/lib_new.so maria_b200/libmaria_atm.so
