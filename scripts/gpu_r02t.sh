# is test_pipelined_realisations_equal_sequential_ones flaky, and since when?  + per-phase clocks of the chain step
TAG=r02t
mkdir -p gpurun_out
cp maria_b200/libmaria_atm.so /tmp/lib_new.so
for v in new arbase; do
  [ $v = new ] && cp /tmp/lib_new.so maria_b200/libmaria_atm.so || cp maria_b200/variants/libmaria_atm_$v.so maria_b200/libmaria_atm.so
  for n in 1 2 3 4 5 6; do
    timeout 300 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "pipelined" > gpurun_out/pipe_${v}_${n}_$TAG.log 2>&1
    echo "$v run $n: $(tail -1 gpurun_out/pipe_${v}_${n}_$TAG.log) $(grep -o 'AssertionError: ([0-9], [0-9])' gpurun_out/pipe_${v}_${n}_$TAG.log | head -1)"
  done
done
cp maria_b200/variants/libmaria_atm_artrace.so maria_b200/libmaria_atm.so
timeout 300 python scripts/chain_trace.py 7 4 0 > gpurun_out/chain_trace_$TAG.txt 2>&1; cat gpurun_out/chain_trace_$TAG.txt | tail -40
cp /tmp/lib_new.so maria_b200/libmaria_atm.so
