TAG=r02x
mkdir -p gpurun_out
cp maria_b200/libmaria_atm.so /tmp/lib_new.so
cp maria_b200/variants/libmaria_atm_artrace.so maria_b200/libmaria_atm.so
timeout 300 python scripts/chain_trace.py 7 4 0 > gpurun_out/chain_trace_$TAG.txt 2>&1; cat gpurun_out/chain_trace_$TAG.txt | tail -30
cp /tmp/lib_new.so maria_b200/libmaria_atm.so
