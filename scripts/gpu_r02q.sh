# AR chain latency experiment: baseline chain kernel vs the 16-byte-load / unrolled-by-four kernel, cluster size and lanes per row
TAG=r02q
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone_$TAG.log 2>&1; tail -3 gpurun_out/pytest_noalloc_alone_$TAG.log
cp maria_b200/libmaria_atm.so /tmp/lib_new.so
cp maria_b200/variants/libmaria_atm_arbase.so maria_b200/libmaria_atm.so
timeout 300 python scripts/chain_bench.py --cl 0 > gpurun_out/chain_base_$TAG.txt 2>&1; grep process gpurun_out/chain_base_$TAG.txt
cp /tmp/lib_new.so maria_b200/libmaria_atm.so
timeout 600 python scripts/chain_bench.py --cl 0 4 6 8 > gpurun_out/chain_new_$TAG.txt 2>&1; grep process gpurun_out/chain_new_$TAG.txt
MATM_AR_MINL_EXPERIMENT=16 timeout 600 python scripts/chain_bench.py --cl 0 8 > gpurun_out/chain_new_l16_$TAG.txt 2>&1; grep process gpurun_out/chain_new_l16_$TAG.txt
MATM_AR_NT_EXPERIMENT=256 timeout 600 python scripts/chain_bench.py --cl 0 8 > gpurun_out/chain_new_nt256_$TAG.txt 2>&1; grep process gpurun_out/chain_new_nt256_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py -m gpu -q -x -k "recurrence or golden or stages or c5 or pipelined or 3d" > gpurun_out/pytest_chain_$TAG.log 2>&1; tail -4 gpurun_out/pytest_chain_$TAG.log
MATM_AR_CL_EXPERIMENT=8 timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py -m gpu -q -x -k "recurrence or golden or stages or c5" > gpurun_out/pytest_chain_cl8_$TAG.log 2>&1; tail -4 gpurun_out/pytest_chain_cl8_$TAG.log
MATM_AR_NT_EXPERIMENT=256 MATM_AR_CL_EXPERIMENT=8 timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "recurrence or c5" > gpurun_out/pytest_chain_nt256cl8_$TAG.log 2>&1; tail -4 gpurun_out/pytest_chain_nt256cl8_$TAG.log
