O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_extend.py -x -q -m gpu -k "decode or mixed" 2>&1 | tail -3
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/decode_c2 python scripts/prof_kernels.py decode 4 > $O/ncu_decode.log 2>&1; echo "ncu decode rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/decode_c4_fp8 python scripts/prof_kernels.py decode_fp8 4 > $O/ncu_decode_fp8.log 2>&1; echo "ncu decode fp8 rc=$?"
