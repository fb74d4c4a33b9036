for v in 21 23; do MCPT_SM_VARIANT=$v timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "variants or same_paths or tessellated or ragged or reference_stream or sample_ranges" 2>&1 | tail -2; done
python tools/tune_gpu.py scene01_lambert.mcpt 5,20,4 7,20,4 21,20,4 23,20,4
for r in 4 12 16 20; do echo "camr $r"; MCPT_SM_CAMR=$r python tools/tune_gpu.py scene01_lambert.mcpt 23,20,4 21,20,4; done
MCPT_SM_CAMR=8 python tools/tune_gpu.py scene01_lambert.mcpt 23,16,4 23,24,4
