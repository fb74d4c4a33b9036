for sc in scene01.mcpt scene02.mcpt scene03.mcpt heightfield:300; do echo "== $sc"; python tools/tune_gpu.py $sc 5,20,4 7,20,4; done
