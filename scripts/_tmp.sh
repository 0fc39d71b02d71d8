export NORI_GPU_LIB=$PWD/nori_b200/libnori_b200_sweep.so
NORI_GPU_FETCH_SHADOW=20 python scripts/sweep.py NORI_GPU_FETCH=12,16,20,24 --spp 1024
python scripts/sweep.py NORI_GPU_NODE_EXIT_CLOSEST=2,3,4,5 --spp 1024
python scripts/sweep.py NORI_GPU_NODE_EXIT_SHADOW=3,4,5 --spp 1024
python scripts/sweep.py NORI_GPU_MAX_LEAF=2,3,4,6,8 --spp 512
python scripts/sweep.py NORI_GPU_SAH_CT=0.6,1.0,1.5 --spp 512
