CTN_DILS=1,8,32,128 timeout 200 python tools/bench_fused.py 64 2>&1 | grep dwpw
CTN_ALLOW_EXPERIMENT_BUILD=1 CTN_LIB_FILE=libctn_exp.so CTN_DILS=1,8,32,128 timeout 200 python tools/bench_fused.py 64 2>&1 | grep dwpw
