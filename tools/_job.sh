export CTN_ALLOW_EXPERIMENT_BUILD=1 CTN_LIB_FILE=libctn_exp.so CTN_TRACE=1
python tools/bench_updc1.py 3 2>&1 | grep -v "^[0-9]* $" | cut -c1-700
