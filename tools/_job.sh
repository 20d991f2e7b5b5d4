export CTN_ALLOW_EXPERIMENT_BUILD=1
for l in libctn_b200.so libctn_exp.so libctn_b200.so libctn_exp.so; do echo $l; CTN_LIB_FILE=$l python tools/profile_step_phases.py 3 | egrep "total|bwd"; done
CTN_LIB_FILE=libctn_exp.so python -m pytest tests/test_gpu_kernels.py tests/test_gpu_train.py -q -m gpu -x 2>&1 | tail -2
