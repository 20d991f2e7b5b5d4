export CTN_ALLOW_EXPERIMENT_BUILD=1
for r in 1 2; do for l in libctn_b200.so libctn_expb.so libctn_expc.so; do echo $l; CTN_LIB_FILE=$l python tools/profile_step_phases.py 3 | egrep "bwd"; done; done
