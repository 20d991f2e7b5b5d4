for f in 0 1 0 1; do echo "pdl force $f"; CTN_X_PDL_FORCE=$f python tools/profile_step_phases.py 3 | egrep "total|fwd|bwd"; done
