for pr in 0 -1 -3 0 -1; do echo "prio $pr"; CTN_X_PRIO=$pr python tools/profile_step_phases.py 3 | egrep "total|fwd|bwd"; done
