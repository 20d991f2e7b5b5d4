"""Internals of the B200-native Conv-TasNet drop-in: ctypes binding of
libctn_b200.so (`_native`), state_dict layout (`_layout`) and the host-side
orchestration of forward / backward (`_engine`)."""
