"""B200-native LD-score engine: host-side mirror of the reference `ld` driver over hand-written sm_100a kernels."""
