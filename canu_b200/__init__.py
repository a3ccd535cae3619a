"""B200-native overlapInCore hot path.  `canu_b200.api` binds the C ABI of liboic_b200.so."""
