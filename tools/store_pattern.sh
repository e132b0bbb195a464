#!/bin/bash
# Store-pattern microbenchmark at the shapes of the fused kernels (build tools/build/store_pattern_bench first).
B=tools/build/store_pattern_bench
$B 200000 910 910 1
$B 200000 256 910 1
$B 200000 256 910 8
$B 200000 256 912 8
