#!/bin/bash
# Direct vector-store microbenchmark at the shapes of the fused kernels (build tools/build/direct_store_bench first).
B=tools/build/direct_store_bench
$B 200000 910 910 1
$B 200000 910 910 5
$B 200000 910 912 5
